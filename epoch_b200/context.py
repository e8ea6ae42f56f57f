"""Host-side mirror of the reference's interface for the PIC hot path.

`PicContext` owns one C-ABI context (one rank == one GPU) and exposes the four
hot entries under the reference's own subroutine names
(`update_eb_fields_half`, `push_particles`, `current_finish`,
`update_eb_fields_final`; epoch3d/src/epoch3d.F90:208-246) plus the setters that
replace reading `MODULE shared_data` globals.  Python is harness glue only: all
arithmetic happens behind the ABI.
"""
import ctypes as C

import numpy as np

from . import _abi

# boundary codes (constants.F90:56-71)
BC = dict(periodic=1, other=2, simple_laser=3, simple_outflow=4, open=5, dump=6,
          zero_gradient=7, clamp=8, reflect=9, conduct=10, thermal=11,
          cpml_laser=12, cpml_outflow=13, mixed=14, heat_bath=15)
SHAPES = dict(tophat=0, triangle=1, bspline3=2)
PNG = {0: 2, 1: 3, 2: 4}  # constants.F90:472-485
FIELD_NAMES = ("ex", "ey", "ez", "bx", "by", "bz", "jx", "jy", "jz")


class EpochError(RuntimeError):
    pass


def _i32(a):
    return (C.c_int32 * len(a))(*[int(x) for x in a])


def _f64(a):
    return (C.c_double * len(a))(*[float(x) for x in a])


def _pad3(a, fill):
    a = list(a)
    return a + [fill] * (3 - len(a))


class PicContext:
    def __init__(self, ndims, shape="triangle", abi=None, device=0, rank=0, nproc=1,
                 nprocdir=(1, 1, 1), coords=(0, 0, 0), nccl_id=None):
        self.abi = abi if abi is not None else _abi.product()
        self.ndims = int(ndims)
        self.shape = SHAPES[shape] if isinstance(shape, str) else int(shape)
        self.ng = PNG[self.shape] + 2
        self.nvar = self.ndims + 4
        self._ctx = C.c_void_p()
        idbuf = None
        if nccl_id is not None:
            idbuf = C.create_string_buffer(bytes(nccl_id), 128)
        rc = self.abi.create(C.byref(self._ctx), self.ndims, self.shape, device, rank, nproc,
                             _i32(_pad3(nprocdir, 1)), _i32(_pad3(coords, 0)), idbuf)
        if rc != 0:
            raise EpochError(f"create failed rc={rc}: {self.abi.last_error(None)}")
        self.n = None
        self.nspecies = 0

    # -- plumbing ---------------------------------------------------------
    def _ck(self, rc):
        if rc != 0:
            msg = self.abi.last_error(self._ctx)
            raise EpochError(f"rc={rc}: {msg.decode() if msg else ''}")

    def close(self):
        if self._ctx:
            self.abi.destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- setters ----------------------------------------------------------
    def set_grid(self, n_local, d, grid_min_local, min_local, max_local, min_global, max_global,
                 n_global=None, n_global_min=None, cpml_thickness=0):
        nd = self.ndims
        self.n = tuple(int(x) for x in n_local[:nd])
        self.d = tuple(float(x) for x in d[:nd])
        n_global = n_global if n_global is not None else n_local
        n_global_min = n_global_min if n_global_min is not None else [1] * nd
        self._ck(self.abi.set_grid(self._ctx, _i32(_pad3(n_local, 1)), _i32(_pad3(n_global, 1)),
                                   _i32(_pad3(n_global_min, 1)), _f64(_pad3(d, 1.0)),
                                   _f64(_pad3(grid_min_local, 0.0)), _f64(_pad3(min_local, 0.0)),
                                   _f64(_pad3(max_local, 0.0)), _f64(_pad3(min_global, 0.0)),
                                   _f64(_pad3(max_global, 0.0)), cpml_thickness))
        self.field_shape = tuple(x + 2 * self.ng for x in self.n)  # Fortran order (x fastest)

    def set_time(self, dt, time, step=0):
        self._ck(self.abi.set_time(self._ctx, dt, time, step))

    def set_field_bcs(self, bcs):
        bcs = [BC[b] if isinstance(b, str) else int(b) for b in bcs]
        self._ck(self.abi.set_field_bcs(self._ctx, _i32((bcs + [1] * 6)[:6])))

    def set_solver(self, field_order=2, maxwell_solver=0, coeffs=None):
        self._ck(self.abi.set_solver(self._ctx, field_order, maxwell_solver,
                                     _f64(coeffs if coeffs is not None else [0.0] * 15)))

    def set_cpml(self, axis, kappa_e, kappa_b, sigma_e, sigma_b, a_e, a_b, min_start, min_end, max_start, max_end,
                 laser_idx_min=-1, laser_idx_max=-1):
        """what set_cpml_helpers leaves on the host for one axis (boundary.F90:1989-2479)"""
        arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (kappa_e, kappa_b, sigma_e, sigma_b, a_e, a_b)]
        assert all(len(a) == self.n[axis] + 2 * self.ng for a in arrs)
        self._ck(self.abi.set_cpml(self._ctx, axis, *[a.ctypes.data_as(C.POINTER(C.c_double)) for a in arrs],
                                   int(min_start), int(min_end), int(max_start), int(max_end),
                                   int(laser_idx_min), int(laser_idx_max)))

    def set_window_fields(self, planes=None):
        """ex_x_max .. bz_x_max of setup_field_boundaries (setup.F90:425-470); None = zeros."""
        planes = planes or {}
        keep = [None if planes.get(k) is None else np.ascontiguousarray(planes[k], dtype=np.float64)
                for k in ("ex", "ey", "ez", "bx", "by", "bz")]
        ptrs = [None if a is None else a.ctypes.data_as(C.POINTER(C.c_double)) for a in keep]
        self._ck(self.abi.set_window_fields(self._ctx, *ptrs))

    def setup_field_boundaries(self):
        """setup_field_boundaries (setup.F90:426-510) from the fields now on the device"""
        self._ck(self.abi.setup_field_boundaries(self._ctx))

    def set_boundary_fields(self, boundary, planes=None):
        """the six planes ex..bz of one boundary (1..6), None = zeros"""
        planes = planes or {}
        keep = [None if planes.get(k) is None else np.ascontiguousarray(planes[k], dtype=np.float64)
                for k in ("ex", "ey", "ez", "bx", "by", "bz")]
        ptrs = [None if a is None else a.ctypes.data_as(C.POINTER(C.c_double)) for a in keep]
        self._ck(self.abi.set_boundary_fields(self._ctx, int(boundary), *ptrs))

    def shift_window(self, x_grid_min_local, x_min_local, x_max_local, x_min, x_max):
        """one cell of shift_window (window.F90:62-94) minus insert_particles"""
        self._ck(self.abi.shift_window(self._ctx, x_grid_min_local, x_min_local, x_max_local, x_min, x_max))

    def set_current_options(self, smooth=False, smooth_its=1, smooth_comp_its=0, strides=None,
                            initial_j=None):
        """current_finish options (current_smooth.F90:29-151): binomial filter passes and
        the subtraction of the host's initial_jx,jy,jz (use_current_correction)."""
        st = np.ascontiguousarray(list(strides) if strides else [], dtype=np.int32)
        j0 = None if initial_j is None else _f64(initial_j)
        self._ck(self.abi.set_current_options(
            self._ctx, int(bool(smooth)), int(smooth_its), int(smooth_comp_its), len(st),
            st.ctypes.data_as(C.POINTER(C.c_int32)) if len(st) else None,
            int(initial_j is not None), j0))

    def set_species(self, isp, charge, mass, weight=1.0, immobile=False, zero_current=False,
                    species_type=0, bcs=None):
        nd = self.ndims
        bcs = bcs if bcs is not None else ["periodic"] * (2 * nd)
        bcs = [BC[b] if isinstance(b, str) else int(b) for b in bcs]
        self._ck(self.abi.set_species(self._ctx, isp, charge, mass, weight, int(immobile),
                                      int(zero_current), species_type, _i32((bcs + [1] * 6)[:6])))
        self.nspecies = max(self.nspecies, isp + 1)

    # -- particles (partlist wire format: pos(ndims), p(3), weight) ---------
    def host_empty(self, shape):
        """float64 array in page-locked host memory (eb200_host_alloc): uploads and downloads of such
        buffers skip the library's staging ring.  Falls back to ordinary memory when the allocation fails."""
        n = int(np.prod(shape))
        p = C.c_void_p()
        if n == 0 or self.abi.host_alloc(C.byref(p), n * 8) != 0 or not p.value:
            return np.empty(shape, dtype=np.float64)
        buf = (C.c_double * n).from_address(p.value)
        arr = np.frombuffer(buf, dtype=np.float64).reshape(shape)
        import weakref
        weakref.finalize(buf, self.abi.host_free, p.value)
        arr.flags.writeable = True
        self.pinned_host = True
        return arr

    def upload_particles(self, isp, packed, append=False):
        packed = np.ascontiguousarray(packed, dtype=np.float64).reshape(-1, self.nvar)
        fn = self.abi.append_particles if append else self.abi.upload_particles
        self._ck(fn(self._ctx, isp, packed.shape[0], packed.ctypes.data_as(C.c_void_p), self.nvar))

    def particle_count(self, isp):
        n = C.c_int64()
        self._ck(self.abi.particle_count(self._ctx, isp, C.byref(n)))
        return n.value

    def download_particles(self, isp, out=None):
        n = self.particle_count(isp)
        if out is None or out.shape[0] < n:
            out = np.empty((n, self.nvar), dtype=np.float64)
        got = C.c_int64()
        self._ck(self.abi.download_particles(self._ctx, isp, 0, n, out.ctypes.data_as(C.c_void_p),
                                             self.nvar, C.byref(got)))
        return out[:got.value]

    def set_thermal_boundary(self, isp, boundary, ext_temp, seed=0):
        """ext_temp_<face> of species isp (boundary 1..6 = x_min .. z_max; mpi_routines.F90:561-584): array shaped
        (3, n_t2 + 2 ng, n_t1 + 2 ng) -- the Fortran ext_temp(1-ng:n_t1+ng, 1-ng:n_t2+ng, 3) in C order; 1D: (3,)"""
        a = np.ascontiguousarray(ext_temp, dtype=np.float64)
        self._ck(self.abi.set_thermal_boundary(self._ctx, isp, int(boundary), a.ctypes.data_as(C.POINTER(C.c_double)), int(seed)))

    def set_probe(self, isp, iprobe, point, normal, ek_min=0.0, ek_max=1e300):
        """particle probe (shared_data.F90:392-404) iprobe = 0..7 of species isp"""
        self._ck(self.abi.set_probe(self._ctx, isp, iprobe, _f64(_pad3(point, 0.0)), _f64(_pad3(normal, 0.0)), ek_min, ek_max))

    def download_probe(self, isp, iprobe, clear=True):
        """the probe's sampled particles since the last clear (rows pos, p(3), weight) and the number sampled"""
        n = C.c_int64()
        self._ck(self.abi.download_probe(self._ctx, isp, iprobe, 0, None, self.nvar, C.byref(n), 0))
        out = np.empty((n.value, self.nvar))
        self._ck(self.abi.download_probe(self._ctx, isp, iprobe, n.value, out.ctypes.data_as(C.POINTER(C.c_double)),
                                         self.nvar, C.byref(n), int(bool(clear))))
        return out, n.value

    def dist_fn(self, isp, direction, ranges, resolution, dgrid, restrictions=None, use_restrictions=None):
        """particle loop of general_dist_fn (io/dist_fn.F90:532-628): this rank's weighted histogram.
        direction: 1-based c_dir_* codes; ranges [(lo, hi)] per dimension.  Returns an array shaped
        resolution[::-1] (C order of the Fortran array(res1, res2, res3))."""
        cur = len(direction)
        res3 = [int(r) for r in resolution] + [1] * (3 - cur)
        rng = np.zeros(6)
        for i, (lo, hi) in enumerate(ranges):
            rng[2 * i], rng[2 * i + 1] = lo, hi
        dg = np.ones(3)
        dg[:cur] = dgrid
        out = np.zeros(res3[::-1])
        rp = up = None
        if restrictions is not None:
            rarr = np.ascontiguousarray(restrictions, dtype=np.float64).reshape(-1)
            uarr = np.ascontiguousarray(use_restrictions, dtype=np.int32)
            assert len(uarr) == self.ndims + 9 and len(rarr) == 2 * len(uarr)
            rp, up = rarr.ctypes.data_as(C.POINTER(C.c_double)), uarr.ctypes.data_as(C.POINTER(C.c_int32))
        self._ck(self.abi.dist_fn(self._ctx, isp, cur, _i32(list(direction) + [1] * (3 - cur)),
                                  rng.ctypes.data_as(C.POINTER(C.c_double)), _i32(res3),
                                  dg.ctypes.data_as(C.POINTER(C.c_double)), rp, up,
                                  out.ctypes.data_as(C.POINTER(C.c_double))))
        return out[(0,) * (3 - cur)] if cur < 3 else out

    # -- fields (Fortran-shaped with ghosts; returned as numpy arrays indexed
    #    [k, j, i] i.e. C order of the column-major reference array) ---------
    def _np_shape(self):
        return tuple(reversed(self.field_shape))

    def upload_fields(self, **fields):
        ptrs = []
        keep = []
        for name in FIELD_NAMES:
            a = fields.get(name)
            if a is None:
                ptrs.append(None)
            else:
                a = np.ascontiguousarray(a, dtype=np.float64)
                assert a.size == int(np.prod(self.field_shape)), (name, a.shape, self.field_shape)
                keep.append(a)
                ptrs.append(a.ctypes.data_as(C.c_void_p))
        self._ck(self.abi.upload_fields(self._ctx, *ptrs))

    def download_fields(self, names=FIELD_NAMES, out=None):
        """`out`: dict of preallocated arrays of the field shape to fill (e.g. from host_empty)"""
        out = dict(out) if out else {}
        ptrs = []
        for name in FIELD_NAMES:
            if name in names:
                if name not in out:
                    out[name] = np.empty(self._np_shape(), dtype=np.float64)
                ptrs.append(out[name].ctypes.data_as(C.c_void_p))
            else:
                ptrs.append(None)
        self._ck(self.abi.download_fields(self._ctx, *ptrs))
        return out

    def interior(self, a):
        ng = self.ng
        return a[tuple(slice(ng, -ng) for _ in range(self.ndims))]

    # -- lasers -------------------------------------------------------------
    def set_laser(self, boundary, lid, amp, pol_angle=0.0, t_start=0.0, t_end=1e300,
                  profile=None, phase=None):
        keep = []

        def ptr(a):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=np.float64)
            keep.append(a)
            return a.ctypes.data_as(C.c_void_p)
        self._ck(self.abi.set_laser(self._ctx, boundary, lid, amp, pol_angle, t_start, t_end,
                                    ptr(profile), ptr(phase)))

    def set_laser_step(self, boundary, lid, t_env, integral_phase):
        self._ck(self.abi.set_laser_step(self._ctx, boundary, lid, t_env, integral_phase))

    # -- the reference's entry points -----------------------------------------
    def update_eb_fields_half(self):
        self._ck(self.abi.update_eb_fields_half(self._ctx))

    def push_particles(self):
        self._ck(self.abi.push_particles(self._ctx))

    def current_finish(self):
        self._ck(self.abi.current_finish(self._ctx))

    def update_eb_fields_final(self):
        self._ck(self.abi.update_eb_fields_final(self._ctx))

    def particle_bcs(self):
        self._ck(self.abi.particle_bcs(self._ctx))

    def efield_bcs(self):
        self._ck(self.abi.efield_bcs(self._ctx))

    def bfield_final_bcs(self):
        self._ck(self.abi.bfield_final_bcs(self._ctx))

    # -- diagnostics ----------------------------------------------------------
    def energy(self):
        fe = C.c_double()
        ps = (C.c_double * max(1, self.nspecies))()
        self._ck(self.abi.energy(self._ctx, C.byref(fe), ps))
        return fe.value, list(ps)[:self.nspecies]

    def charge_residual(self):
        a, b = C.c_double(), C.c_double()
        self._ck(self.abi.charge_residual(self._ctx, C.byref(a), C.byref(b)))
        return a.value, b.value

    def charge_residual_static(self):
        """charge_residual on the library's static device scratch: no allocation (operators' addition)"""
        a, b = C.c_double(), C.c_double()
        self._ck(self.abi.charge_residual_static(self._ctx, C.byref(a), C.byref(b)))
        return a.value, b.value

    # -- load balancing (housekeeping/balance.F90) ---------------------------------
    def get_load(self, grid_min_global, n_global):
        """this rank's load_x, load_y, load_z of part_load_func (before the ALLREDUCE)"""
        out = [np.zeros(int(n_global[a]) + 2 * self.ng, dtype=np.int64) for a in range(self.ndims)]
        ptr = [o.ctypes.data_as(C.POINTER(C.c_int64)) for o in out] + [None] * (3 - self.ndims)
        self._ck(self.abi.get_load(self._ctx, _f64(_pad3(grid_min_global, 0.0)), *ptr))
        return out

    def redistribute(self, old_min, old_max, new_min, new_max, grid_min_global, grid_min_local, min_local, max_local):
        """old_min .. new_max: per axis lists of the 1-based cell ranges of every processor coordinate"""
        flat = [_i32([int(v) for ax in t for v in ax]) for t in (old_min, old_max, new_min, new_max)]
        self._ck(self.abi.redistribute(self._ctx, *flat, _f64(_pad3(grid_min_global, 0.0)),
                                       _f64(_pad3(grid_min_local, 0.0)), _f64(_pad3(min_local, 0.0)),
                                       _f64(_pad3(max_local, 0.0))))

    DERIVED = {"mass_density": 0, "ekbar": 1, "ekflux": 2, "poynt_flux": 3, "charge_density": 4,
               "number_density": 5, "ppc": 6, "average_weight": 7, "temperature": 8,
               "species_current": 9, "average_momentum": 10}

    def calc_derived(self, kind, ispecies=-1, direction=0):
        """calc_<kind>(data_array, current_species, direction) of io/calc_df.F90; returns the
        array with ghosts (ispecies -1 = summed over species, direction 1..3 / -1..-3 / 0)"""
        out = np.empty(self._np_shape(), dtype=np.float64)
        k = self.DERIVED[kind] if isinstance(kind, str) else int(kind)
        self._ck(self.abi.calc_derived(self._ctx, k, int(ispecies), int(direction),
                                       out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def set_option(self, key, value):
        self._ck(self.abi.set_option(self._ctx, key.encode(), str(value).encode()))

    def sync(self):
        self._ck(self.abi.sync(self._ctx))

    def mark(self, phase, begin):
        self._ck(self.abi.mark(self._ctx, phase.encode(), int(bool(begin))))

    def timer(self, phase, reset=False):
        ms, n = C.c_double(), C.c_int64()
        self._ck(self.abi.timer(self._ctx, phase.encode(), int(reset), C.byref(ms), C.byref(n)))
        return ms.value, n.value
