"""Mirror of PROGRAM pic's control flow around the hot path.

Reproduces, for a single deck-like description, what the Fortran driver does
between `read_deck` and the end of the main loop as far as the hot path is
concerned: grid setup (housekeeping/setup.F90:174-231), time-step selection
(`set_dt`, setup.F90:700-767; 1D :566-620; 2D analogue), the initial boundary
calls (epoch3d.F90:141-158), and the per-step sequence with its time
bookkeeping (epoch3d.F90:184-249).  The arithmetic of the
hot path itself is behind `PicContext` (C ABI); this file is host logic only and
is shared by the tests (driving oracle and CUDA library alike) and by bench.py.
"""
import math

import numpy as np

from .context import BC, PicContext, SHAPES, PNG
from . import decomp

# constants.F90:144-151
PI = 3.141592653589793238462643383279503
Q0 = 1.602176565e-19
M0 = 9.10938291e-31
C = 2.99792458e8
KB = 1.3806488e-23
EPS0 = 8.854187817620389850536563031710750e-12
C_TINY = 1e-300


def gauss(x, x0, w):
    """evaluator_blocks.F90:1009-1012"""
    return np.exp(-((x - x0) / w) ** 2)


class LaserSpec:
    """deck `laser` block (deck_laser_block.f90): amplitude, omega, envelope."""

    def __init__(self, boundary, lid=1, intensity_w_cm2=None, amp=None, lambda_=None, omega=None,
                 pol_angle=0.0, t_start=0.0, t_end=None, t_profile=None, profile=None, phase=None):
        self.boundary = boundary  # 1..6
        self.id = lid
        if amp is None:
            # deck_laser_block.f90:135-139
            amp = math.sqrt(intensity_w_cm2 / (C * EPS0 / 2.0)) * 100.0
        self.amp = amp
        self.omega = omega if omega is not None else 2.0 * PI * C / lambda_  # laser.f90:325-328
        self.pol_angle = pol_angle
        self.t_start = t_start
        self.t_end = t_end
        self.t_profile = t_profile  # callable(time) or None (=1)
        self.profile = profile      # callable(*tangential coordinate arrays) or None
        self.phase = phase


class SpeciesSpec:
    def __init__(self, name, charge, mass, bcs=None, zero_current=False, immobile=False,
                 density=0.0, temperature=(0.0, 0.0, 0.0)):
        self.name, self.charge, self.mass = name, charge, mass
        self.bcs, self.zero_current, self.immobile = bcs, zero_current, immobile
        self.density = density            # for set_plasma_frequency_dt only
        self.temperature = temperature


# constants.F90:125-132
MAXWELL_SOLVERS = {"custom": -1, "yee": 0, "lehe": 1, "lehe_x": 2, "lehe_y": 3, "lehe_z": 4, "cowan": 5, "pukhov": 6}


class Simulation:
    def __init__(self, ndims, n_global, mins, maxs, shape="triangle", bc_field=None, species=(),
                 lasers=(), dt_multiplier=0.95, field_order=2, t_end=None, abi=None, device=0,
                 rank=0, nproc=1, nprocdir=(1, 1, 1), nccl_id=None, dt=None, maxwell_solver="yee",
                 stencil=None, cpml_thickness=6, cpml_kappa_max=20.0, cpml_a_max=0.15, cpml_sigma_max=0.7,
                 options=None):
        self.ndims = nd = ndims
        self.shape = SHAPES[shape] if isinstance(shape, str) else shape
        self.n_global = tuple(n_global[:nd])
        self.mins, self.maxs = tuple(mins[:nd]), tuple(maxs[:nd])
        self.bc_field = [BC[b] if isinstance(b, str) else b for b in
                         (bc_field or ["periodic"] * (2 * nd))]
        self.species, self.lasers = list(species), list(lasers)
        self.field_order = field_order
        self.t_end = t_end
        self.rank, self.nproc = rank, nproc
        self.nprocdir = tuple(nprocdir)
        self.coords = decomp.rank_to_coords(rank, self.nprocdir)
        # CPML: any cpml_* field boundary widens the grid by cpml_thickness cells on BOTH sides of
        # EVERY axis (setup.F90:400-414), else the thickness is zero (:254-258)
        bcn = [self._norm_field_bc(b) for b in self.bc_field]
        self.cpml_boundaries = any(b in (BC["cpml_laser"], BC["cpml_outflow"]) for b in bcn)
        self.cpml_thickness = t = int(cpml_thickness) if self.cpml_boundaries else 0
        self.cpml_kappa_max, self.cpml_a_max, self.cpml_sigma_max = cpml_kappa_max, cpml_a_max, cpml_sigma_max
        self.n_global = tuple(n + 2 * t for n in self.n_global)
        # setup_grid: setup.F90:174-231
        self.length = [self.maxs[d] - self.mins[d] for d in range(nd)]
        self.d = [self.length[d] / float(self.n_global[d] - 2 * t) for d in range(nd)]
        self.grid_min = [(self.mins[d] - self.d[d] * t) + self.d[d] / 2.0 for d in range(nd)]
        # split_domain: near-even split (mpi_routines.F90:44-270)
        self.cell_min, self.cell_max = [], []
        for d in range(nd):
            lo, hi = decomp.cell_range(self.n_global[d], self.nprocdir[d], self.coords[d])
            self.cell_min.append(lo)
            self.cell_max.append(hi)
        self.n_local = [self.cell_max[d] - self.cell_min[d] + 1 for d in range(nd)]
        # cell_x_min / cell_x_max ... of every processor coordinate (mpi_routines.F90:444-476)
        self.cell_mins = [[decomp.cell_range(self.n_global[d], self.nprocdir[d], p)[0] for p in range(self.nprocdir[d])]
                          for d in range(nd)]
        self.cell_maxs = [[decomp.cell_range(self.n_global[d], self.nprocdir[d], p)[1] for p in range(self.nprocdir[d])]
                          for d in range(nd)]
        # setup_grid_x: utilities.f90:343-373 (x_global(ix) = x_grid_min + (ix-1)*dx)
        self.grid_min_local = [self.grid_min[d] + (self.cell_min[d] - 1) * self.d[d] for d in range(nd)]
        grid_max_local = [self.grid_min[d] + (self.cell_max[d] - 1) * self.d[d] for d in range(nd)]
        self.fng = field_order // 2      # set_field_order, fields.f90:34-49
        cp = [self.set_cpml_helpers(d) for d in range(nd)] if self.cpml_boundaries else None
        off = [(c["min_offset"], c["max_offset"]) for c in cp] if cp else [(0, 0)] * nd
        self.min_local = [self.grid_min_local[d] + (off[d][0] - 0.5) * self.d[d] for d in range(nd)]
        self.max_local = [grid_max_local[d] - (off[d][1] - 0.5) * self.d[d] for d in range(nd)]
        self._cpml_off = off

        self.ctx = PicContext(nd, self.shape, abi=abi, device=device, rank=rank, nproc=nproc,
                              nprocdir=self.nprocdir, coords=self.coords, nccl_id=nccl_id)
        for k, v in (options or {}).items():      # library tuning options (eb200_set_option), before any allocation
            self.ctx.set_option(k, v)
        self.ctx.set_grid(self.n_local, self.d, self.grid_min_local, self.min_local, self.max_local,
                          self.mins, self.maxs, n_global=self.n_global, n_global_min=self.cell_min,
                          cpml_thickness=self.cpml_thickness)
        self.ctx.set_field_bcs(self.bc_field)
        if cp:
            for d in range(nd):
                self.ctx.set_cpml(d, **{k: v for k, v in cp[d].items() if k not in ("min_offset", "max_offset")})
        self.maxwell_solver = MAXWELL_SOLVERS[maxwell_solver] if isinstance(maxwell_solver, str) else maxwell_solver
        if self.maxwell_solver == 1:          # "lehe" is the deck's alias of lehe_x (deck_control_block.F90:121-122)
            self.maxwell_solver = 2
        self.stencil = dict(stencil or {})    # custom solver: betaxy ... deltaz (+ dt) from the deck's stencil block
        self.dt_multiplier = dt_multiplier
        self.dt = dt if dt is not None else self.set_dt()
        self.coeffs = self.set_maxwell_solver()
        self.ctx.set_solver(field_order, self.maxwell_solver, self.coeffs)
        for i, sp in enumerate(self.species):
            self.ctx.set_species(i, sp.charge, sp.mass, 1.0, sp.immobile, sp.zero_current, 0, sp.bcs)
        self.time = 0.0
        self.step = 0
        self._setup_lasers()
        self._started = False

    # -- set_dt -----------------------------------------------------------------
    def set_dt(self):
        nd, d = self.ndims, self.d
        cfl = {2: 1.0, 4: 6.0 / 7.0, 6: 120.0 / 149.0}[self.field_order]  # fields.f90:34-49
        if nd == 1:
            solver = d[0] / C
        elif nd == 2:
            solver = d[0] * d[1] / math.sqrt(d[0] ** 2 + d[1] ** 2) / C
        else:
            solver = d[0] * d[1] * d[2] / math.sqrt((d[0] * d[1]) ** 2 + (d[1] * d[2]) ** 2
                                                   + (d[2] * d[0]) ** 2) / C
        dt = cfl * solver
        ms = self.maxwell_solver
        if ms in (2, 3, 4):
            # R. Lehe: dt = MIN(d_a, transverse CFL) / c  (setup.F90:713-721; 2D/1D analogues)
            a = ms - 2
            if a >= nd:
                raise ValueError("lehe solver axis outside the simulation's dimensions")
            others = [d[k] for k in range(nd) if k != a]
            if len(others) == 2:
                trans = others[0] * others[1] / math.sqrt(others[0] ** 2 + others[1] ** 2)
            elif len(others) == 1:
                trans = others[0]
            else:
                trans = d[a]
            dt = min(d[a], trans) / C
        elif ms in (5, 6):
            dt = min(d) / C                    # Cowan, Pukhov (setup.F90:723-727)
        bcs = [self._norm_field_bc(b) for b in self.bc_field]
        any_open = any(b in (BC["simple_laser"], BC["simple_outflow"]) for b in bcs)
        if any_open:
            dt = min(dt, solver)
        if ms == -1:                           # custom stencil: the deck gives dt (setup.F90:734-746) ...
            dt = self.stencil["dt"]
            if self.dt_multiplier < 1.0:       # ... and a multiplier below one would detune it (:735-746: warns, resets)
                self.dt_multiplier = 1.0
        # set_plasma_frequency_dt (setup.F90:640-696; 1D :514-563)
        min_dt = 1000000.0
        k_max = 2.0 * PI / d[0]
        for sp in self.species:
            omega2 = Q0 ** 2 / sp.mass / EPS0 * sp.density + 3.0 * k_max ** 2 * KB / sp.mass * max(sp.temperature)
            if omega2 > C_TINY:
                min_dt = min(min_dt, 2.0 * PI / math.sqrt(omega2))
        dt_plasma = min_dt / 2.0
        if dt_plasma > C_TINY:
            dt = min(dt, dt_plasma)
        # set_laser_dt (laser.f90:452-505)
        if self.lasers:
            dt_laser = min(2.0 * PI / L.omega for L in self.lasers) / 2.0
            if dt_laser > C_TINY:
                dt = min(dt, dt_laser)
        return self.dt_multiplier * dt

    # -- set_cpml_helpers ----------------------------------------------------------
    def set_cpml_helpers(self, axis):
        """One axis of set_cpml_helpers (boundary.F90:1989-2479; 1D :815-966): profiles over
        (1-ng : n+ng), local slab ranges, offsets of x_min_local / x_max_local, laser planes."""
        from .context import PNG
        ng = PNG[self.shape] + 2
        n, d = self.n_local[axis], self.d[axis]
        gmin, gmax, nglob = self.cell_min[axis], self.cell_max[axis], self.n_global[axis]
        t, fng = self.cpml_thickness, self.fng
        cpml_m, cpml_ma = 3, 1
        kap_e, kap_b = np.ones(n + 2 * ng), np.ones(n + 2 * ng)
        sig_e, sig_b, a_e, a_b = (np.zeros(n + 2 * ng) for _ in range(4))
        at = lambda i: i + ng - 1                     # array position of local index i
        sigma_maxval = self.cpml_sigma_max * C * 0.8 * (cpml_m + 1.0) / d
        out = dict(min_start=n + 1, min_end=0, max_start=n + 1, max_end=0, min_offset=0, max_offset=0,
                   laser_idx_min=-1, laser_idx_max=-1)
        bc_min = self._norm_field_bc(self.bc_field[2 * axis])
        bc_max = self._norm_field_bc(self.bc_field[2 * axis + 1])
        cp = (BC["cpml_laser"], BC["cpml_outflow"])

        def profile(x_pos):
            return (1.0 + (self.cpml_kappa_max - 1.0) * x_pos ** cpml_m, sigma_maxval * x_pos ** cpml_m,
                    self.cpml_a_max * (1.0 - x_pos) ** cpml_ma)
        if bc_min in cp:
            if gmin <= t:
                out["min_start"] = 1
                if gmax >= t:
                    out["min_end"] = t - gmin + 1
                    out["min_offset"] = t - gmin + 1
                else:
                    out["min_end"] = n
                    out["min_offset"] = t
                for i in range(out["min_start"], out["min_end"] + 1):
                    ig = i + gmin - 1
                    kap_e[at(i)], sig_e[at(i)], a_e[at(i)] = profile(1.0 - float(ig - 1) / float(t))
                    kap_b[at(i)], sig_b[at(i)], a_b[at(i)] = profile(1.0 - (float(ig) - 0.5) / float(t))
            if gmin <= t + fng + 1 <= gmax:
                out["laser_idx_min"] = t + fng + 1 - gmin
        if bc_max in cp:
            if gmax >= nglob - t + 1:
                out["max_end"] = n
                if gmin <= nglob - t + 1:
                    out["max_start"] = nglob - t + 1 - gmin + 1
                    out["max_offset"] = t - nglob + gmax
                else:
                    out["max_start"] = 1
                    out["max_offset"] = t
                for i in range(out["max_start"], out["max_end"] + 1):
                    ig = nglob - (i + gmin - 1) + 1
                    kap_e[at(i)], sig_e[at(i)], a_e[at(i)] = profile(1.0 - float(ig - 1) / float(t))
                    kap_b[at(i - 1)], sig_b[at(i - 1)], a_b[at(i - 1)] = profile(1.0 - (float(ig) - 0.5) / float(t))
            if gmin <= nglob - t - fng + 2 <= gmax:
                out["laser_idx_max"] = nglob - t - fng + 2 - gmin
        out.update(kappa_e=kap_e, kappa_b=kap_b, sigma_e=sig_e, sigma_b=sig_b, a_e=a_e, a_b=a_b)
        return out

    # -- set_maxwell_solver ------------------------------------------------------
    def set_maxwell_solver(self):
        """Coefficients of the extended B-update stencils (fields.f90:53-179; 2D :51-100; 1D
        :50-72) in the ABI's order: alpha x,y,z | beta xy,xz,yx,yz,zx,zy | gamma x,y,z |
        delta x,y,z.  Terms of directions the simulation does not have stay zero."""
        nd, ms = self.ndims, self.maxwell_solver
        d = list(self.d) + [1.0] * (3 - nd)
        al = [1.0, 1.0, 1.0]
        be = {k: 0.0 for k in ("xy", "xz", "yx", "yz", "zx", "zy")}
        ga = [0.0, 0.0, 0.0]
        de = [0.0, 0.0, 0.0]
        ax = "xyz"
        has = lambda k: ax.index(k[0]) < nd and ax.index(k[1]) < nd
        if ms == 0:
            return [0.0] * 15
        if ms == -1:
            for k in be:
                be[k] = float(self.stencil.get("beta" + k, 0.0)) if has(k) else 0.0
            for i in range(nd):
                de[i] = float(self.stencil.get("delta" + ax[i], 0.0))
                if nd == 3:
                    ga[i] = float(self.stencil.get("gamma" + ax[i], 0.0))
        elif ms in (2, 3, 4):
            a = ms - 2
            dx_cdt = d[a] / (C * self.dt)
            for o in range(nd):
                if o == a:
                    continue
                be[ax[a] + ax[o]] = 0.125 * (d[a] / d[o]) ** 2
                be[ax[o] + ax[a]] = 0.125
            de[a] = 0.25 * (1.0 - dx_cdt ** 2 * math.sin(0.5 * PI / dx_cdt) ** 2)
        elif ms in (5, 6):
            if nd == 1 or (ms == 5 and nd != 3):
                raise ValueError("this maxwell_solver does not exist in this dimensionality")
            delta = min(d[:nd])
            cc = [(delta / d[i]) ** 2 for i in range(3)]
            if ms == 5:      # Cowan
                cx1 = 1.0 / (cc[0] * cc[1] + cc[1] * cc[2] + cc[0] * cc[2])
                cx2 = 1.0 - cc[0] * cc[1] * cc[2] * cx1
                be["yx"] = 0.125 * cc[0] * cx2
                be["xy"] = 0.125 * cc[1] * cx2
                be["xz"] = 0.125 * cc[2] * cx2
                ga[0] = cc[1] * cc[2] * (0.0625 - 0.125 * cc[1] * cc[2] * cx1)
                ga[1] = cc[0] * cc[2] * (0.0625 - 0.125 * cc[0] * cc[2] * cx1)
                ga[2] = cc[0] * cc[1] * (0.0625 - 0.125 * cc[0] * cc[1] * cx1)
            else:            # Pukhov
                be["yx"] = 0.125 * cc[0]
                be["xy"] = 0.125 * cc[1]
                be["xz"] = 0.125 * cc[2] if nd == 3 else 0.0
            if nd == 3:
                be["zx"], be["zy"], be["yz"] = be["yx"], be["xy"], be["xz"]
        al[0] = 1.0 - 2.0 * be["xy"] - 2.0 * be["xz"] - 4.0 * ga[0] - 3.0 * de[0]
        al[1] = 1.0 - 2.0 * be["yx"] - 2.0 * be["yz"] - 4.0 * ga[1] - 3.0 * de[1]
        al[2] = 1.0 - 2.0 * be["zx"] - 2.0 * be["zy"] - 4.0 * ga[2] - 3.0 * de[2]
        return al + [be["xy"], be["xz"], be["yx"], be["yz"], be["zx"], be["zy"]] + ga + de

    @staticmethod
    def _norm_field_bc(b):
        # setup_boundaries, boundary.F90:45-58
        if b == BC["other"] or b == BC["reflect"]:
            return BC["clamp"]
        if b == BC["open"]:
            return BC["simple_outflow"]
        return b

    # -- lasers -----------------------------------------------------------------
    def _tangential_coords(self, boundary):
        """cell-centre coordinates over indices 0..n of the tangential dims, as
        evaluated by laser_update_profile (laser.f90:232-268) on this rank."""
        a = (boundary - 1) // 2
        axes = [d for d in range(self.ndims) if d != a]
        lines = [self.grid_min_local[d] + (np.arange(0, self.n_local[d] + 1) - 1) * self.d[d] for d in axes]
        if not lines:
            return []
        if len(lines) == 1:
            return [lines[0]]
        g0, g1 = np.meshgrid(lines[0], lines[1], indexing="ij")  # [i, j]
        return [g0, g1]

    def _setup_lasers(self):
        for L in self.lasers:
            a = (L.boundary - 1) // 2
            on_bnd = self.coords[a] == (0 if (L.boundary - 1) % 2 == 0 else self.nprocdir[a] - 1)
            L.active = on_bnd
            if not on_bnd:
                continue
            coords = self._tangential_coords(L.boundary)

            def plane(fn):
                if fn is None:
                    return None
                v = np.asarray(fn(*coords), dtype=np.float64) if coords else np.asarray([fn()], dtype=np.float64)
                if coords:
                    v = np.broadcast_to(v, coords[0].shape)
                # storage: lower axis fastest
                return np.ascontiguousarray(v.T if v.ndim == 2 else v)
            t_end = L.t_end if L.t_end is not None else (self.t_end if self.t_end is not None else 1e300)
            self.ctx.set_laser(L.boundary, L.id, L.amp, L.pol_angle, L.t_start, t_end,
                               plane(L.profile), plane(L.phase))

    def _laser_step(self):
        # update_laser_omegas + laser_time_profile (laser.f90:207-226,334-346)
        for L in self.lasers:
            if not L.active:
                continue
            t_env = 1.0 if L.t_profile is None else float(L.t_profile(self.time))
            self.ctx.set_laser_step(L.boundary, L.id, t_env, L.omega * self.time)

    # -- PROGRAM pic ------------------------------------------------------------
    def start(self):
        """epoch3d.F90:141-158: particle_bcs; efield_bcs; half-dt bfield_final_bcs."""
        ctx = self.ctx
        ctx.set_time(self.dt, self.time, self.step)
        if not getattr(self, "_bnd_fields_done", False):
            ctx.setup_field_boundaries()      # after_deck_last (setup.F90:241): planes of the initial fields
            self._bnd_fields_done = True
        ctx.particle_bcs()
        ctx.efield_bcs()
        dt_store = self.dt
        self.time = self.time + dt_store / 2.0
        ctx.set_time(dt_store / 2.0, self.time, self.step)
        self._laser_step()
        ctx.bfield_final_bcs()
        ctx.set_time(dt_store, self.time, self.step)
        self._started = True

    # -- distribution functions (io/dist_fn.F90) -----------------------------------
    DIRS = ("px", "py", "pz", "en", "gamma_m1", "xy_angle", "yz_angle", "zx_angle", "mod_p")

    def dir_code(self, name):
        """c_dir_* of constants.F90:202-213 for this dimensionality"""
        if name in ("x", "y", "z"):
            return "xyz".index(name) + 1
        return self.ndims + 1 + self.DIRS.index(name)

    def general_dist_fn(self, isp, direction, ranges_in, resolution_in, restrictions=None):
        """general_dist_fn (io/dist_fn.F90:118-762) up to the local array: the host part before the particle
        loop -- spatial ranges snapped to cells and clipped to this rank (:204-335), the -1..1 fall-back of an
        empty range and dgrid (:490-518) -- then the loop on the device.  direction: names ("x", "px", "en", ...);
        restrictions: {name: (lo, hi)}.  Returns (array, [grid per dimension]); the MPI_REDUCE over ranks that share
        the spatial range stays with the caller."""
        nd = self.ndims
        c_tiny = 2.2250738585072014e-308
        ranges, resolution, dgrid, codes = [], [], [], []
        outside = False
        for name, (r1, r2), res in zip(direction, ranges_in, resolution_in):
            code = self.dir_code(name)
            codes.append(code)
            if code <= nd:
                d = code - 1
                dx = self.d[d]
                gmin_l = self.grid_min_local[d]
                gmax_l = self.grid_min_local[d] + (self.n_local[d] - 1) * dx
                if abs(r1 - r2) <= c_tiny:
                    r1, r2 = gmin_l - 0.5 * dx, gmax_l + 0.5 * dx
                else:
                    r1, r2 = max(r1, self.mins[d]), min(r2, self.maxs[d])
                    r1 = self.min_local[d] + math.floor((r1 - self.min_local[d]) / dx) * dx
                    r2 = self.min_local[d] + math.ceil((r2 - self.min_local[d]) / dx) * dx
                    r1, r2 = max(r1, gmin_l - 0.5 * dx), min(r2, gmax_l + 0.5 * dx)
                res = int(round((r2 - r1) / dx))        # NINT
                if res <= 0:
                    outside, res = True, 0
                dg = dx
            else:
                if abs(r1 - r2) <= c_tiny:
                    r1, r2 = -1.0, 1.0
                dg = (r2 - r1) / float(res)
            ranges.append((r1, r2)); resolution.append(res); dgrid.append(dg)
        grids = [np.asarray([r[0] + 0.5 * g + i * g for i in range(n)]) for r, g, n in zip(ranges, dgrid, resolution)]
        if outside:
            return np.zeros((1,) * len(direction)), grids
        rs = us = None
        if restrictions:
            rs, us = np.zeros((nd + 9, 2)), np.zeros(nd + 9, dtype=np.int32)
            for name, (lo, hi) in restrictions.items():
                k = self.dir_code(name) - 1
                rs[k], us[k] = (lo, hi), 1
        return self.ctx.dist_fn(isp, codes, ranges, resolution, dgrid, rs, us), grids

    # -- load balancing (housekeeping/balance.F90) --------------------------------
    def balance_workload(self, allreduce=None, new_tables=None):
        """The re-slicing part of balance_workload (balance.F90:175-319) with over_ride: global
        load histograms from the device (get_load), calculate_breaks per decomposed axis, then
        redistribute_domain on the device.  allreduce(np.int64 array) -> sum over ranks.
        new_tables = (mins, maxs) overrides the computed breaks (tests)."""
        from . import balance
        nd = self.ndims
        if new_tables is None:
            local = self.ctx.get_load(self.grid_min, self.n_global)
            loads = balance.global_load(local, self.n_global, self.ctx.ng, allreduce)
            new_mins, new_maxs = [], []
            for a in range(nd):
                mn, mx = balance.calculate_breaks(loads[a], self.nprocdir[a], self.ctx.ng, balance.ncell_min(PNG[self.shape]))
                new_mins.append(mn)
                new_maxs.append(mx)
        else:
            new_mins, new_maxs = [list(map(int, t)) for t in new_tables[0]], [list(map(int, t)) for t in new_tables[1]]
        if new_mins == self.cell_mins and new_maxs == self.cell_maxs:
            return False
        old_mins, old_maxs = self.cell_mins, self.cell_maxs
        old_n_local = list(self.n_local)
        self.cell_mins, self.cell_maxs = new_mins, new_maxs
        self.cell_min = [new_mins[d][self.coords[d]] for d in range(nd)]
        self.cell_max = [new_maxs[d][self.coords[d]] for d in range(nd)]
        self.n_local = [self.cell_max[d] - self.cell_min[d] + 1 for d in range(nd)]
        self.grid_min_local = [self.grid_min[d] + (self.cell_min[d] - 1) * self.d[d] for d in range(nd)]
        grid_max_local = [self.grid_min[d] + (self.cell_max[d] - 1) * self.d[d] for d in range(nd)]
        cp = [self.set_cpml_helpers(d) for d in range(nd)] if self.cpml_boundaries else None
        off = [(c["min_offset"], c["max_offset"]) for c in cp] if cp else [(0, 0)] * nd
        self._cpml_off = off
        self.min_local = [self.grid_min_local[d] + (off[d][0] - 0.5) * self.d[d] for d in range(nd)]
        self.max_local = [grid_max_local[d] - (off[d][1] - 0.5) * self.d[d] for d in range(nd)]
        self.ctx.redistribute(old_mins, old_maxs, new_mins, new_maxs, self.grid_min, self.grid_min_local,
                              self.min_local, self.max_local)
        self.ctx.n = tuple(self.n_local)
        self.ctx.field_shape = tuple(x + 2 * self.ctx.ng for x in self.ctx.n)
        # what the host redistributes itself and hands over again
        if cp:
            for d in range(nd):
                self.ctx.set_cpml(d, **{k: v for k, v in cp[d].items() if k not in ("min_offset", "max_offset")})
        self._setup_lasers()
        if getattr(self, "window", None) is not None:
            self.ctx.set_window_fields(None)
        # the planes of setup_field_boundaries: EPOCH's host remaps its own copies (balance.F90:847-1010); this mirror
        # keeps none, so re-sliced planes are handed over as zeros (decks with initial fields at an open boundary
        # AND load balancing need the Fortran host)
        for b in range(2 * nd):
            others = [a for a in range(nd) if a != b // 2]
            if any(old_n_local[a] != self.n_local[a] for a in others):     # (planes of unchanged extent stay on the device)
                self.ctx.set_boundary_fields(b + 1, None)
        return True

    # -- moving window (housekeeping/window.F90) ---------------------------------
    def enable_window(self, v_x, start_time=0.0, stop_time=1e300, insert=None, planes=None):
        """move_window = T.  `insert(sim, x0)` stands for insert_particles (:198-358): it returns,
        per species, the particles of the incoming cell column [x0, x0 + dx) of this rank (or None);
        in EPOCH this is the deck's density / temperature expressions and the host RNG."""
        self.window = dict(v_x=v_x, start=start_time, stop=stop_time, insert=insert, started=False, frac=0.0)
        self.window_offset = 0.0
        self.ctx.set_window_fields(planes)
        # global grid bookkeeping of shift_window: x_grid_min (centre of global cell 1), xb_min, x_min, x_max
        self._x_grid_min = self.grid_min[0]
        self._xb_min = self.mins[0] - self.d[0] * self.cpml_thickness      # xb_global(1): outer edge of the layer
        self._x_min, self._x_max = self.mins[0], self.maxs[0]

    def _shift_window(self, cells):
        dx = self.d[0]
        for _ in range(cells):
            if self.window["insert"] is not None and self.coords[0] == self.nprocdir[0] - 1:
                x_grid_max = self._x_grid_min + (self.n_global[0] - 1) * dx
                x0 = x_grid_max + 0.5 * dx
                new = self.window["insert"](self, x0)
                for isp, P in enumerate(new or []):
                    if P is not None and len(P):
                        self.ctx.upload_particles(isp, P, append=True)
            # shift the box (window.F90:74-86) and setup_grid_x (utilities.f90:343-373)
            self._x_grid_min = self._x_grid_min + dx
            self._xb_min = self._xb_min + dx
            self._x_min = self._xb_min + dx * self.cpml_thickness
            self._x_max = (self._xb_min + self.n_global[0] * dx) - dx * self.cpml_thickness
            self.grid_min[0] = self._x_grid_min
            self.grid_min_local[0] = self._x_grid_min + (self.cell_min[0] - 1) * dx
            grid_max_local = self._x_grid_min + (self.cell_max[0] - 1) * dx
            self.min_local[0] = self.grid_min_local[0] + (self._cpml_off[0][0] - 0.5) * dx
            self.max_local[0] = grid_max_local - (self._cpml_off[0][1] - 0.5) * dx
            self.mins = (self._x_min,) + tuple(self.mins[1:])
            self.maxs = (self._x_max,) + tuple(self.maxs[1:])
            self.ctx.shift_window(self.grid_min_local[0], self.min_local[0], self.max_local[0], self._x_min, self._x_max)

    def moving_window(self):
        """moving_window (window.F90:388-436), called where epoch3d.F90:248 does"""
        w = getattr(self, "window", None)
        if w is None:
            return
        if not w["started"]:
            if w["start"] <= self.time < w["stop"]:
                w["frac"] = 0.0
                w["started"] = True
        if w["started"]:
            if self.time >= w["stop"] or w["v_x"] <= 0.0:
                return
            w["frac"] = w["frac"] + self.dt * w["v_x"] / self.d[0]
            cells = int(math.floor(w["frac"]))
            if cells > 0:
                self.window_offset = self.window_offset + float(cells) * self.d[0]
                self._shift_window(cells)
                self.ctx.particle_bcs()
                w["frac"] = w["frac"] - float(cells)

    def finished(self):
        return self.t_end is not None and self.time >= self.t_end

    def step_once(self, output=None, push=True, after_half=None):
        """One iteration of the main DO loop (epoch3d.F90:184-249).  `output` is
        called where output_routines is (time = (n+1)*dt, fields at n+1/2)."""
        if not self._started:
            self.start()
        ctx = self.ctx
        ctx.set_time(self.dt, self.time, self.step)
        ctx.update_eb_fields_half()
        if after_half is not None:
            # E, B at t+dt/2 and particles at x(t): div E = rho(x + half drift)/eps0 here
            after_half(self)
        if push and self.species:
            ctx.push_particles()
            ctx.current_finish()
        self.step += 1
        self.time = self.time + self.dt / 2.0
        if output is not None:
            output(self)
        self.time = self.time + self.dt / 2.0
        ctx.set_time(self.dt, self.time, self.step)
        self._laser_step()
        ctx.update_eb_fields_final()
        self.moving_window()             # epoch3d.F90:248
