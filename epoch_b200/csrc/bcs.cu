// bcs.cu -- field halo exchange, physical field boundaries, laser/outflow planes
// and the current ghost-sum.  Replaces, from epochNd/src/boundary.F90: field_bc /
// do_field_mpi_with_lengths (:145-152,:321-468), field_zero_gradient (:623-698),
// field_clamp_zero (:702-783), particle_reflection_bcs (:787-919),
// particle_periodic_bcs (:923-1084), efield_bcs (:1148), bfield_bcs (:1206),
// bfield_final_bcs (:1267); outflow_bcs_* of laser.f90:509-973; and current_finish
// of housekeeping/current_smooth.F90:29-45.
//
// All kernels are "one thread per grid line along the boundary-normal axis": a
// line is a fixed position in the two other axes (ghosts included, so that the
// reference's x-then-y-then-z ordering fills edges and corners).  For y/z normals
// consecutive threads walk consecutive x => coalesced; x-normal slabs are ng-wide
// strided pieces and small.
#include "ctx.cuh"
#include <algorithm>

namespace eb {

struct LineP {
  int axis, ng, nn;           // normal axis, ghost width, interior cells along it
  long long sa;               // stride along the axis
  int e1, e2;                 // extents of the two other axes (ghosts included)
  long long s1, s2;           // their strides
  long long nlines;
};

static LineP line_params(const eb200_ctx* c, int axis) {
  const GridP& g = c->g;
  LineP p;
  const int a1 = (axis + 1) % 3, a2 = (axis + 2) % 3;
  p.axis = axis; p.ng = g.ng; p.nn = g.n[axis]; p.sa = g.st[axis];
  p.e1 = g.n[a1] + 2 * g.ngd[a1]; p.e2 = g.n[a2] + 2 * g.ngd[a2];
  p.s1 = g.st[a1]; p.s2 = g.st[a2];
  p.nlines = (long long)p.e1 * p.e2;
  // make the x axis (stride 1) the fastest varying line coordinate when it is a1 or a2
  if (p.s2 < p.s1) { int t = p.e1; p.e1 = p.e2; p.e2 = t; long long s = p.s1; p.s1 = p.s2; p.s2 = s; }
  return p;
}
__device__ __forceinline__ long long line_base(const LineP& p, long long t) {
  return (t % p.e1) * p.s1 + (t / p.e1) * p.s2;
}
// element with Fortran coordinate c along the axis
#define AT(base, cc) ((base) + (long long)((cc) + p.ng - 1) * p.sa)

struct Arr3 { double* a[3]; };

static inline dim3 line_grid(const LineP& p, int ncomp) {
  return dim3((unsigned)((p.nlines + 127) / 128), (unsigned)ncomp, 1);
}

// ---- periodic halo on a single rank --------------------------------------------
__global__ void k_halo_periodic(LineP p, Arr3 A, int do_max, int do_min) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.nlines) return;
  double* a = A.a[blockIdx.y];
  const long long b = line_base(p, t);
  if (do_max) for (int i = 1; i <= p.ng; ++i) a[AT(b, p.nn + i)] = a[AT(b, i)];
  if (do_min) for (int i = 1; i <= p.ng; ++i) a[AT(b, i - p.ng)] = a[AT(b, p.nn + i - p.ng)];
}
// ---- pack / unpack / add of ng-thick slabs (multi-rank) -------------------------
// slab = coordinates c0+1 .. c0+ng along the axis; buffer layout [comp][i][line]
__global__ void k_pack(LineP p, Arr3 A, int c0, double* buf) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.nlines) return;
  const double* a = A.a[blockIdx.y];
  const long long b = line_base(p, t);
  double* o = buf + (long long)blockIdx.y * p.ng * p.nlines;
  for (int i = 1; i <= p.ng; ++i) o[(long long)(i - 1) * p.nlines + t] = a[AT(b, c0 + i)];
}
__global__ void k_unpack(LineP p, Arr3 A, int c0, const double* buf, int add) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.nlines) return;
  double* a = A.a[blockIdx.y];
  const long long b = line_base(p, t);
  const double* o = buf + (long long)blockIdx.y * p.ng * p.nlines;
  for (int i = 1; i <= p.ng; ++i) {
    const double v = o[(long long)(i - 1) * p.nlines + t];
    if (add) a[AT(b, c0 + i)] += v; else a[AT(b, c0 + i)] = v;
  }
}

// both sides in one launch (blockIdx.z = side): side 0 uses (c0a, bufa), side 1 (c0b, bufb); on = which sides exist
__global__ void k_pack2(LineP p, Arr3 A, int c0a, int c0b, double* bufa, double* bufb) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.nlines) return;
  const double* a = A.a[blockIdx.y];
  const long long b = line_base(p, t);
  const int c0 = blockIdx.z ? c0b : c0a;
  double* o = (blockIdx.z ? bufb : bufa) + (long long)blockIdx.y * p.ng * p.nlines;
  for (int i = 1; i <= p.ng; ++i) o[(long long)(i - 1) * p.nlines + t] = a[AT(b, c0 + i)];
}
__global__ void k_unpack2(LineP p, Arr3 A, int c0a, int c0b, const double* bufa, const double* bufb, int add, int on_a, int on_b) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.nlines) return;
  if (blockIdx.z ? !on_b : !on_a) return;
  double* a = A.a[blockIdx.y];
  const long long b = line_base(p, t);
  const int c0 = blockIdx.z ? c0b : c0a;
  const double* o = (blockIdx.z ? bufb : bufa) + (long long)blockIdx.y * p.ng * p.nlines;
  for (int i = 1; i <= p.ng; ++i) {
    const double v = o[(long long)(i - 1) * p.nlines + t];
    if (add) a[AT(b, c0 + i)] += v; else a[AT(b, c0 + i)] = v;
  }
}
static inline dim3 line_grid2(const LineP& p, int ncomp) { dim3 g = line_grid(p, ncomp); g.z = 2; return g; }

static int32_t ensure_halo(eb200_ctx* c, size_t doubles) {
  if (c->halo_bytes >= doubles * sizeof(double)) return 0;
  if (c->halo_send) cudaFree(c->halo_send);
  if (c->halo_recv) cudaFree(c->halo_recv);
  c->halo_send = c->halo_recv = nullptr; c->halo_bytes = 0;
  EB_CUDA(c, cudaMalloc(&c->halo_send, doubles * sizeof(double)));
  EB_CUDA(c, cudaMalloc(&c->halo_recv, doubles * sizeof(double)));
  c->halo_bytes = doubles * sizeof(double);
  return 0;
}

// do_field_mpi_with_lengths: x then y then z, full extents in the other axes
static int32_t field_bc_arrays(eb200_ctx* c, Arr3 A, int ncomp);
int32_t field_bc(eb200_ctx* c, int first_comp, int ncomp) {
  Arr3 A;
  for (int k = 0; k < 3; ++k) A.a[k] = c->f.f[first_comp + (k < ncomp ? k : 0)];
  return field_bc_arrays(c, A, ncomp);
}
int32_t field_bc_array(eb200_ctx* c, double* a) {
  Arr3 A;
  for (int k = 0; k < 3; ++k) A.a[k] = a;
  return field_bc_arrays(c, A, 1);
}
static int32_t field_bc_arrays(eb200_ctx* c, Arr3 A, int ncomp) {
  ScopedPhase ph(c, "halo");
  for (int ax = 0; ax < c->ndims; ++ax) {
    const LineP p = line_params(c, ax);
    const bool per = c->bc_field[2 * ax] == EB200_BC_PERIODIC;
    // a rank fills its max-side ghosts unless it sits on a non-periodic physical boundary
    const bool fill_max = !c->max_bnd[ax] || c->bc_field[2 * ax + 1] == EB200_BC_PERIODIC;
    const bool fill_min = !c->min_bnd[ax] || c->bc_field[2 * ax] == EB200_BC_PERIODIC;
    if (c->nprocdir[ax] == 1) {
      if (!per) continue;
      k_halo_periodic<<<line_grid(p, ncomp), 128, 0, c->stream>>>(p, A, fill_max, fill_min);
      c->launches++;
    } else {
      const size_t cnt = (size_t)ncomp * p.ng * p.nlines;
      EB_TRY(ensure_halo(c, 2 * cnt));
      const int lo = fill_min ? c->nbr_lo[ax] : -1, hi = fill_max ? c->nbr_hi[ax] : -1;
      // A(1:ng) goes to the low neighbour and A(nn+1-ng:nn) to the high one; A(nn+1:nn+ng) comes from the high
      // neighbour and A(1-ng:0) from the low one: one pack, ONE exchange (both directions, all components), one unpack
      double *s_lo = c->halo_send, *s_hi = c->halo_send + cnt, *r_hi = c->halo_recv, *r_lo = c->halo_recv + cnt;
      k_pack2<<<line_grid2(p, ncomp), 128, 0, c->stream>>>(p, A, 0, p.nn - p.ng, s_lo, s_hi);
      EB_TRY(comm_exchange2(c, s_lo, lo, cnt, s_hi, hi, cnt, r_hi, hi, cnt, r_lo, lo, cnt));
      k_unpack2<<<line_grid2(p, ncomp), 128, 0, c->stream>>>(p, A, p.nn, -p.ng, r_hi, r_lo, 0, hi >= 0, lo >= 0);
      c->launches += 2;
    }
  }
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

// ---- clamp / zero-gradient ---------------------------------------------------------
// mode per component: 0 none, 1 clamp (odd mirror), 2 zero gradient (even mirror)
struct MirrorP { int mode[3]; int stg[3]; };
__global__ void k_mirror(LineP p, Arr3 A, MirrorP m, int is_max) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.nlines) return;
  const int cidx = blockIdx.y;
  const int mode = m.mode[cidx];
  if (mode == 0) return;
  double* a = A.a[cidx];
  const long long b = line_base(p, t);
  const double sgn = mode == 1 ? -1.0 : 1.0;
  const int ng = p.ng, nn = p.nn;
  if (!is_max) {
    if (m.stg[cidx]) {
      for (int i = 1; i <= ng - 1; ++i) a[AT(b, i - ng)] = sgn * a[AT(b, ng - i)];
      if (mode == 1) a[AT(b, 0)] = 0.0;
    } else {
      for (int i = 1; i <= ng; ++i) a[AT(b, i - ng)] = sgn * a[AT(b, ng + 1 - i)];
    }
  } else {
    if (m.stg[cidx]) {
      if (mode == 1) a[AT(b, nn)] = 0.0;
      for (int i = 1; i <= ng - 1; ++i) a[AT(b, nn + i)] = sgn * a[AT(b, nn - i)];
    } else {
      for (int i = 1; i <= ng; ++i) a[AT(b, nn + i)] = sgn * a[AT(b, nn + 1 - i)];
    }
  }
}

static inline bool stagger(int dir, int comp) {   // setup.F90:134-144
  if (comp < 3) return dir == comp;               // ex,ey,ez
  return dir != comp - 3;                         // bx,by,bz
}

// efield_bcs / bfield_bcs physical part.  The reference loops "conduct" boundaries
// first and then clamp / zero-gradient ones; a boundary has exactly one code so
// the per-boundary work is independent and is issued in boundary order.
static int32_t physical_bcs(eb200_ctx* c, bool efield) {
  const int first = efield ? EX : BX;
  Arr3 A;
  for (int k = 0; k < 3; ++k) A.a[k] = c->f.f[first + k];
  for (int pass = 0; pass < 2; ++pass)
    for (int b = 0; b < 2 * c->ndims; ++b) {
      const int ax = b / 2, is_max = b & 1;
      if (is_max ? !c->max_bnd[ax] : !c->min_bnd[ax]) continue;
      const int bc = c->bc_field[b];
      MirrorP m;
      bool any = false;
      for (int k = 0; k < 3; ++k) {
        m.stg[k] = stagger(ax, first + k);
        m.mode[k] = 0;
        if (pass == 0 && bc == EB200_BC_CONDUCT) {
          // E: normal comp clamped, tangential zero-gradient; B: the opposite
          const bool normal = k == ax;
          m.mode[k] = efield ? (normal ? 1 : 2) : (normal ? 2 : 1);
        }
        if (pass == 1 && (bc == EB200_BC_CLAMP || bc == EB200_BC_SIMPLE_LASER || bc == EB200_BC_SIMPLE_OUTFLOW)) m.mode[k] = 1;
        if (pass == 1 && (bc == EB200_BC_ZERO_GRADIENT || bc == EB200_BC_CPML_LASER || bc == EB200_BC_CPML_OUTFLOW)) m.mode[k] = 2;
        any = any || m.mode[k] != 0;
      }
      if (!any) continue;
      const LineP p = line_params(c, ax);
      k_mirror<<<line_grid(p, 3), 128, 0, c->stream>>>(p, A, m, is_max);
      c->launches++;
    }
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

// field_zero_gradient(array, c_stagger_centre, b) on every boundary (the tail of the calc_df routines)
int32_t zero_gradient_centre(eb200_ctx* c, double* a) {
  Arr3 A; A.a[0] = A.a[1] = A.a[2] = a;
  MirrorP m;
  for (int k = 0; k < 3; ++k) { m.mode[k] = k == 0 ? 2 : 0; m.stg[k] = 0; }
  for (int b = 0; b < 2 * c->ndims; ++b) {
    const int ax = b / 2, is_max = b & 1;
    if (c->bc_field[b] == EB200_BC_PERIODIC) continue;
    if (is_max ? !c->max_bnd[ax] : !c->min_bnd[ax]) continue;
    const LineP p = line_params(c, ax);
    k_mirror<<<line_grid(p, 1), 128, 0, c->stream>>>(p, A, m, is_max);
    c->launches++;
  }
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

int32_t efield_bcs(eb200_ctx* c) {
  EB_TRY(field_bc(c, EX, 3));
  ScopedPhase ph(c, "bcs");
  return physical_bcs(c, true);
}
int32_t bfield_bcs(eb200_ctx* c, bool mpi_only) {
  EB_TRY(field_bc(c, BX, 3));
  if (mpi_only) return 0;
  ScopedPhase ph(c, "bcs");
  return physical_bcs(c, false);
}

// ---- laser / outflow planes ---------------------------------------------------------
constexpr int MAX_LASERS = 8;
struct OutflowP {
  int a, t1, t2, is_max, lp, nd;
  int nlo, nhi, lo_axis, hi_axis;
  double l1, l2, sum, diff, dt_eps;
  int nl;
  double amp[MAX_LASERS], ph0[MAX_LASERS], cpol[MAX_LASERS], spol[MAX_LASERS];
  const double* profile[MAX_LASERS];
  const double* phase[MAX_LASERS];
  const double* planes;      // six boundary planes ex..bz of setup_field_boundaries (or null = zero)
  long long plane_cnt;
  int ext_lo;                // extent of the plane's fast axis incl. ghosts
};
__global__ void k_outflow(GridP g, FieldP F, OutflowP p) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)p.nlo * p.nhi) return;
  const int pl = (int)(t % p.nlo), q = (int)(t / p.nlo);
  int cc[3] = {1, 1, 1};
  cc[p.a] = p.lp;
  if (p.lo_axis < p.nd) cc[p.lo_axis] = pl;
  if (p.hi_axis < p.nd) cc[p.hi_axis] = q;
  const long long i = g.idx(cc[0], cc[1], cc[2]);
  const long long sa = g.st[p.a];
  double* Ba = F.f[BX + p.a]; double* B1 = F.f[BX + p.t1]; double* B2 = F.f[BX + p.t2];
  const double* E1 = F.f[EX + p.t1]; const double* E2 = F.f[EX + p.t2];
  const double* J1 = F.f[JX + p.t1]; const double* J2 = F.f[JX + p.t2];
  double source1 = 0.0, source2 = 0.0;
  for (int l = 0; l < p.nl; ++l) {
    const double prof = p.profile[l] ? p.profile[l][t] : 1.0;
    const double ph = p.phase[l] ? p.phase[l][t] : 0.0;
    const double base = p.amp[l] * prof * sin(p.ph0[l] + ph);
    source1 = source1 + base * p.cpol[l];
    source2 = source2 + base * p.spol[l];
  }
  // boundary planes (laser.f90:536,559-572): the normal B is set to its plane, the tangential pairs add
  // 2 (E_plane +- c B_plane)
  double PBa = 0.0, PE1 = 0.0, PE2 = 0.0, PB1 = 0.0, PB2 = 0.0;
  if (p.planes) {
    const long long l = (long long)(cc[p.lo_axis] + g.ngd[p.lo_axis] - 1) + (long long)p.ext_lo * (cc[p.hi_axis] + g.ngd[p.hi_axis] - 1);
    PBa = p.planes[(BX + p.a) * p.plane_cnt + l];
    PE1 = p.planes[(EX + p.t1) * p.plane_cnt + l]; PE2 = p.planes[(EX + p.t2) * p.plane_cnt + l];
    PB1 = p.planes[(BX + p.t1) * p.plane_cnt + l]; PB2 = p.planes[(BX + p.t2) * p.plane_cnt + l];
  }
  if (!p.is_max) {
    Ba[i - sa] = PBa;
    double v2 = 4.0 * source1 + 2.0 * (PE1 + C0 * PB2);
    v2 = v2 - 2.0 * E1[i];
    if (p.t2 < p.nd) v2 = v2 - p.l2 * (Ba[i] - Ba[i - g.st[p.t2]]);
    v2 = v2 + p.dt_eps * J1[i] + p.diff * B2[i];
    B2[i - sa] = p.sum * v2;
    double v1 = -4.0 * source2 - 2.0 * (PE2 - C0 * PB1);
    v1 = v1 + 2.0 * E2[i];
    if (p.t1 < p.nd) v1 = v1 - p.l1 * (Ba[i] - Ba[i - g.st[p.t1]]);
    v1 = v1 - p.dt_eps * J2[i] + p.diff * B1[i];
    B1[i - sa] = p.sum * v1;
  } else {
    Ba[i + sa] = PBa;
    double v2 = -4.0 * source1 - 2.0 * (PE1 - C0 * PB2);
    v2 = v2 + 2.0 * E1[i];
    if (p.t2 < p.nd) v2 = v2 + p.l2 * (Ba[i] - Ba[i - g.st[p.t2]]);
    v2 = v2 - p.dt_eps * J1[i] + p.diff * B2[i - sa];
    B2[i] = p.sum * v2;
    double v1 = 4.0 * source2 + 2.0 * (PE2 + C0 * PB1);
    v1 = v1 - 2.0 * E2[i];
    if (p.t1 < p.nd) v1 = v1 + p.l1 * (Ba[i] - Ba[i - g.st[p.t1]]);
    v1 = v1 + p.dt_eps * J2[i] + p.diff * B1[i - sa];
    B1[i] = p.sum * v1;
  }
}

static int32_t outflow_bcs(eb200_ctx* c, int b) {
  const GridP& g = c->g;
  OutflowP p;
  p.a = b / 2; p.t1 = (p.a + 1) % 3; p.t2 = (p.a + 2) % 3; p.is_max = b & 1; p.nd = c->ndims;
  p.lp = p.is_max ? g.n[p.a] : 1;
  if (c->bc_field[b] == EB200_BC_CPML_LASER) p.lp = c->cpml_laser_idx[b];     // laser.f90:529-531
  const double dtc2 = c->dt * (C0 * C0);
  const double la = dtc2 / g.d[p.a];
  p.l1 = p.t1 < c->ndims ? dtc2 / g.d[p.t1] : 0.0;
  p.l2 = p.t2 < c->ndims ? dtc2 / g.d[p.t2] : 0.0;
  p.sum = 1.0 / (la + C0); p.diff = la - C0; p.dt_eps = c->dt / EPS0;
  p.lo_axis = p.t1 < p.t2 ? p.t1 : p.t2; p.hi_axis = p.t1 < p.t2 ? p.t2 : p.t1;
  p.nlo = p.lo_axis < c->ndims ? g.n[p.lo_axis] + 1 : 1;
  p.nhi = p.hi_axis < c->ndims ? g.n[p.hi_axis] + 1 : 1;
  p.nl = 0;
  if (c->add_laser[b])
    for (const Laser& L : c->lasers[b]) {
      if (!(c->time >= L.t_start && c->time <= L.t_end)) continue;   // laser.f90:542
      if (p.nl >= MAX_LASERS) return fail(c, EB200_E_UNSUPPORTED, "more than 8 active lasers on one boundary");
      p.amp[p.nl] = L.t_env * L.amp; p.ph0[p.nl] = L.integral_phase;
      p.cpol[p.nl] = cos(L.pol); p.spol[p.nl] = sin(L.pol);
      p.profile[p.nl] = L.profile; p.phase[p.nl] = L.phase;
      p.nl++;
    }
  if (c->bnd_dropped[b])
    return fail(c, EB200_E_STATE, "the boundary planes of this side were dropped by redistribute: call set_boundary_fields");
  p.planes = c->bnd_planes[b];
  p.ext_lo = g.n[p.lo_axis] + 2 * g.ngd[p.lo_axis];
  p.plane_cnt = g.total / (g.n[p.a] + 2 * g.ngd[p.a]);
  const long long cnt = (long long)p.nlo * p.nhi;
  k_outflow<<<(unsigned)((cnt + 127) / 128), 128, 0, c->stream>>>(c->g, c->f, p);
  c->launches++;
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

// setup_field_boundaries (housekeeping/setup.F90:426-510): thread per point of the plane normal to axis a
// (other axes incl. ghosts, lower axis fastest); n0 = the boundary plane (1 / n, or one beyond the laser plane)
__global__ void k_setup_bnd(GridP g, FieldP F, int a, int n0, double* planes, long long cnt) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= cnt) return;
  const int lo = a == 0 ? 1 : 0, hi = a == 2 ? 1 : 2;
  const int ext_lo = g.n[lo] + 2 * g.ngd[lo];
  int cc[3];
  cc[a] = n0; cc[lo] = 1 - g.ngd[lo] + (int)(t % ext_lo); cc[hi] = 1 - g.ngd[hi] + (int)(t / ext_lo);
  const long long i = g.idx(cc[0], cc[1], cc[2]);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const double* E = F.f[EX + k]; const double* B = F.f[BX + k];
    planes[(EX + k) * cnt + t] = k == a ? 0.5 * (E[i] + E[i - g.st[a]]) : E[i];
    planes[(BX + k) * cnt + t] = k == a ? B[i] : 0.5 * (B[i] + B[i - g.st[a]]);
  }
}
int32_t setup_field_boundaries(eb200_ctx* c) {
  const GridP& g = c->g;
  for (int b = 0; b < 2 * c->ndims; ++b) {
    const int a = b / 2;
    const bool is_max = b & 1;
    int n0 = is_max ? g.n[a] : 1;
    if (c->bc_field[b] == EB200_BC_CPML_LASER) n0 = is_max ? c->cpml_laser_idx[b] + 1 : c->cpml_laser_idx[b] - 1;
    const long long cnt = g.total / (g.n[a] + 2 * g.ngd[a]);
    if (!c->bnd_planes[b]) EB_CUDA(c, cudaMalloc(&c->bnd_planes[b], 6 * cnt * sizeof(double)));
    c->bnd_dropped[b] = false;
    k_setup_bnd<<<(unsigned)((cnt + 127) / 128), 128, 0, c->stream>>>(c->g, c->f, a, n0, c->bnd_planes[b], cnt);
    c->launches++;
  }
  EB_CUDA(c, cudaGetLastError());
  return 0;
}
int32_t set_boundary_fields(eb200_ctx* c, int b, const double* const planes[6]) {
  const GridP& g = c->g;
  const int a = b / 2;
  const size_t cnt = (size_t)(g.total / (g.n[a] + 2 * g.ngd[a]));
  if (!c->bnd_planes[b]) EB_CUDA(c, cudaMalloc(&c->bnd_planes[b], 6 * cnt * sizeof(double)));
  c->bnd_dropped[b] = false;
  EB_CUDA(c, cudaMemsetAsync(c->bnd_planes[b], 0, 6 * cnt * sizeof(double), c->stream));
  for (int k = 0; k < 6; ++k)
    if (planes[k])
      EB_CUDA(c, cudaMemcpyAsync(c->bnd_planes[b] + k * cnt, planes[k], cnt * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int32_t bfield_final_bcs(eb200_ctx* c) {
  EB_TRY(bfield_bcs(c, false));
  {
    ScopedPhase ph(c, "bcs");
    for (int b = 0; b < 2 * c->ndims; ++b) {
      const int ax = b / 2;
      if ((b & 1) ? !c->max_bnd[ax] : !c->min_bnd[ax]) continue;
      if (c->add_laser[b] || c->bc_field[b] == EB200_BC_SIMPLE_OUTFLOW) EB_TRY(outflow_bcs(c, b));
    }
  }
  return bfield_bcs(c, true);
}

// ---- current ghost-sum ---------------------------------------------------------------
__global__ void k_reflect_sum(LineP p, double* a, int rmin, int rmax, int flip) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.nlines) return;
  const long long b = line_base(p, t);
  const int ng = p.ng, nn = p.nn;
  if (rmin) {
    if (flip) for (int i = 1; i <= ng - 1; ++i) { a[AT(b, i)] = a[AT(b, i)] - a[AT(b, -i)]; a[AT(b, -i)] = 0.0; }
    else for (int i = 1; i <= ng - 1; ++i) { a[AT(b, i)] = a[AT(b, i)] + a[AT(b, 1 - i)]; a[AT(b, 1 - i)] = 0.0; }
  }
  if (rmax) {
    if (flip) for (int i = 1; i <= ng; ++i) { a[AT(b, nn - i)] = a[AT(b, nn - i)] - a[AT(b, nn + i)]; a[AT(b, nn + i)] = 0.0; }
    else for (int i = 1; i <= ng; ++i) { a[AT(b, nn + 1 - i)] = a[AT(b, nn + 1 - i)] + a[AT(b, nn + i)]; a[AT(b, nn + i)] = 0.0; }
  }
}
__global__ void k_periodic_sum(LineP p, double* a) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.nlines) return;
  const long long b = line_base(p, t);
  const int ng = p.ng, nn = p.nn;
  for (int i = 1; i <= ng; ++i) a[AT(b, i)] = a[AT(b, i)] + a[AT(b, nn + i)];
  for (int i = 1; i <= ng; ++i) a[AT(b, nn - ng + i)] = a[AT(b, nn - ng + i)] + a[AT(b, i - ng)];
}

// particle_clear_bcs (boundary.F90:1088-1119): the ghost zones of one axis
__global__ void k_clear_ghost_lines(LineP p, double* a) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.nlines) return;
  const long long b = line_base(p, t);
  for (int i = 1; i <= p.ng; ++i) { a[AT(b, 1 - i)] = 0.0; a[AT(b, p.nn + i)] = 0.0; }
}

static int bc_allspecies_of(const eb200_ctx* c, int b) {
  int bc = -1;
  for (const SpeciesD& sp : c->species) if (sp.defined) { bc = sp.bc[b]; break; }
  return bc < 0 ? EB200_BC_PERIODIC : bc;
}
// ANY(bc_allspecies == c_bc_mixed) (deck_species_block.F90:231-245)
bool bc_mixed(const eb200_ctx* c) {
  for (int b = 0; b < 2 * c->ndims; ++b) {
    int bc = -1;
    for (const SpeciesD& sp : c->species) {
      if (!sp.defined) continue;
      if (bc < 0) bc = sp.bc[b];
      else if (bc != sp.bc[b]) return true;
    }
  }
  return false;
}

// processor_summation_bcs (boundary.F90:1123-1144) on one array.  bc6: the species form
// (current_bcs(species) with mixed per-species boundaries): that species' codes, ghosts cleared afterwards.
// processor_summation_bcs for ncomp arrays at once (flip[k] = the wall-normal axis whose reflection flips the sign of array k)
static int32_t summation_bcs_n(eb200_ctx* c, Arr3 A, int ncomp, const int* flip, const int* bc6) {
  auto bc_allspecies = [bc6](const eb200_ctx* cc, int b) { return bc6 ? bc6[b] : bc_allspecies_of(cc, b); };
  for (int ax = 0; ax < c->ndims; ++ax) {
    const bool rmin = c->min_bnd[ax] && bc_allspecies(c, 2 * ax) == EB200_BC_REFLECT;
    const bool rmax = c->max_bnd[ax] && bc_allspecies(c, 2 * ax + 1) == EB200_BC_REFLECT;
    if (!rmin && !rmax) continue;
    const LineP p = line_params(c, ax);
    for (int k = 0; k < ncomp; ++k) {
      k_reflect_sum<<<line_grid(p, 1), 128, 0, c->stream>>>(p, A.a[k], rmin, rmax, flip[k] == ax);
      c->launches++;
    }
  }
  for (int ax = 0; ax < c->ndims; ++ax) {
    const LineP p = line_params(c, ax);
    // neighbour_local: no transfer through a non-periodic physical boundary
    const bool has_lo = !(c->min_bnd[ax] && bc_allspecies(c, 2 * ax) != EB200_BC_PERIODIC);
    const bool has_hi = !(c->max_bnd[ax] && bc_allspecies(c, 2 * ax + 1) != EB200_BC_PERIODIC);
    if (c->nprocdir[ax] == 1) {
      if (has_lo && has_hi)
        for (int k = 0; k < ncomp; ++k) { k_periodic_sum<<<line_grid(p, 1), 128, 0, c->stream>>>(p, A.a[k]); c->launches++; }
    } else {
      const size_t cnt = (size_t)ncomp * p.ng * p.nlines;
      EB_TRY(ensure_halo(c, 2 * cnt));
      const int lo = has_lo ? c->nbr_lo[ax] : -1, hi = has_hi ? c->nbr_hi[ax] : -1;
      // the ghost slab A(nn+1:nn+ng) goes to +1 and A(1-ng:0) to -1; what comes from -1 is added into A(1:ng), what
      // comes from +1 into A(nn+1-ng:nn) (boundary.F90:976-989): one exchange for both directions and all arrays
      double *s_hi = c->halo_send, *s_lo = c->halo_send + cnt, *r_lo = c->halo_recv, *r_hi = c->halo_recv + cnt;
      k_pack2<<<line_grid2(p, ncomp), 128, 0, c->stream>>>(p, A, p.nn, -p.ng, s_hi, s_lo);
      EB_TRY(comm_exchange2(c, s_hi, hi, cnt, s_lo, lo, cnt, r_lo, lo, cnt, r_hi, hi, cnt));
      if (p.nn >= 2 * p.ng) {
        k_unpack2<<<line_grid2(p, ncomp), 128, 0, c->stream>>>(p, A, 0, p.nn - p.ng, r_lo, r_hi, 1, lo >= 0, hi >= 0);
      } else {
        // a sub-domain thinner than two ghost widths: A(1:ng) and A(nn+1-ng:nn) overlap, so the two additions
        // must not run side by side -- one after the other, in the reference's order (boundary.F90:976-989)
        k_unpack2<<<line_grid2(p, ncomp), 128, 0, c->stream>>>(p, A, 0, p.nn - p.ng, r_lo, r_hi, 1, lo >= 0, 0);
        k_unpack2<<<line_grid2(p, ncomp), 128, 0, c->stream>>>(p, A, 0, p.nn - p.ng, r_lo, r_hi, 1, 0, hi >= 0);
        c->launches++;
      }
      c->launches += 2;
    }
  }
  if (bc6) {
    for (int ax = 0; ax < c->ndims; ++ax) {
      const LineP p = line_params(c, ax);
      for (int k = 0; k < ncomp; ++k) { k_clear_ghost_lines<<<line_grid(p, 1), 128, 0, c->stream>>>(p, A.a[k]); c->launches++; }
    }
  }
  EB_CUDA(c, cudaGetLastError());
  return 0;
}
int32_t summation_bcs(eb200_ctx* c, double* arr, int flip_dir, const int* bc6) {
  Arr3 A; A.a[0] = A.a[1] = A.a[2] = arr;
  const int flip[3] = {flip_dir, flip_dir, flip_dir};
  return summation_bcs_n(c, A, 1, flip, bc6);
}

// ---- smooth_current (current_smooth.F90:49-151, the strided compensated binomial filter) ----
// one pass: interior of dst = alpha * src + beta * (the 2 * ND neighbours of src at +-stride),
// neighbours summed in the reference's order (x-, x+, y-, y+, z-, z+)
template <int ND>
__global__ void __launch_bounds__(256) k_smooth_pass(GridP g, Arr3 src, Arr3 dst, double alpha, double beta, int stride,
                                                     long long count) {
  const long long cs = stride;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < count; t += (long long)gridDim.x * blockDim.x) {
    const int ix = 1 + (int)(t % g.n[0]);
    const long long r = t / g.n[0];
    const int iy = ND >= 2 ? 1 + (int)(r % g.n[1]) : 1;
    const int iz = ND >= 3 ? 1 + (int)(r / g.n[1]) : 1;
    const long long i = g.idx(ix, iy, iz);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double* w = src.a[k];
      double nb = w[i - cs * g.st[0]] + w[i + cs * g.st[0]];
      if (ND >= 2) { nb = nb + w[i - cs * g.st[1]]; nb = nb + w[i + cs * g.st[1]]; }
      if (ND >= 3) { nb = nb + w[i - cs * g.st[2]]; nb = nb + w[i + cs * g.st[2]]; }
      dst.a[k][i] = __dadd_rn(__dmul_rn(alpha, w[i]), __dmul_rn(nb, beta));
    }
  }
}
template <int ND>
__global__ void __launch_bounds__(256) k_copy_interior(GridP g, Arr3 src, Arr3 dst, long long count) {
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < count; t += (long long)gridDim.x * blockDim.x) {
    const int ix = 1 + (int)(t % g.n[0]);
    const long long r = t / g.n[0];
    const int iy = ND >= 2 ? 1 + (int)(r % g.n[1]) : 1;
    const int iz = ND >= 3 ? 1 + (int)(r / g.n[1]) : 1;
    const long long i = g.idx(ix, iy, iz);
#pragma unroll
    for (int k = 0; k < 3; ++k) dst.a[k][i] = src.a[k][i];
  }
}
__global__ void k_sub_scalar(double* a, long long n, double v) {
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x)
    a[t] = a[t] - v;
}

static int32_t smooth_current(eb200_ctx* c) {
  const long long total = c->g.total;
  if (!c->smooth_wk) EB_CUDA(c, cudaMalloc(&c->smooth_wk, (size_t)3 * total * sizeof(double)));
  Arr3 J, W;
  for (int k = 0; k < 3; ++k) { J.a[k] = c->f.f[JX + k]; W.a[k] = c->smooth_wk + (size_t)k * total; }
  // wk = array, ghosts included (ng_l = MAX(sng, jng) = jng since strides <= ng)
  EB_CUDA(c, cudaMemcpyAsync(c->smooth_wk, c->f.f[JX], (size_t)3 * total * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  long long count = c->g.n[0];
  if (c->ndims >= 2) count *= c->g.n[1];
  if (c->ndims >= 3) count *= c->g.n[2];
  const int block = 256;
  const int grid = (int)std::min<long long>((count + block - 1) / block, 148LL * 32);
  double alpha = 0.5;
  const double beta = (1.0 - alpha) / (2.0 * c->ndims);     // computed once, as in the reference
  for (int it = 1; it <= c->smooth_its + c->smooth_comp_its; ++it) {
    for (int stride : c->smooth_strides) {
      EB_TRY(field_bc_arrays(c, W, 3));
#define L(ND)                                                                                              \
      if (c->ndims == ND) {                                                                                \
        k_smooth_pass<ND><<<grid, block, 0, c->stream>>>(c->g, W, J, alpha, beta, stride, count);           \
        k_copy_interior<ND><<<grid, block, 0, c->stream>>>(c->g, J, W, count);                             \
      }
      L(1) L(2) L(3)
#undef L
      c->launches += 2;
    }
    // the compensation weight takes effect only for passes after its + 1 (reference behaviour)
    if (it > c->smooth_its) alpha = (double)c->smooth_its * 0.5 + 1.0;
  }
  // array(interior) = wk(interior): already equal after the last pass
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

// current_finish (current_smooth.F90:29-45)
int32_t current_finish(eb200_ctx* c) {
  {
    ScopedPhase ph(c, "halo");
    if (!bc_mixed(c)) {       // (mixed: done species by species inside push_particles, boundary.F90:1130-1136)
      Arr3 J; J.a[0] = c->f.f[JX]; J.a[1] = c->f.f[JY]; J.a[2] = c->f.f[JZ];
      const int flip[3] = {0, 1, 2};
      EB_TRY(summation_bcs_n(c, J, 3, flip, nullptr));
    }
  }
  EB_TRY(field_bc(c, JX, 3));
  if (c->smooth) EB_TRY(smooth_current(c));
  if (c->use_current_correction) {
    ScopedPhase ph(c, "fields");
    for (int k = 0; k < 3; ++k) {
      k_sub_scalar<<<148 * 8, 256, 0, c->stream>>>(c->f.f[JX + k], c->g.total, c->initial_j[k]);
      c->launches++;
    }
    EB_CUDA(c, cudaGetLastError());
  }
  return 0;
}

}  // namespace eb
