// fdtd.cu -- Yee FDTD half-step kernels (orders 2/4/6 and the extended second-order
// stencils of set_maxwell_solver), fp64, HBM-bandwidth bound.
// Replaces update_e_field / update_b_field of epochNd/src/fields.f90 (3D :183-424,
// :428-815; non-CPML Yee branches).  One thread per cell over the reference's loop
// range 0..n in every existing dimension, x fastest so that loads/stores coalesce;
// the +-1 neighbour reads in y and z hit L1/L2 (planes of 522x522 doubles = 2.2 MB
// stay resident in the 126 MB L2), so DRAM traffic is the algorithmic 96 B (E) /
// 72 B (B) per cell.
#include "ctx.cuh"

namespace eb {

struct FdtdP {
  double cx[3], cy[3], cz[3];   // per-order coefficients c1,c2,c3 times cn* (E) or hdt* (B)
  double fac;
  double sc[15];                // extended-stencil coefficients (set_maxwell_solver, fields.f90:53-179)
  // CPML: coefficient of a cell = (base / kappa(index along the axis)) * c_k  (fields.f90:192-310, :437-630)
  double base[3], ck[3];
  const double* kap[3];         // kappa_e (E update) or kappa_b (B update), element 0 <-> index 1-ng
  int ng;
  int ex0, ex1, ey0, ey1, ez0, ez1;   // loop extents (counts) and origins
  long long count;
};

// per-cell coefficients of axis d: plain (p.cx/cy/cz) or kappa-scaled
template <int ORD, bool CPML>
__device__ __forceinline__ void axis_coef(const FdtdP& p, int d, const double* plain, int idx, double* out) {
  if (!CPML) { out[0] = plain[0]; out[1] = plain[1]; out[2] = plain[2]; return; }
  const double cp = p.base[d] / p.kap[d][idx + p.ng - 1];
  out[0] = ORD == 2 ? cp : p.ck[0] * cp;
  out[1] = p.ck[1] * cp;
  out[2] = p.ck[2] * cp;
}

template <int ND, int ORD, bool CPML>
__global__ void __launch_bounds__(256) k_update_e(GridP g, FieldP F, FdtdP p) {
  const long long sx = g.st[0], sy = g.st[1], sz = g.st[2];
  const int nx1 = g.n[0] + 1;
  const int ny1 = ND >= 2 ? g.n[1] + 1 : 1;
  double* __restrict__ ex = F.f[EX]; double* __restrict__ ey = F.f[EY]; double* __restrict__ ez = F.f[EZ];
  const double* __restrict__ bx = F.f[BX]; const double* __restrict__ by = F.f[BY]; const double* __restrict__ bz = F.f[BZ];
  const double* __restrict__ jx = F.f[JX]; const double* __restrict__ jy = F.f[JY]; const double* __restrict__ jz = F.f[JZ];
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < p.count;
       t += (long long)gridDim.x * blockDim.x) {
    const int ix = (int)(t % nx1);
    const long long r = t / nx1;
    const int iy = ND >= 2 ? (int)(r % ny1) : 1;
    const int iz = ND >= 3 ? (int)(r / ny1) : 1;
    const long long i = g.idx(ix, iy, iz);
    double cx[3], cy[3] = {0, 0, 0}, cz[3] = {0, 0, 0};
    axis_coef<ORD, CPML>(p, 0, p.cx, ix, cx);
    if (ND >= 2) axis_coef<ORD, CPML>(p, 1, p.cy, iy, cy);
    if (ND >= 3) axis_coef<ORD, CPML>(p, 2, p.cz, iz, cz);
    double vx = ex[i], vy = ey[i], vz = ez[i];
    const double bxi = bx[i], byi = by[i], bzi = bz[i];
    if (ND >= 2) {
      vx = vx + cy[0] * (bzi - bz[i - sy]);
      if (ORD >= 4) vx = vx + cy[1] * (bz[i + sy] - bz[i - 2 * sy]);
      if (ORD >= 6) vx = vx + cy[2] * (bz[i + 2 * sy] - bz[i - 3 * sy]);
    }
    if (ND >= 3) {
      vx = vx - cz[0] * (byi - by[i - sz]);
      if (ORD >= 4) vx = vx - cz[1] * (by[i + sz] - by[i - 2 * sz]);
      if (ORD >= 6) vx = vx - cz[2] * (by[i + 2 * sz] - by[i - 3 * sz]);
    }
    vx = vx - p.fac * jx[i];
    if (ND >= 3) {
      vy = vy + cz[0] * (bxi - bx[i - sz]);
      if (ORD >= 4) vy = vy + cz[1] * (bx[i + sz] - bx[i - 2 * sz]);
      if (ORD >= 6) vy = vy + cz[2] * (bx[i + 2 * sz] - bx[i - 3 * sz]);
    }
    vy = vy - cx[0] * (bzi - bz[i - sx]);
    if (ORD >= 4) vy = vy - cx[1] * (bz[i + sx] - bz[i - 2 * sx]);
    if (ORD >= 6) vy = vy - cx[2] * (bz[i + 2 * sx] - bz[i - 3 * sx]);
    vy = vy - p.fac * jy[i];
    vz = vz + cx[0] * (byi - by[i - sx]);
    if (ORD >= 4) vz = vz + cx[1] * (by[i + sx] - by[i - 2 * sx]);
    if (ORD >= 6) vz = vz + cx[2] * (by[i + 2 * sx] - by[i - 3 * sx]);
    if (ND >= 2) {
      vz = vz - cy[0] * (bxi - bx[i - sy]);
      if (ORD >= 4) vz = vz - cy[1] * (bx[i + sy] - bx[i - 2 * sy]);
      if (ORD >= 6) vz = vz - cy[2] * (bx[i + 2 * sy] - bx[i - 3 * sy]);
    }
    vz = vz - p.fac * jz[i];
    ex[i] = vx; ey[i] = vy; ez[i] = vz;
  }
}

template <int ND, int ORD, bool CPML>
__global__ void __launch_bounds__(256) k_update_b(GridP g, FieldP F, FdtdP p) {
  const long long sx = g.st[0], sy = g.st[1], sz = g.st[2];
  const int nx1 = g.n[0] + 1;
  const int ny1 = ND >= 2 ? g.n[1] + 1 : 1;
  const double* __restrict__ ex = F.f[EX]; const double* __restrict__ ey = F.f[EY]; const double* __restrict__ ez = F.f[EZ];
  double* __restrict__ bx = F.f[BX]; double* __restrict__ by = F.f[BY]; double* __restrict__ bz = F.f[BZ];
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < p.count;
       t += (long long)gridDim.x * blockDim.x) {
    const int ix = (int)(t % nx1);
    const long long r = t / nx1;
    const int iy = ND >= 2 ? (int)(r % ny1) : 1;
    const int iz = ND >= 3 ? (int)(r / ny1) : 1;
    const long long i = g.idx(ix, iy, iz);
    double cx[3], cy[3] = {0, 0, 0}, cz[3] = {0, 0, 0};
    axis_coef<ORD, CPML>(p, 0, p.cx, ix, cx);
    if (ND >= 2) axis_coef<ORD, CPML>(p, 1, p.cy, iy, cy);
    if (ND >= 3) axis_coef<ORD, CPML>(p, 2, p.cz, iz, cz);
    double vx = bx[i], vy = by[i], vz = bz[i];
    const double exi = ex[i], eyi = ey[i], ezi = ez[i];
    if (ND >= 2) {
      vx = vx - cy[0] * (ez[i + sy] - ezi);
      if (ORD >= 4) vx = vx - cy[1] * (ez[i + 2 * sy] - ez[i - sy]);
      if (ORD >= 6) vx = vx - cy[2] * (ez[i + 3 * sy] - ez[i - 2 * sy]);
    }
    if (ND >= 3) {
      vx = vx + cz[0] * (ey[i + sz] - eyi);
      if (ORD >= 4) vx = vx + cz[1] * (ey[i + 2 * sz] - ey[i - sz]);
      if (ORD >= 6) vx = vx + cz[2] * (ey[i + 3 * sz] - ey[i - 2 * sz]);
    }
    if (ND >= 3) {
      vy = vy - cz[0] * (ex[i + sz] - exi);
      if (ORD >= 4) vy = vy - cz[1] * (ex[i + 2 * sz] - ex[i - sz]);
      if (ORD >= 6) vy = vy - cz[2] * (ex[i + 3 * sz] - ex[i - 2 * sz]);
    }
    vy = vy + cx[0] * (ez[i + sx] - ezi);
    if (ORD >= 4) vy = vy + cx[1] * (ez[i + 2 * sx] - ez[i - sx]);
    if (ORD >= 6) vy = vy + cx[2] * (ez[i + 3 * sx] - ez[i - 2 * sx]);
    vz = vz - cx[0] * (ey[i + sx] - eyi);
    if (ORD >= 4) vz = vz - cx[1] * (ey[i + 2 * sx] - ey[i - sx]);
    if (ORD >= 6) vz = vz - cx[2] * (ey[i + 3 * sx] - ey[i - 2 * sx]);
    if (ND >= 2) {
      vz = vz + cy[0] * (ex[i + sy] - exi);
      if (ORD >= 4) vz = vz + cy[1] * (ex[i + 2 * sy] - ex[i - sy]);
      if (ORD >= 6) vz = vz + cy[2] * (ex[i + 3 * sy] - ex[i - 2 * sy]);
    }
    bx[i] = vx; by[i] = vy; bz[i] = vz;
  }
}

// Extended-stencil difference along axis a (Lehe / Cowan / Pukhov / custom solvers,
// fields.f90:655-732; 2D :426-470; 1D :296-306): the plain forward difference weighted by
// alpha plus the same difference taken one cell aside in the other two directions (beta),
// on the four diagonals (gamma), and the three-cell-wide difference (delta).  so1 / so2 are
// the strides of the other axes in index order; H1 / H2 say whether they exist.
template <bool H1, bool H2>
__device__ __forceinline__ double ext_diff(const double* __restrict__ f, long long i, long long sa, long long so1,
                                           long long so2, double alpha, double beta1, double beta2, double gamma,
                                           double delta) {
  double r = alpha * (f[i + sa] - f[i]);
  if (H1) r += beta1 * (f[i + sa + so1] - f[i + so1] + f[i + sa - so1] - f[i - so1]);
  if (H2) r += beta2 * (f[i + sa + so2] - f[i + so2] + f[i + sa - so2] - f[i - so2]);
  if (H1 && H2)
    r += gamma * (f[i + sa + so1 - so2] - f[i + so1 - so2] + f[i + sa - so1 - so2] - f[i - so1 - so2]
                + f[i + sa + so1 + so2] - f[i + so1 + so2] + f[i + sa - so1 + so2] - f[i - so1 + so2]);
  r += delta * (f[i + 2 * sa] - f[i - sa]);
  return r;
}

// update_b_field with an extended stencil (field_order 2 only, as in the reference).  The
// wider stencil re-reads neighbouring lines; they come from L1/L2, DRAM traffic stays 72 B/cell.
template <int ND, bool CPML>
__global__ void __launch_bounds__(256) k_update_b_ext(GridP g, FieldP F, FdtdP p) {
  constexpr int ORD = 2;
  const long long sx = g.st[0], sy = g.st[1], sz = g.st[2];
  const int nx1 = g.n[0] + 1;
  const int ny1 = ND >= 2 ? g.n[1] + 1 : 1;
  const double* __restrict__ ex = F.f[EX]; const double* __restrict__ ey = F.f[EY]; const double* __restrict__ ez = F.f[EZ];
  double* __restrict__ bx = F.f[BX]; double* __restrict__ by = F.f[BY]; double* __restrict__ bz = F.f[BZ];
  constexpr bool HY = ND >= 2, HZ = ND >= 3;
  // coefficient layout: alpha x,y,z | beta xy,xz,yx,yz,zx,zy | gamma x,y,z | delta x,y,z
  const double* q = p.sc;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < p.count;
       t += (long long)gridDim.x * blockDim.x) {
    const int ix = (int)(t % nx1);
    const long long r = t / nx1;
    const int iy = ND >= 2 ? (int)(r % ny1) : 1;
    const int iz = ND >= 3 ? (int)(r / ny1) : 1;
    const long long i = g.idx(ix, iy, iz);
    double cx[3], cy[3] = {0, 0, 0}, cz[3] = {0, 0, 0};
    axis_coef<ORD, CPML>(p, 0, p.cx, ix, cx);
    if (ND >= 2) axis_coef<ORD, CPML>(p, 1, p.cy, iy, cy);
    if (ND >= 3) axis_coef<ORD, CPML>(p, 2, p.cz, iz, cz);
    double vx = bx[i], vy = by[i], vz = bz[i];
    if (ND >= 2) vx = vx - cy[0] * ext_diff<true, HZ>(ez, i, sy, sx, sz, q[1], q[5], q[6], q[10], q[13]);
    if (ND >= 3) vx = vx + cz[0] * ext_diff<true, true>(ey, i, sz, sx, sy, q[2], q[7], q[8], q[11], q[14]);
    if (ND >= 3) vy = vy - cz[0] * ext_diff<true, true>(ex, i, sz, sx, sy, q[2], q[7], q[8], q[11], q[14]);
    vy = vy + cx[0] * ext_diff<HY, HZ>(ez, i, sx, sy, sz, q[0], q[3], q[4], q[9], q[12]);
    vz = vz - cx[0] * ext_diff<HY, HZ>(ey, i, sx, sy, sz, q[0], q[3], q[4], q[9], q[12]);
    if (ND >= 2) vz = vz + cy[0] * ext_diff<true, HZ>(ex, i, sy, sx, sz, q[1], q[5], q[6], q[10], q[13]);
    bx[i] = vx; by[i] = vy; bz[i] = vz;
  }
}

static FdtdP make_params(eb200_ctx* c, bool efield) {
  FdtdP p;
  // update_eb_fields_half / _final scalars, fields.f90:821-830
  const double hdt = 0.5 * c->dt;
  double c1 = 1, c2 = 0, c3 = 0;
  if (c->field_order == 4) { c1 = 9.0 / 8.0; c2 = -1.0 / 24.0; }
  if (c->field_order == 6) { c1 = 75.0 / 64.0; c2 = -25.0 / 384.0; c3 = 3.0 / 640.0; }
  double* cc[3] = {p.cx, p.cy, p.cz};
  for (int d = 0; d < 3; ++d) {
    const double hdtd = hdt / c->g.d[d];
    const double base = efield ? hdtd * (C0 * C0) : hdtd;
    cc[d][0] = c->field_order == 2 ? base : c1 * base;
    cc[d][1] = c2 * base;
    cc[d][2] = c3 * base;
  }
  p.fac = hdt / EPS0;
  for (int k = 0; k < 15; ++k) p.sc[k] = c->coeffs[k];
  p.ck[0] = c1; p.ck[1] = c2; p.ck[2] = c3; p.ng = c->ng;
  for (int d = 0; d < 3; ++d) {
    const double hdtd = hdt / c->g.d[d];
    p.base[d] = efield ? hdtd * (C0 * C0) : hdtd;
    const size_t len = (size_t)(c->g.n[d] + 2 * c->ng);
    p.kap[d] = c->cpml_prof[d] ? c->cpml_prof[d] + (efield ? 0 : len) : nullptr;
  }
  long long cnt = c->g.n[0] + 1;
  if (c->ndims >= 2) cnt *= c->g.n[1] + 1;
  if (c->ndims >= 3) cnt *= c->g.n[2] + 1;
  p.count = cnt;
  return p;
}

// One CTA per 256 consecutive cells, no grid-stride loop: the CTAs in flight then cover a window of about one x-y
// plane that moves through the volume in z, so the z neighbour plane a cell reads comes back from L2 a plane later.
// (A capped grid striding through the volume spread every wave over the whole box: ncu at 512^3 showed 8.2 doubles
// read per cell in the B update instead of the algorithmic 6 and 11.1 instead of 9 in the E update, at 95-100 % of
// the DRAM bandwidth -- profiles/r02_dram_512_summary.txt.)
static inline unsigned grid_for(long long count, int block) {
  long long b = (count + block - 1) / block;
  if (b > 0x7fffffffLL) b = 0x7fffffffLL;      // (the kernels keep their stride loop for boxes beyond 2^39 cells)
  if (b < 1) b = 1;
  return (unsigned)b;
}

template <bool E>
static int32_t launch(eb200_ctx* c) {
  FdtdP p = make_params(c, E);
  const int block = 256; const unsigned grid = grid_for(p.count, block);
  const bool cpml = c->cpml_on;
  if (cpml)
    for (int d = 0; d < c->ndims; ++d)
      if (!p.kap[d]) return fail(c, EB200_E_STATE, "CPML boundaries: set_cpml has not been called for every axis");
#define L(ND, ORD)                                                                        \
  if (c->ndims == ND && c->field_order == ORD) {                                          \
    if (E) { if (cpml) k_update_e<ND, ORD, true><<<grid, block, 0, c->stream>>>(c->g, c->f, p);    \
             else k_update_e<ND, ORD, false><<<grid, block, 0, c->stream>>>(c->g, c->f, p); }      \
    else { if (cpml) k_update_b<ND, ORD, true><<<grid, block, 0, c->stream>>>(c->g, c->f, p);      \
           else k_update_b<ND, ORD, false><<<grid, block, 0, c->stream>>>(c->g, c->f, p); }        \
  }
#define LX(ND)                                                                            \
  if (c->ndims == ND) {                                                                   \
    if (cpml) k_update_b_ext<ND, true><<<grid, block, 0, c->stream>>>(c->g, c->f, p);     \
    else k_update_b_ext<ND, false><<<grid, block, 0, c->stream>>>(c->g, c->f, p);         \
  }
  if (!E && c->maxwell_solver != 0 && c->field_order == 2) {
    // fields.f90:637,655: the extended stencils exist in the second-order B update only
    LX(1) LX(2) LX(3)
  } else {
    L(1, 2) L(1, 4) L(1, 6) L(2, 2) L(2, 4) L(2, 6) L(3, 2) L(3, 4) L(3, 6)
  }
#undef L
#undef LX
  c->launches++;
  EB_CUDA(c, cudaGetLastError());
  // cpml_advance_e_currents / cpml_advance_b_currents(hdt) close the CPML branch of the update
  if (cpml) EB_TRY(cpml_advance_currents(c, E, 0.5 * c->dt));
  return 0;
}

// ---- CPML auxiliary currents (boundary.F90 cpml_advance_e_currents / _b_currents) --------------
// One launch per CPML side.  Axis a, b = a + 1, c = a + 2 (cyclic):
//   E: psi_e{b}{a} <- d_a B_c, E_b -= fac * psi;   psi_e{c}{a} <- d_a B_b, E_c += fac * psi
//   B: psi_b{b}{a} <- d_a E_c, B_b += t * psi;     psi_b{c}{a} <- d_a E_b, B_c -= t * psi
// over slab indices [start, end] along a and the interior 1..n of the other axes.
struct CpmlP {
  int a, start, nslab, efield, ng;
  int n1, n2, ax1, ax2;              // the other two axes (index order) and their interior extents
  double tstep, fac, inv_d;
  const double *kappa, *sigma, *acoef;
  double *psi0, *psi1, *F0, *F1;
  const double *S0, *S1;
  long long count;
};
__global__ void __launch_bounds__(128) k_cpml_currents(GridP g, CpmlP p) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.count) return;
  // the x-most of (slab index, other axes) varies fastest
  int cc[3] = {1, 1, 1};
  long long r = t;
  for (int d = 0; d < 3; ++d) {
    int ext;
    if (d == p.a) ext = p.nslab; else if (d == p.ax1) ext = p.n1; else ext = p.n2;
    const int v = (int)(r % ext); r /= ext;
    cc[d] = d == p.a ? p.start + v : 1 + v;
  }
  const int ipos = cc[p.a];
  const long long i = g.idx(cc[0], cc[1], cc[2]);
  const long long sa = g.st[p.a];
  const double kappa = p.kappa[ipos + p.ng - 1], sigma = p.sigma[ipos + p.ng - 1], acoeff = p.acoef[ipos + p.ng - 1];
  const double bcoeff = exp(-(sigma / kappa + acoeff) * p.tstep);
  const double ccoeff_d = (bcoeff - 1.0) * sigma / kappa / (sigma + kappa * acoeff) * p.inv_d;
  if (p.efield) {
    const double q0 = bcoeff * p.psi0[i] + ccoeff_d * (p.S0[i] - p.S0[i - sa]);
    p.psi0[i] = q0; p.F0[i] = p.F0[i] - p.fac * q0;
    const double q1 = bcoeff * p.psi1[i] + ccoeff_d * (p.S1[i] - p.S1[i - sa]);
    p.psi1[i] = q1; p.F1[i] = p.F1[i] + p.fac * q1;
  } else {
    const double q0 = bcoeff * p.psi0[i] + ccoeff_d * (p.S0[i + sa] - p.S0[i]);
    p.psi0[i] = q0; p.F0[i] = p.F0[i] + p.tstep * q0;
    const double q1 = bcoeff * p.psi1[i] + ccoeff_d * (p.S1[i + sa] - p.S1[i]);
    p.psi1[i] = q1; p.F1[i] = p.F1[i] - p.tstep * q1;
  }
}

int32_t cpml_advance_currents(eb200_ctx* c, bool efield, double tstep) {
  const GridP& g = c->g;
  for (int a = 0; a < c->ndims; ++a)
    for (int side = 0; side < 2; ++side) {
      const int bnd = 2 * a + side;
      if (c->bc_field[bnd] != EB200_BC_CPML_LASER && c->bc_field[bnd] != EB200_BC_CPML_OUTFLOW) continue;
      int start = c->cpml_start[bnd], end = c->cpml_end[bnd];
      if (!efield && side == 1) { start -= 1; end -= 1; }      // B on the max side: start-1 .. end-1
      if (end < start) continue;
      const int b = (a + 1) % 3, cdir = (a + 2) % 3;
      CpmlP p;
      p.a = a; p.start = start; p.nslab = end - start + 1; p.efield = efield; p.ng = c->ng;
      p.ax1 = a == 0 ? 1 : 0; p.ax2 = a == 2 ? 1 : 2;
      p.n1 = p.ax1 < c->ndims ? g.n[p.ax1] : 1; p.n2 = p.ax2 < c->ndims ? g.n[p.ax2] : 1;
      p.tstep = tstep; p.fac = tstep * (C0 * C0); p.inv_d = 1.0 / g.d[a];
      const size_t len = (size_t)(g.n[a] + 2 * c->ng);
      const double* prof = c->cpml_prof[a];
      p.kappa = prof + (efield ? 0 : len); p.sigma = prof + (efield ? 2 : 3) * len; p.acoef = prof + (efield ? 4 : 5) * len;
      double* psi = c->cpml_psi + (size_t)(4 * a) * g.total;
      if (efield) {
        p.psi0 = psi; p.psi1 = psi + g.total;
        p.F0 = c->f.f[EX + b]; p.F1 = c->f.f[EX + cdir];
        p.S0 = c->f.f[BX + cdir]; p.S1 = c->f.f[BX + b];
      } else {
        p.psi0 = psi + 2 * g.total; p.psi1 = psi + 3 * g.total;
        p.F0 = c->f.f[BX + b]; p.F1 = c->f.f[BX + cdir];
        p.S0 = c->f.f[EX + cdir]; p.S1 = c->f.f[EX + b];
      }
      p.count = (long long)p.nslab * p.n1 * p.n2;
      k_cpml_currents<<<(unsigned)((p.count + 127) / 128), 128, 0, c->stream>>>(c->g, p);
      c->launches++;
    }
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

int32_t update_e_field(eb200_ctx* c) { ScopedPhase ph(c, "fields"); return launch<true>(c); }
int32_t update_b_field(eb200_ctx* c) { ScopedPhase ph(c, "fields"); return launch<false>(c); }

}  // namespace eb
