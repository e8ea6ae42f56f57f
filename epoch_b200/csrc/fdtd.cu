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
  int ex0, ex1, ey0, ey1, ez0, ez1;   // loop extents (counts) and origins
  long long count;
};

template <int ND, int ORD>
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
    double vx = ex[i], vy = ey[i], vz = ez[i];
    const double bxi = bx[i], byi = by[i], bzi = bz[i];
    if (ND >= 2) {
      vx = vx + p.cy[0] * (bzi - bz[i - sy]);
      if (ORD >= 4) vx = vx + p.cy[1] * (bz[i + sy] - bz[i - 2 * sy]);
      if (ORD >= 6) vx = vx + p.cy[2] * (bz[i + 2 * sy] - bz[i - 3 * sy]);
    }
    if (ND >= 3) {
      vx = vx - p.cz[0] * (byi - by[i - sz]);
      if (ORD >= 4) vx = vx - p.cz[1] * (by[i + sz] - by[i - 2 * sz]);
      if (ORD >= 6) vx = vx - p.cz[2] * (by[i + 2 * sz] - by[i - 3 * sz]);
    }
    vx = vx - p.fac * jx[i];
    if (ND >= 3) {
      vy = vy + p.cz[0] * (bxi - bx[i - sz]);
      if (ORD >= 4) vy = vy + p.cz[1] * (bx[i + sz] - bx[i - 2 * sz]);
      if (ORD >= 6) vy = vy + p.cz[2] * (bx[i + 2 * sz] - bx[i - 3 * sz]);
    }
    vy = vy - p.cx[0] * (bzi - bz[i - sx]);
    if (ORD >= 4) vy = vy - p.cx[1] * (bz[i + sx] - bz[i - 2 * sx]);
    if (ORD >= 6) vy = vy - p.cx[2] * (bz[i + 2 * sx] - bz[i - 3 * sx]);
    vy = vy - p.fac * jy[i];
    vz = vz + p.cx[0] * (byi - by[i - sx]);
    if (ORD >= 4) vz = vz + p.cx[1] * (by[i + sx] - by[i - 2 * sx]);
    if (ORD >= 6) vz = vz + p.cx[2] * (by[i + 2 * sx] - by[i - 3 * sx]);
    if (ND >= 2) {
      vz = vz - p.cy[0] * (bxi - bx[i - sy]);
      if (ORD >= 4) vz = vz - p.cy[1] * (bx[i + sy] - bx[i - 2 * sy]);
      if (ORD >= 6) vz = vz - p.cy[2] * (bx[i + 2 * sy] - bx[i - 3 * sy]);
    }
    vz = vz - p.fac * jz[i];
    ex[i] = vx; ey[i] = vy; ez[i] = vz;
  }
}

template <int ND, int ORD>
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
    double vx = bx[i], vy = by[i], vz = bz[i];
    const double exi = ex[i], eyi = ey[i], ezi = ez[i];
    if (ND >= 2) {
      vx = vx - p.cy[0] * (ez[i + sy] - ezi);
      if (ORD >= 4) vx = vx - p.cy[1] * (ez[i + 2 * sy] - ez[i - sy]);
      if (ORD >= 6) vx = vx - p.cy[2] * (ez[i + 3 * sy] - ez[i - 2 * sy]);
    }
    if (ND >= 3) {
      vx = vx + p.cz[0] * (ey[i + sz] - eyi);
      if (ORD >= 4) vx = vx + p.cz[1] * (ey[i + 2 * sz] - ey[i - sz]);
      if (ORD >= 6) vx = vx + p.cz[2] * (ey[i + 3 * sz] - ey[i - 2 * sz]);
    }
    if (ND >= 3) {
      vy = vy - p.cz[0] * (ex[i + sz] - exi);
      if (ORD >= 4) vy = vy - p.cz[1] * (ex[i + 2 * sz] - ex[i - sz]);
      if (ORD >= 6) vy = vy - p.cz[2] * (ex[i + 3 * sz] - ex[i - 2 * sz]);
    }
    vy = vy + p.cx[0] * (ez[i + sx] - ezi);
    if (ORD >= 4) vy = vy + p.cx[1] * (ez[i + 2 * sx] - ez[i - sx]);
    if (ORD >= 6) vy = vy + p.cx[2] * (ez[i + 3 * sx] - ez[i - 2 * sx]);
    vz = vz - p.cx[0] * (ey[i + sx] - eyi);
    if (ORD >= 4) vz = vz - p.cx[1] * (ey[i + 2 * sx] - ey[i - sx]);
    if (ORD >= 6) vz = vz - p.cx[2] * (ey[i + 3 * sx] - ey[i - 2 * sx]);
    if (ND >= 2) {
      vz = vz + p.cy[0] * (ex[i + sy] - exi);
      if (ORD >= 4) vz = vz + p.cy[1] * (ex[i + 2 * sy] - ex[i - sy]);
      if (ORD >= 6) vz = vz + p.cy[2] * (ex[i + 3 * sy] - ex[i - 2 * sy]);
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
template <int ND>
__global__ void __launch_bounds__(256) k_update_b_ext(GridP g, FieldP F, FdtdP p) {
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
    double vx = bx[i], vy = by[i], vz = bz[i];
    if (ND >= 2) vx = vx - p.cy[0] * ext_diff<true, HZ>(ez, i, sy, sx, sz, q[1], q[5], q[6], q[10], q[13]);
    if (ND >= 3) vx = vx + p.cz[0] * ext_diff<true, true>(ey, i, sz, sx, sy, q[2], q[7], q[8], q[11], q[14]);
    if (ND >= 3) vy = vy - p.cz[0] * ext_diff<true, true>(ex, i, sz, sx, sy, q[2], q[7], q[8], q[11], q[14]);
    vy = vy + p.cx[0] * ext_diff<HY, HZ>(ez, i, sx, sy, sz, q[0], q[3], q[4], q[9], q[12]);
    vz = vz - p.cx[0] * ext_diff<HY, HZ>(ey, i, sx, sy, sz, q[0], q[3], q[4], q[9], q[12]);
    if (ND >= 2) vz = vz + p.cy[0] * ext_diff<true, HZ>(ex, i, sy, sx, sz, q[1], q[5], q[6], q[10], q[13]);
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
  long long cnt = c->g.n[0] + 1;
  if (c->ndims >= 2) cnt *= c->g.n[1] + 1;
  if (c->ndims >= 3) cnt *= c->g.n[2] + 1;
  p.count = cnt;
  return p;
}

static inline int grid_for(long long count, int block) {
  long long b = (count + block - 1) / block;
  const long long cap = 148LL * 8 * 4;   // a few waves of 148 SMs x 8 resident CTAs
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

template <bool E>
static int32_t launch(eb200_ctx* c) {
  FdtdP p = make_params(c, E);
  const int block = 256, grid = grid_for(p.count, block);
#define L(ND, ORD)                                                                  \
  if (c->ndims == ND && c->field_order == ORD) {                                    \
    if (E) k_update_e<ND, ORD><<<grid, block, 0, c->stream>>>(c->g, c->f, p);       \
    else k_update_b<ND, ORD><<<grid, block, 0, c->stream>>>(c->g, c->f, p);         \
  }
  if (!E && c->maxwell_solver != 0 && c->field_order == 2) {
    // fields.f90:637,655: the extended stencils exist in the second-order B update only
    if (c->ndims == 1) k_update_b_ext<1><<<grid, block, 0, c->stream>>>(c->g, c->f, p);
    if (c->ndims == 2) k_update_b_ext<2><<<grid, block, 0, c->stream>>>(c->g, c->f, p);
    if (c->ndims == 3) k_update_b_ext<3><<<grid, block, 0, c->stream>>>(c->g, c->f, p);
  } else {
  L(1, 2) L(1, 4) L(1, 6) L(2, 2) L(2, 4) L(2, 6) L(3, 2) L(3, 4) L(3, 6)
  }
#undef L
  c->launches++;
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

int32_t update_e_field(eb200_ctx* c) { ScopedPhase ph(c, "fields"); return launch<true>(c); }
int32_t update_b_field(eb200_ctx* c) { ScopedPhase ph(c, "fields"); return launch<false>(c); }

}  // namespace eb
