// window.cu -- the device part of the moving window (housekeeping/window.F90): one call shifts
// the window by one cell, as one iteration of shift_window's loop does (window.F90:62-94):
//   * the x grid of this rank moves by dx (setup_grid_x with the new x_grid_min; the host passes
//     the new scalars so that their rounding is the host's),
//   * remove_particles (window.F90:362-384): on an x_min boundary rank the particles with
//     pos_x < x_min are dropped,
//   * shift_fields (window.F90:98-194): every grid array moves one cell towards -x, its halos are
//     refreshed and, on an x_max boundary rank, the incoming cells are set from the boundary
//     planes e*_x_max / b*_x_max of setup_field_boundaries (setup.F90:425-470).
// insert_particles (new plasma in the incoming cell column) needs the deck's density /
// temperature expressions and the host RNG: the host builds those particles and hands them over
// with eb200_append_particles before this call, in the reference's order.
#include "ctx.cuh"

namespace eb {

// field(i) = field(i + 1), i = 1-ng .. nx+ng-1, for every line along x (x is the fastest axis):
// a block owns a line and walks it in chunks; each chunk is read (one element ahead) before any
// of it is written, and the element a chunk reads from the next chunk is written only later.
__global__ void __launch_bounds__(256) k_shift_x(double* __restrict__ a, int len, long long nlines) {
  for (long long line = blockIdx.x; line < nlines; line += gridDim.x) {
    double* l = a + line * len;
    for (int b = 0; b < len - 1; b += blockDim.x) {
      const int i = b + threadIdx.x;
      const double v = i < len - 1 ? l[i + 1] : 0.0;
      __syncthreads();
      if (i < len - 1) l[i] = v;
      __syncthreads();
    }
  }
}

// "Fix incoming field cell" (window.F90:147-165), one thread per (j, k) of the full extents
struct WinP { double* f[6]; const double* pl[6]; int nx; long long sx, sy; long long nlines; };
__global__ void k_window_incoming(WinP w) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= w.nlines) return;
  const long long b = t * w.sy;               // element (1-ng, j, k); x index i lives at b + (i + ng - 1)
  const long long o = b - w.sx;               // so that o + i is x index i - ... (sx = ng - 1 offset, see caller)
  double *ex = w.f[0], *ey = w.f[1], *ez = w.f[2], *bx = w.f[3], *by = w.f[4], *bz = w.f[5];
  const int nx = w.nx;
#define AT(i) (o + (i))
  ex[AT(nx)] = w.pl[0][t];
  ex[AT(nx + 1)] = w.pl[0][t];
  ey[AT(nx + 1)] = w.pl[1][t];
  ez[AT(nx + 1)] = w.pl[2][t];
  ex[AT(nx - 1)] = 0.5 * (ex[AT(nx - 2)] + ex[AT(nx)]);
  ey[AT(nx)] = 0.5 * (ey[AT(nx - 1)] + ey[AT(nx + 1)]);
  ez[AT(nx)] = 0.5 * (ez[AT(nx - 1)] + ez[AT(nx + 1)]);
  bx[AT(nx + 1)] = w.pl[3][t];
  by[AT(nx)] = w.pl[4][t];
  bz[AT(nx)] = w.pl[5][t];
  bx[AT(nx)] = 0.5 * (bx[AT(nx - 1)] + bx[AT(nx + 1)]);
  by[AT(nx - 1)] = 0.5 * (by[AT(nx - 2)] + by[AT(nx)]);
  bz[AT(nx - 1)] = 0.5 * (bz[AT(nx - 2)] + bz[AT(nx)]);
#undef AT
}

// cpml_psi(nx+1) = cpml_psi(nx) on the x_max boundary rank (window.F90:151-169), for npsi arrays
__global__ void k_window_psi(double* psi, long long total, int npsi, int nx, long long sx, long long sy, long long nlines) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nlines * npsi) return;
  double* a = psi + (t / nlines) * total + (t % nlines) * sy - sx;
  a[nx + 1] = a[nx];
}

__global__ void k_window_remove(const double* __restrict__ x, uint32_t* key, long long n, double x_min) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && x[i] < x_min) key[i] = KEY_INVALID;
}

int32_t set_window_planes(eb200_ctx* c, const double* const planes[6]) {
  const GridP& g = c->g;
  const size_t cnt = (size_t)(g.total / (g.n[0] + 2 * g.ngd[0]));     // lines along x = the y-z extents with ghosts
  if (!c->win_planes) EB_CUDA(c, cudaMalloc(&c->win_planes, 6 * cnt * sizeof(double)));
  EB_CUDA(c, cudaMemsetAsync(c->win_planes, 0, 6 * cnt * sizeof(double), c->stream));
  for (int k = 0; k < 6; ++k)
    if (planes && planes[k])
      EB_CUDA(c, cudaMemcpyAsync(c->win_planes + k * cnt, planes[k], cnt * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int32_t shift_window(eb200_ctx* c, double grid_min_local, double min_local, double max_local, double gmin, double gmax) {
  ScopedPhase ph(c, "window");
  GridP& g = c->g;
  // setup_grid_x
  g.gml[0] = grid_min_local; g.min_local[0] = min_local; g.max_local[0] = max_local;
  g.gmin[0] = gmin; g.gmax[0] = gmax;
  // remove_particles
  if (c->min_bnd[0]) {
    for (SpeciesD& sp : c->species) {
      if (!sp.defined || sp.n == 0) continue;
      k_window_remove<<<(unsigned)((sp.n + 255) / 256), 256, 0, c->stream>>>(sp.a[0], sp.key, sp.n, gmin);
      c->launches++;
    }
  }
  for (SpeciesD& sp : c->species) sp.sorted = false;       // every cell index moved by one
  // shift_fields
  const int len = g.n[0] + 2 * g.ngd[0];
  const long long nlines = g.total / len;
  const int grid = (int)std::min<long long>(nlines, 148LL * 16);
  for (int f = 0; f < NFIELD; ++f) {
    k_shift_x<<<grid, 256, 0, c->stream>>>(c->f.f[f], len, nlines);
    c->launches++;
    EB_TRY(field_bc(c, f, 1));
  }
  // the psi arrays of the CPML layers move with the fields (window.F90:114-129)
  const int npsi = c->cpml_psi ? 4 * c->ndims : 0;
  for (int k = 0; k < npsi; ++k) {
    double* a = c->cpml_psi + (size_t)k * g.total;
    k_shift_x<<<grid, 256, 0, c->stream>>>(a, len, nlines);
    c->launches++;
    EB_TRY(field_bc_array(c, a));
  }
  if (c->max_bnd[0]) {
    if (!c->win_planes) EB_TRY(set_window_planes(c, nullptr));
    WinP w;
    for (int k = 0; k < 6; ++k) { w.f[k] = c->f.f[k]; w.pl[k] = c->win_planes + (size_t)k * nlines; }
    w.nx = g.n[0]; w.sy = len; w.sx = 1 - g.ngd[0]; w.nlines = nlines;   // o + i = b + (i + ng - 1)
    k_window_incoming<<<(unsigned)((nlines + 127) / 128), 128, 0, c->stream>>>(w);
    c->launches++;
    if (npsi) {
      k_window_psi<<<(unsigned)((nlines * npsi + 127) / 128), 128, 0, c->stream>>>(c->cpml_psi, g.total, npsi, g.n[0], w.sx, w.sy, nlines);
      c->launches++;
    }
  }
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

}  // namespace eb
