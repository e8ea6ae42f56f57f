// reduce.cu -- device reductions used by the host's energy line and by the parity
// harness: calc_total_energy_sum (io/calc_df.F90:1190-1288) and the Gauss-law
// residual div E - rho/eps0 with rho deposited by the normalised weights of
// calc_charge_density (io/calc_df.F90:439-519, include/*/gxfac.inc).
#include "push_common.cuh"

namespace eb {

__device__ __forceinline__ double block_sum(double v) {
  __shared__ double sh[32];
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) sh[wid] = v;
  __syncthreads();
  v = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.0;
  if (wid == 0) for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  __syncthreads();
  return v;
}

__global__ void k_kinetic(long long n, const double* px, const double* py, const double* pz, const double* w,
                          double part_mc, double* out) {
  double acc = 0.0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const double fac = part_mc * w[i] * C0;
    const double ux = px[i] / part_mc, uy = py[i] / part_mc, uz = pz[i] / part_mc;
    const double u2 = ux * ux + uy * uy + uz * uz;
    const double gamma_rel = sqrt(u2 + 1.0);
    acc += (u2 / (gamma_rel + 1.0)) * fac;
  }
  acc = block_sum(acc);
  if (threadIdx.x == 0) atomicAdd(out, acc);
}

__global__ void k_field_energy(GridP g, FieldP F, double* out) {
  const long long cnt = (long long)g.n[0] * g.n[1] * g.n[2];
  double acc = 0.0;
  const double c2 = C0 * C0;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < cnt; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % g.n[0]) + 1;
    const long long r = t / g.n[0];
    const int j = (int)(r % g.n[1]) + 1, k = (int)(r / g.n[1]) + 1;
    const long long q = g.idx(i, j, k);
    const double ex = F.f[EX][q], ey = F.f[EY][q], ez = F.f[EZ][q];
    const double bx = F.f[BX][q], by = F.f[BY][q], bz = F.f[BZ][q];
    acc += ex * ex + ey * ey + ez * ez + c2 * (bx * bx + by * by + bz * bz);
  }
  acc = block_sum(acc);
  if (threadIdx.x == 0) atomicAdd(out, acc);
}

int32_t energy(eb200_ctx* c, double* fe, double* per_species) {
  const int ns = (int)c->species.size();
  if (ns + 1 > 64) return fail(c, EB200_E_UNSUPPORTED, "more than 63 species");
  EB_CUDA(c, cudaMemsetAsync(c->d_red, 0, 64 * sizeof(double), c->stream));
  k_field_energy<<<148 * 4, 256, 0, c->stream>>>(c->g, c->f, c->d_red);
  c->launches++;
  for (int s = 0; s < ns; ++s) {
    const SpeciesD& sp = c->species[s];
    if (!sp.defined || sp.zero_current || sp.n == 0) continue;
    k_kinetic<<<148 * 4, 256, 0, c->stream>>>(sp.n, sp.a[3], sp.a[4], sp.a[5], sp.a[6], C0 * sp.mass, c->d_red + 1 + s);
    c->launches++;
  }
  EB_CUDA(c, cudaMemcpyAsync(c->h_red, c->d_red, 64 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  double vol = 1.0;
  for (int d = 0; d < c->ndims; ++d) vol *= c->g.d[d];
  if (fe) *fe = 0.5 * EPS0 * c->h_red[0] * vol;
  if (per_species) for (int s = 0; s < ns; ++s) per_species[s] = c->h_red[1 + s];
  return 0;
}

// normalised weights (include/*/gxfac.inc)
template <int SHAPE>
__device__ __forceinline__ void weights_fac(double f, double* w) {
  if (SHAPE == SHAPE_TOPHAT) { w[0] = 0.5 + f; w[1] = 0.5 - f; }
  else if (SHAPE == SHAPE_TRIANGLE) {
    const double c2 = f * f;
    w[0] = 0.5 * (0.25 + c2 + f); w[1] = 0.75 - c2; w[2] = 0.5 * (0.25 + c2 - f);
  } else {
    const double fac1 = 1.0 / 24.0, fac2 = 1.0 / 6.0, fac3 = 115.0 / 48.0;
    const double c2 = f * f;
    double a = 0.5 + f; a = a * a; w[0] = fac1 * (a * a);
    w[1] = fac2 * (1.1875 + 2.75 * f + c2 * (1.5 - f - c2));
    w[2] = 0.25 * (fac3 + c2 * (c2 - 2.5));
    w[3] = fac2 * (1.1875 - 2.75 * f + c2 * (1.5 + f - c2));
    double b = 0.5 - f; b = b * b; w[4] = fac1 * (b * b);
  }
}

// rho at the half-drifted positions x + u c dt / (2 gamma): the positions the
// Esirkepov current links (see oracle/orc_particles.cpp charge_residual)
template <int ND, int SHAPE>
__global__ void k_rho(GridP g, PushP p, const double* X, const double* Y, const double* Z, const double* PX,
                      const double* PY, const double* PZ, const double* W, double qvol, double* rho) {
  constexpr int NW = Sh<SHAPE>::NW, SFM = Sh<SHAPE>::SFMIN;
  const long long ip = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (ip >= p.n) return;
  const double* POS[3] = {X, Y, Z};
  const double u0[3] = {PX[ip] * p.ipart_mc, PY[ip] * p.ipart_mc, PZ[ip] * p.ipart_mc};
  const double root = p.dtco2 / sqrt(u0[0] * u0[0] + u0[1] * u0[1] + u0[2] * u0[2] + 1.0);
  int cell[3] = {1, 1, 1};
  double w[3][NW];
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    double cr = ((POS[d][ip] - p.gml[d]) + u0[d] * root) / g.d[d];
    if (SHAPE == SHAPE_TOPHAT) cr -= 0.5;
    double frac;
    cell[d] = cell_round(cr, frac);
    weights_fac<SHAPE>(frac, w[d]);
  }
  const double wdata = qvol * W[ip];
#pragma unroll
  for (int k = 0; k < (ND >= 3 ? NW : 1); ++k)
#pragma unroll
    for (int j = 0; j < (ND >= 2 ? NW : 1); ++j)
#pragma unroll
      for (int i = 0; i < NW; ++i) {
        const long long q = g.idx(cell[0] + SFM + i, ND >= 2 ? cell[1] + SFM + j : 1, ND >= 3 ? cell[2] + SFM + k : 1);
        atomicAdd(rho + q, w[0][i] * (ND >= 2 ? w[1][j] : 1.0) * (ND >= 3 ? w[2][k] : 1.0) * wdata);
      }
}

__device__ __forceinline__ void atomic_max_pos(double* addr, double v) {
  // v >= 0: the IEEE bit patterns of non-negative doubles order like integers
  atomicMax((unsigned long long*)addr, (unsigned long long)__double_as_longlong(v));
}

__global__ void k_residual(GridP g, FieldP F, const double* rho, double* ref, int mode /*0 plain,1 capture,2 minus ref*/,
                           double* out) {
  const long long cnt = (long long)g.n[0] * g.n[1] * g.n[2];
  double mr = 0.0, mrho = 0.0;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < cnt; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % g.n[0]) + 1;
    const long long r = t / g.n[0];
    const int j = (int)(r % g.n[1]) + 1, k = (int)(r / g.n[1]) + 1;
    const long long q = g.idx(i, j, k);
    double div = (F.f[EX][q] - F.f[EX][q - g.st[0]]) / g.d[0];
    if (g.ndims >= 2) div += (F.f[EY][q] - F.f[EY][q - g.st[1]]) / g.d[1];
    if (g.ndims >= 3) div += (F.f[EZ][q] - F.f[EZ][q - g.st[2]]) / g.d[2];
    const double rr = rho[q] / EPS0;
    double res = div - rr;
    if (mode == 1) ref[q] = res;
    if (mode == 2) res -= ref[q];
    mr = fmax(mr, fabs(res));
    mrho = fmax(mrho, fabs(rr));
  }
  atomic_max_pos(out, mr);
  atomic_max_pos(out + 1, mrho);
}

int32_t charge_residual(eb200_ctx* c, double* max_res, double* max_rho, bool capture) {
  const GridP& g = c->g;
  if (!c->rho) EB_CUDA(c, cudaMalloc(&c->rho, (size_t)g.total * sizeof(double)));
  EB_CUDA(c, cudaMemsetAsync(c->rho, 0, (size_t)g.total * sizeof(double), c->stream));
  double idvol = 1.0;
  for (int d = 0; d < c->ndims; ++d) idvol /= g.d[d];
  for (SpeciesD& sp : c->species) {
    if (!sp.defined || sp.n == 0) continue;
    PushP p;
    p.n = sp.n;
    for (int d = 0; d < 3; ++d) { p.idx[d] = 1.0 / g.d[d]; p.gml[d] = g.gml[d]; }
    p.dtco2 = C0 * (c->dt / 2.0);
    p.ipart_mc = 1.0 / (C0 * sp.mass);
    const unsigned grid = (unsigned)((sp.n + 127) / 128);
#define L(ND, SH)                                                                                  \
    if (c->ndims == ND && c->shape == SH)                                                          \
      k_rho<ND, SH><<<grid, 128, 0, c->stream>>>(g, p, sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], \
                                                  sp.a[5], sp.a[6], sp.charge * idvol, c->rho);
    L(1, 0) L(1, 1) L(1, 2) L(2, 0) L(2, 1) L(2, 2) L(3, 0) L(3, 1) L(3, 2)
#undef L
    c->launches++;
  }
  // fold ghost contributions exactly like the current (reflect / periodic / neighbours)
  EB_TRY(summation_bcs(c, c->rho, -1));
  int mode = 0;
  if (capture) {
    if (!c->res_ref) EB_CUDA(c, cudaMalloc(&c->res_ref, (size_t)g.total * sizeof(double)));
    mode = 1;
  } else if (c->res_ref) mode = 2;
  EB_CUDA(c, cudaMemsetAsync(c->d_red, 0, 2 * sizeof(double), c->stream));
  k_residual<<<148 * 4, 256, 0, c->stream>>>(g, c->f, c->rho, c->res_ref, mode, c->d_red);
  c->launches++;
  EB_CUDA(c, cudaMemcpyAsync(c->h_red, c->d_red, 2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (max_res) *max_res = c->h_red[0];
  if (max_rho) *max_rho = c->h_red[1];
  return 0;
}

// ---------------------------------------------------------------------------------------------
// charge_residual without any allocation (eb200_charge_residual_static; no counterpart in the
// reference, added for callers that may not call cudaMalloc:
// CUDA graph capture, tight serving loops).  The rho scratch of charge_residual -- one double per
// grid point incl. ghosts, cudaMalloc'ed lazily there -- is this file-scope device array instead:
// STATIC_RHO_MAX doubles = 16 MiB, held by every GPU that loads the library until exit.
// All calls, from every context on a device, share the array; it keeps the last call's contents, so
// it is cleared at the start of every call, and calls that overlap on the GPU (two contexts = two
// streams) race.  In host code the array's NAME is a host-side stand-in, not a device address: the
// address is fetched once per device, with that device current, by cudaGetSymbolAddress.
// ---------------------------------------------------------------------------------------------
constexpr long long STATIC_RHO_MAX = 1LL << 21;              // grid points incl. ghosts (2 097 152: e.g. 118^3 cells, triangle)
static __device__ __align__(256) double g_static_rho[STATIC_RHO_MAX];
constexpr int STATIC_MAX_DEVICES = 64;
static double* g_static_rho_addr[STATIC_MAX_DEVICES] = {nullptr};

long long charge_residual_static_max() { return STATIC_RHO_MAX; }

static int32_t static_rho_address(eb200_ctx* c, double** out) {
  int dev = -1;
  EB_CUDA(c, cudaGetDevice(&dev));                            // (CHECK_CTX has made the context's device current)
  if (dev < 0 || dev >= STATIC_MAX_DEVICES) return fail(c, EB200_E_UNSUPPORTED, "charge_residual_static: device ordinal beyond 63");
  if (!g_static_rho_addr[dev]) {
    double* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_static_rho);
    if (e != cudaSuccess || !ptr) {
      cudaGetLastError();
      return fail(c, EB200_E_CUDA, std::string("cudaGetSymbolAddress(g_static_rho): ") + cudaGetErrorString(e));
    }
    g_static_rho_addr[dev] = ptr;
  }
  *out = g_static_rho_addr[dev];
  return 0;
}

// the body of charge_residual (mode 0 / 2: plain, or minus the captured reference) on the static scratch;
// calls no cudaMalloc / cudaFree: d_red, h_red exist since create, the single-rank ghost sums are kernels only
int32_t charge_residual_static(eb200_ctx* c, double* max_res, double* max_rho) {
  const GridP& g = c->g;
  if (g.total > STATIC_RHO_MAX)
    return fail(c, EB200_E_ARG, "charge_residual_static: grid of " + std::to_string(g.total) + " points (ghosts included) exceeds the static scratch of " +
                                std::to_string(STATIC_RHO_MAX) + " points; use charge_residual");
  if (c->nproc > 1)
    return fail(c, EB200_E_UNSUPPORTED, "charge_residual_static: single-rank contexts only (the multi-rank ghost sum sizes its halo buffers on demand)");
  EB_TRY(settle_particles(c));          // (a pending deferred permute is finished by copy kernels inside the particle arrays: no allocation)
  double* rho = nullptr;
  EB_TRY(static_rho_address(c, &rho));
  EB_CUDA(c, cudaMemsetAsync(rho, 0, (size_t)g.total * sizeof(double), c->stream));     // the array keeps the last call's contents
  double idvol = 1.0;
  for (int d = 0; d < c->ndims; ++d) idvol /= g.d[d];
  for (SpeciesD& sp : c->species) {
    if (!sp.defined || sp.n == 0) continue;
    PushP p;
    p.n = sp.n;
    for (int d = 0; d < 3; ++d) { p.idx[d] = 1.0 / g.d[d]; p.gml[d] = g.gml[d]; }
    p.dtco2 = C0 * (c->dt / 2.0);
    p.ipart_mc = 1.0 / (C0 * sp.mass);
    const unsigned grid = (unsigned)((sp.n + 127) / 128);
#define L(ND, SH)                                                                                  \
    if (c->ndims == ND && c->shape == SH)                                                          \
      k_rho<ND, SH><<<grid, 128, 0, c->stream>>>(g, p, sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], \
                                                  sp.a[5], sp.a[6], sp.charge * idvol, rho);
    L(1, 0) L(1, 1) L(1, 2) L(2, 0) L(2, 1) L(2, 2) L(3, 0) L(3, 1) L(3, 2)
#undef L
    c->launches++;
  }
  EB_TRY(summation_bcs(c, rho, -1));
  EB_CUDA(c, cudaMemsetAsync(c->d_red, 0, 2 * sizeof(double), c->stream));
  k_residual<<<148 * 4, 256, 0, c->stream>>>(g, c->f, rho, c->res_ref, c->res_ref ? 2 : 0, c->d_red);
  c->launches++;
  EB_CUDA(c, cudaMemcpyAsync(c->h_red, c->d_red, 2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (max_res) *max_res = c->h_red[0];
  if (max_rho) *max_rho = c->h_red[1];
  return 0;
}

}  // namespace eb
