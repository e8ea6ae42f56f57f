// particles.cu -- particle container (cell-sorted structure of arrays, fp64),
// push_particles orchestration, per-step counting sort, particle boundary
// conditions and migration.  Replaces the linked-list container and wire format of
// housekeeping/partlist.F90 (:353-553,:808-862), push_particles' driver part
// (particles.F90:147-212,609-612) and particle_bcs (boundary.F90:1316-1972).
#include "push_common.cuh"
#include <cub/device/device_scan.cuh>
#include <algorithm>

namespace eb {

int32_t launch_push_baseline(eb200_ctx* c, const SpeciesD& sp, const PushP& p);   // push_baseline.cu
int32_t launch_push_split(eb200_ctx* c, SpeciesD& sp, const PushP& p, bool* handled, const void* keysink);  // push_tiled.cu
size_t keysink_size();
void keysink_fill(void* dst, const BcP3& bc, uint32_t* key, uint32_t* count, uint32_t nbucket);
void keysink_set_migration(void* dst, int axis, double mn, double mx, uint32_t* list_lo, uint32_t* list_hi,
                           unsigned long long* cnt, uint32_t cap);

// ---------------------------------------------------------------------------------
// storage
// ---------------------------------------------------------------------------------
void set_view(SpeciesD& sp, long long off) {
  sp.off = off;
  for (int k = 0; k < 7; ++k) sp.a[k] = sp.base[k] ? sp.base[k] + off : nullptr;
  sp.key = sp.key_base ? sp.key_base + off : nullptr;
}

template <class T>
static int32_t regrow(eb200_ctx* c, T*& base, const T* view, long long old_n, long long new_cap, long long new_off) {
  T* np = nullptr;
  EB_CUDA(c, cudaMalloc(&np, (size_t)new_cap * sizeof(T)));
  if (view && old_n > 0) EB_CUDA(c, cudaMemcpyAsync(np + new_off, view, (size_t)old_n * sizeof(T), cudaMemcpyDeviceToDevice, c->stream));
  if (base) { EB_CUDA(c, cudaStreamSynchronize(c->stream)); cudaFree(base); }
  base = np;
  return 0;
}

// room for `need` live slots (+ the head room of the deferred permute: win slots between the low and
// the high state, gap slots for migration arrivals at either end).  The live slots [0, sp.n) of
// the current view are kept and end up in the low state.
int32_t species_reserve(eb200_ctx* c, SpeciesD& sp, long long need, bool exact) {
  if (sp.base[6] && need + sp.win + 2 * sp.gap <= sp.cap) return 0;
  EB_TRY(settle_species(c, sp));
  const bool tiled = c->push_kernel != 1;
  // "auto" head room of the deferred permute: an eighth of the species, and at least 4.5 layers of tiles along the
  // slowest axis -- a particle that crosses a tile boundary along that axis moves by one layer of tiles in the sorted
  // order and the displacement limit is a quarter of the window (on a thin z slab of a big x-y face, the strong-
  // scaling cut, a layer is a large part of the species: 512 x 512 x 64 cells at 16 ppc = 16.7 M of 268 M particles)
  long long win_auto = std::max<long long>(need / 8, 1 << 16);
  if (c->grid_set && c->g.ncell_ext > 0) {
    const GridP& g = c->g;
    const double cells = (double)g.n[0] * g.n[1] * g.n[2];
    const double layer = (double)g.ncell_ext / (double)std::max(1, g.ntile[c->ndims - 1]) * ((double)need / std::max(cells, 1.0));
    win_auto = std::max<long long>(win_auto, (long long)(4.5 * layer));
    win_auto = std::min<long long>(win_auto, std::max<long long>(need, 1 << 16));
  }
  long long win = !tiled ? 0 : (c->defer_window >= 0 ? c->defer_window : win_auto);
  const long long gap = 0;      // (slots kept free below the low state: none needed, arrivals are appended)
  const long long body = exact ? std::max<long long>(need, 1024) : std::max<long long>(need + need / 32 + 1024, 4096);
  const long long ncap = body + win + 2 * gap;
  // attributes live in a[0..ndims-1] (pos), a[3..5] (p), a[6] (w)
  for (int k = 0; k < 7; ++k) {
    if (k < 3 && k >= c->ndims) continue;
    EB_TRY(regrow(c, sp.base[k], (const double*)sp.a[k], sp.n, ncap, gap));
  }
  EB_TRY(regrow(c, sp.key_base, (const uint32_t*)sp.key, sp.n, ncap, gap));
  {
    uint32_t* nd = nullptr;
    EB_CUDA(c, cudaMalloc(&nd, (size_t)ncap * sizeof(uint32_t)));
    if (sp.dest) { EB_CUDA(c, cudaStreamSynchronize(c->stream)); cudaFree(sp.dest); }
    sp.dest = nd;
  }
  sp.cap = ncap; sp.win = win; sp.gap = gap;
  set_view(sp, gap);
  return 0;
}

// AoS (partlist wire format) <-> SoA
__global__ void k_unpack_aos(const double* __restrict__ packed, int nd, long long n, long long off,
                             double* x, double* y, double* z, double* px, double* py, double* pz, double* w,
                             uint32_t* key) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* q = packed + i * (nd + 4);
  x[off + i] = q[0];
  if (nd >= 2) y[off + i] = q[1];
  if (nd >= 3) z[off + i] = q[2];
  px[off + i] = q[nd]; py[off + i] = q[nd + 1]; pz[off + i] = q[nd + 2];
  w[off + i] = q[nd + 3];
  key[off + i] = 0;
}
__global__ void k_pack_aos(double* __restrict__ packed, int nd, long long n, long long off,
                           const double* x, const double* y, const double* z, const double* px,
                           const double* py, const double* pz, const double* w) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double* q = packed + i * (nd + 4);
  q[0] = x[off + i];
  if (nd >= 2) q[1] = y[off + i];
  if (nd >= 3) q[2] = z[off + i];
  q[nd] = px[off + i]; q[nd + 1] = py[off + i]; q[nd + 2] = pz[off + i];
  q[nd + 3] = w[off + i];
}

static int32_t ensure_stage(eb200_ctx* c, long long doubles) {
  if (c->mig_cap >= doubles) return 0;
  for (int s = 0; s < 2; ++s) {
    if (c->mig_send[s]) cudaFree(c->mig_send[s]);
    if (c->mig_recv[s]) cudaFree(c->mig_recv[s]);
    c->mig_send[s] = c->mig_recv[s] = nullptr;
  }
  c->mig_cap = 0;
  for (int s = 0; s < 2; ++s) {
    EB_CUDA(c, cudaMalloc(&c->mig_send[s], (size_t)doubles * sizeof(double)));
    EB_CUDA(c, cudaMalloc(&c->mig_recv[s], (size_t)doubles * sizeof(double)));
  }
  c->mig_cap = doubles;
  return 0;
}

int32_t add_particles(eb200_ctx* c, int isp, long long n, const double* packed, int nvar, bool replace) {
  SpeciesD& sp = c->species[isp];
  if (replace) { sp.n = 0; sp.was_ordered = false; }
  if (n == 0) { sp.sorted = false; return 0; }
  EB_TRY(species_reserve(c, sp, sp.n + n));
  const size_t rec = (size_t)nvar * sizeof(double);
  EB_TRY(ensure_stage(c, std::max<long long>((long long)(c->pin_bytes / sizeof(double)), 1 << 16)));
  // chunks of whole particles go pinned ring -> device staging -> SoA arrays (stream-ordered)
  double* stage = c->mig_recv[0];
  const long long n0 = sp.n;
  EB_TRY(staged_h2d(c, packed, (size_t)n * rec, rec,
      [stage](size_t, size_t) { return (void*)stage; },
      [c, &sp, stage, rec, n0](size_t o, size_t mbytes) -> int32_t {
        const long long m = (long long)(mbytes / rec), first = (long long)(o / rec);
        k_unpack_aos<<<(unsigned)((m + 255) / 256), 256, 0, c->stream>>>(stage, c->ndims, m, n0 + first,
            sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], sp.a[5], sp.a[6], sp.key);
        c->launches++;
        EB_CUDA(c, cudaGetLastError());
        return 0;
      }));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  sp.n += n;
  sp.sorted = false;
  return 0;
}

int32_t get_particles(eb200_ctx* c, int isp, long long first, long long cap, double* packed, int nvar, long long* n_out) {
  SpeciesD& sp = c->species[isp];
  const long long n = std::max<long long>(0, std::min(sp.n - first, cap));
  const size_t rec = (size_t)nvar * sizeof(double);
  EB_TRY(ensure_stage(c, std::max<long long>((long long)(c->pin_bytes / sizeof(double)), 1 << 16)));
  double* stage = c->mig_send[0];
  EB_TRY(staged_d2h(c, packed, (size_t)n * rec, rec,
      [c, &sp, stage, rec, first](size_t o, size_t mbytes) -> const void* {
        const long long m = (long long)(mbytes / rec), f0 = first + (long long)(o / rec);
        k_pack_aos<<<(unsigned)((m + 255) / 256), 256, 0, c->stream>>>(stage, c->ndims, m, f0,
            sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], sp.a[5], sp.a[6]);
        c->launches++;
        return cudaGetLastError() == cudaSuccess ? (const void*)stage : nullptr;
      }));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (n_out) *n_out = n;
  return 0;
}

// ---------------------------------------------------------------------------------
// per-step counting sort by the cell the NEXT push will gather from
// ---------------------------------------------------------------------------------
// key = extended-grid linear index of (cell_x1, cell_y1, cell_z1) evaluated exactly
// as the push does (half drift with the stored momentum, particles.F90:242-291).
// particle boundary condition of one axis (particle_bcs, boundary.F90:1316-1972)

// Fused pass over the particles of a species: (i) particle_bcs of the axes that are not
// decomposed (no migration there: periodic wrap, reflection, open-boundary removal),
// (ii) sort key of the cell the next push gathers from, (iii) warp-aggregated histogram of
// the keys.  Removed particles get KEY_INVALID and are counted in the last bucket.
template <int ND, int SHAPE>
__global__ void k_keys(GridP g, PushP p, BcP3 bc, double* __restrict__ X, double* __restrict__ Y,
                       double* __restrict__ Z, double* __restrict__ PX, double* __restrict__ PY,
                       double* __restrict__ PZ, uint32_t* key, uint32_t* count, uint32_t nbucket, long long i0 = 0) {
  const long long ip = i0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool act = ip < p.n;
  uint32_t k = nbucket - 1;
  if (act && key[ip] != KEY_INVALID) {
    double* POS[3] = {X, Y, Z};
    double* MOM[3] = {PX, PY, PZ};
    double pos[3] = {0, 0, 0}, mom[3];
    mom[0] = PX[ip]; mom[1] = PY[ip]; mom[2] = PZ[ip];
#pragma unroll
    for (int d = 0; d < ND; ++d) pos[d] = POS[d][ip];
    const int ch = bc_and_key<ND, SHAPE>(g, p, bc, pos, mom, k);
    if (ch < 0) { k = nbucket - 1; key[ip] = KEY_INVALID; }
    else {
      key[ip] = k;
#pragma unroll
      for (int d = 0; d < ND; ++d)
        if (ch & (1 << d)) { POS[d][ip] = pos[d]; MOM[d][ip] = mom[d]; }
    }
  }
  // histogram: sorted input => few distinct keys per warp
  const unsigned am = __ballot_sync(0xffffffffu, act);
  if (!act) return;
  const unsigned peers = __match_any_sync(am, k);
  if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(count + k, (uint32_t)__popc(peers));
}

// warp-aggregated histogram / rank: sorted input => few distinct keys per warp
__global__ void k_hist(const uint32_t* __restrict__ key, long long n, uint32_t* count, uint32_t nbucket) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool act = i < n;
  const unsigned am = __ballot_sync(0xffffffffu, act);
  if (!act) return;
  uint32_t k = key[i];
  if (k == KEY_INVALID) k = nbucket - 1;
  const unsigned peers = __match_any_sync(am, k);
  const int leader = __ffs(peers) - 1;
  if ((threadIdx.x & 31) == leader) atomicAdd(count + k, (uint32_t)__popc(peers));
}
__global__ void k_dest(const uint32_t* __restrict__ key, long long n, const uint32_t* __restrict__ start,
                       uint32_t* cursor, uint32_t* dest, uint32_t nbucket) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool act = i < n;
  const unsigned am = __ballot_sync(0xffffffffu, act);
  if (!act) return;
  uint32_t k = key[i];
  if (k == KEY_INVALID) k = nbucket - 1;
  const unsigned peers = __match_any_sync(am, k);
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(peers) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(cursor + k, (uint32_t)__popc(peers));
  base = __shfl_sync(peers, base, leader);
  dest[i] = start[k] + base + (uint32_t)__popc(peers & ((1u << lane) - 1u));
}
__global__ void k_permute(const double* __restrict__ src, double* __restrict__ dst,
                          const uint32_t* __restrict__ dest, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[dest[i]] = src[i];
}
// k_dest of the deferred permute: stores the INVERSE map (src[sorted index] = slot).  A slot displaced by
// more than `limit` (a particle wrapped around a periodic boundary of the slowest tile axis, a migration
// arrival, ...) is copied to the species' side buffer instead -- src = n + index there; far_count counts
// them (beyond far_cap they stay in place and the host falls back to a permute).
struct Attr7 { const double* a[7]; };
// A warp takes SRC_U consecutive groups of 32 slots and keeps their memory round trips in flight together (the key
// loads, then one cursor atomic per (group, key) and the start[] loads side by side, then the stores): with one
// group per warp the three dependent round trips left the kernel latency bound at 1.8 TB/s of DRAM traffic
// (ncu, 512^3: 20.4 GB in 11.2 ms, profiles/r02_dram_512_summary.txt).
#ifndef EB_SRC_U
#define EB_SRC_U 4
#endif
constexpr int SRC_U = EB_SRC_U;
__global__ void __launch_bounds__(256) k_src(const uint32_t* __restrict__ key, long long n, const uint32_t* __restrict__ start,
                      uint32_t* cursor, uint32_t* src, uint32_t nbucket, unsigned long long* far_count,
                      Attr7 at, double* __restrict__ far, long long far_cap, long long limit) {
  const int lane = threadIdx.x & 31;
  const long long wbase = (((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32LL * SRC_U);
  if (wbase >= n) return;
  uint32_t k[SRC_U], st[SRC_U], base[SRC_U];
  unsigned peers[SRC_U];
  bool live[SRC_U];
#pragma unroll
  for (int u = 0; u < SRC_U; ++u) {
    const long long i = wbase + u * 32 + lane;
    k[u] = i < n ? key[i] : KEY_INVALID;
  }
#pragma unroll
  for (int u = 0; u < SRC_U; ++u) {
    const long long i = wbase + u * 32 + lane;
    const bool act = i < n;
    const unsigned am = __ballot_sync(0xffffffffu, act);
    live[u] = act && k[u] != KEY_INVALID;
    if (!live[u]) k[u] = nbucket - 1;
    peers[u] = 0; base[u] = 0; st[u] = 0;
    if (act) {
      peers[u] = __match_any_sync(am, k[u]);
      if (lane == __ffs(peers[u]) - 1) base[u] = atomicAdd(cursor + k[u], (uint32_t)__popc(peers[u]));
      st[u] = start[k[u]];
    }
  }
#pragma unroll
  for (int u = 0; u < SRC_U; ++u) {
    const long long i = wbase + u * 32 + lane;
    if (i >= n) continue;
    const uint32_t b = __shfl_sync(peers[u], base[u], __ffs(peers[u]) - 1);
    const uint32_t d = st[u] + b + (uint32_t)__popc(peers[u] & ((1u << lane) - 1u));
    if (live[u]) {
      uint32_t s = (uint32_t)i;
      const long long df = (long long)d - i;
      if (df > limit || -df > limit) {
        // (an unordered load displaces everything: once the buffer is full nobody queues at the counter any more;
        //  a count >= far_cap tells the host to permute at once)
        const unsigned long long slot = *far_count < (unsigned long long)far_cap ? atomicAdd(far_count, 1ULL) : (unsigned long long)far_cap;
        if ((long long)slot < far_cap) {
#pragma unroll
          for (int a = 0; a < 7; ++a)
            if (at.a[a]) far[(size_t)a * far_cap + slot] = at.a[a][i];
          s = (uint32_t)(n + (long long)slot);
        }
      }
      src[d] = s;
    }
  }
}
// dst[i] = from[src[i]] for i in [i0, i1); src >= n_main points into the side buffer
__global__ void k_gather_copy(const double* __restrict__ from, const double* __restrict__ far, long long n_main,
                              double* __restrict__ dst, const uint32_t* __restrict__ src, long long i0, long long i1) {
  const long long i = i0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= i1) return;
  const long long s = (long long)src[i];
  dst[i] = s >= n_main ? far[s - n_main] : from[s];
}
// Deterministic mode: the atomic cursors of k_src leave the order INSIDE a cell to the warp scheduler.
// Thread per cell: re-order the cell's src entries by the particles' content (x, then y, z, px, py, pz,
// w, then slot) so that the sorted order -- and with it every serial sum of the deposit -- is a function
// of the particle set alone.
struct Canon {
  Attr7 at; const double* far; long long far_cap, n_main;
  __device__ __forceinline__ double get(int a, uint32_t s) const {
    return (long long)s >= n_main ? far[(size_t)a * far_cap + ((long long)s - n_main)] : at.a[a][s];
  }
  __device__ bool less(uint32_t sa, uint32_t sb, double xa, double xb) const {
    if (xa != xb) return xa < xb;
#pragma unroll 1
    for (int a = 1; a < 7; ++a) {
      if (!at.a[a]) continue;
      const double va = get(a, sa), vb = get(a, sb);
      if (va != vb) return va < vb;
    }
    return sa < sb;
  }
};
__global__ void k_canon(Canon cn, const uint32_t* __restrict__ start, long long nbucket_live, uint32_t* src) {
  const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= nbucket_live) return;
  const uint32_t b = start[cell], e = start[cell + 1];
  // insertion sort; the keys of the entries already placed sit in a small window of local memory
  constexpr int WN = 48;
  double kx[WN];
  for (uint32_t i = b; i < e; ++i) {
    const uint32_t si = src[i];
    const double xi = cn.get(0, si);
    uint32_t j = i;
    while (j > b) {
      const uint32_t sj = src[j - 1];
      const double xj = (j - 1 - b) < (uint32_t)WN ? kx[j - 1 - b] : cn.get(0, sj);
      if (!cn.less(si, sj, xi, xj)) break;
      src[j] = sj;
      if ((j - b) < (uint32_t)WN) kx[j - b] = xj;
      --j;
    }
    src[j] = si;
    if ((j - b) < (uint32_t)WN) kx[j - b] = xi;
  }
}
__global__ void k_tile_start(const uint32_t* __restrict__ cell_start, uint32_t* ts, long long ntile1, int cpt) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < ntile1) ts[t] = cell_start[t * cpt];
}
__global__ void k_fill_u32(uint32_t* a, long long n, uint32_t v) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = v;
}

static BcP make_bc(const eb200_ctx* c, const SpeciesD& sp, int ax) {
  const GridP& g = c->g;
  BcP b;
  b.axis = ax; b.min_bnd = c->min_bnd[ax]; b.max_bnd = c->max_bnd[ax];
  b.bc_min = sp.bc[2 * ax]; b.bc_max = sp.bc[2 * ax + 1];
  b.multi = c->nprocdir[ax] > 1;
  b.min_local = g.min_local[ax]; b.max_local = g.max_local[ax];
  b.gmin = g.gmin[ax]; b.gmax = g.gmax[ax];
  const double boundary_shift = g.d[ax] * (double)((1 + c->png + c->cpml_thickness) / 2);   // boundary.F90:1336-1338
  b.min_outer = b.gmin - boundary_shift; b.max_outer = b.gmax + boundary_shift;
  b.shift = (b.gmax - b.gmin) + 2.0 * g.d[ax] * (double)c->cpml_thickness;                  // x_shift
  const int fmin = c->bc_field[2 * ax], fmax = c->bc_field[2 * ax + 1];
  b.cpml_min = fmin == EB200_BC_CPML_LASER || fmin == EB200_BC_CPML_OUTFLOW;
  b.cpml_max = fmax == EB200_BC_CPML_LASER || fmax == EB200_BC_CPML_OUTFLOW;
  return b;
}

static PushP make_push_params(eb200_ctx* c, const SpeciesD& sp, long long n) {
  PushP p;
  const int nd = c->ndims;
  const double fac = shape_fac(c->shape, nd);
  p.n = n;
  for (int d = 0; d < 3; ++d) { p.idx[d] = 1.0 / c->g.d[d]; p.gml[d] = c->g.gml[d]; }
  const double idt = 1.0 / c->dt;
  p.dto2 = c->dt / 2.0;
  p.dtco2 = C0 * p.dto2;
  const double dtfac = 0.5 * c->dt * fac;
  p.q = sp.charge;
  p.part_mc = C0 * sp.mass;
  p.ipart_mc = 1.0 / p.part_mc;
  p.cmratio = p.q * dtfac * p.ipart_mc;
  p.ccmratio = C0 * p.cmratio;
  if (nd == 3) { p.fc[0] = idt * p.idx[1] * p.idx[2] * fac; p.fc[1] = idt * p.idx[0] * p.idx[2] * fac; p.fc[2] = idt * p.idx[0] * p.idx[1] * fac; }
  else if (nd == 2) { p.fc[0] = idt * p.idx[1] * fac; p.fc[1] = idt * p.idx[0] * fac; p.fc[2] = p.idx[0] * p.idx[1] * fac; }
  else { p.fc[0] = idt * fac; p.fc[1] = p.idx[0] * fac; p.fc[2] = 0.0; }
  p.zero_current = sp.zero_current;
  return p;
}

static int32_t ensure_cells(eb200_ctx* c, SpeciesD& sp) {
  const long long need = c->g.ncell_ext + 2;
  if (c->cell_cap < need) {
    if (c->cell_count) cudaFree(c->cell_count);
    c->cell_count = nullptr; c->cell_cap = 0;
    EB_CUDA(c, cudaMalloc(&c->cell_count, (size_t)need * sizeof(uint32_t)));
    c->cell_cap = need;
  }
  if (sp.cell_cap < need) {
    if (sp.cell_start) cudaFree(sp.cell_start);
    sp.cell_start = nullptr; sp.cell_cap = 0;
    EB_CUDA(c, cudaMalloc(&sp.cell_start, (size_t)need * sizeof(uint32_t)));
    sp.cell_cap = need;
  }
  return 0;
}

static int32_t ensure_spare(eb200_ctx* c, long long cap) {
  if (c->attr_scratch_cap == cap) return 0;   // spare and attribute must be interchangeable: same capacity
  if (c->attr_scratch) { EB_CUDA(c, cudaStreamSynchronize(c->stream)); cudaFree(c->attr_scratch); }
  c->attr_scratch = nullptr; c->attr_scratch_cap = 0;
  EB_CUDA(c, cudaMalloc(&c->attr_scratch, (size_t)cap * sizeof(double)));
  c->attr_scratch_cap = cap;
  return 0;
}

// the deferred permute moves the live slots from the current state to the other one
struct DeferPlan { bool fwd; long long off_new, room; };
static DeferPlan defer_plan(const SpeciesD& sp) {
  DeferPlan d;
  d.fwd = sp.off >= sp.gap + sp.win;            // high -> low: sorted index i lands BELOW its old slots
  d.off_new = d.fwd ? sp.gap : sp.gap + sp.win;
  const long long W = d.fwd ? sp.off - d.off_new : d.off_new - sp.off;
  // A kernel that writes the sorted indices [a, b) overwrites the old slots [a - W, b - W) (fwd; [a + W, b + W)
  // otherwise).  With chunks taken in increasing (decreasing) order those slots have all been read by earlier
  // kernels iff b - a <= W - max displacement.
  d.room = W - (d.fwd ? sp.disp_fwd : sp.disp_bwd);
  return d;
}

// finish a deferred permute without a push: gather-copy every attribute into the other state
int32_t settle_species(eb200_ctx* c, SpeciesD& sp) {
  if (!sp.deferred) return 0;
  ScopedPhase ph(c, "sort");
  const DeferPlan d = defer_plan(sp);
  const long long chunk = d.room;
  for (int k = 0; k < 7; ++k) {
    if (k < 3 && k >= c->ndims) continue;
    double* dst = sp.base[k] + d.off_new;
    const long long nch = (sp.n + chunk - 1) / chunk;
    for (long long ic = 0; ic < nch; ++ic) {
      const long long q = d.fwd ? ic : nch - 1 - ic;
      const long long i0 = q * chunk, i1 = std::min(sp.n, i0 + chunk);
      k_gather_copy<<<(unsigned)((i1 - i0 + 255) / 256), 256, 0, c->stream>>>(sp.a[k], sp.far + (size_t)k * sp.far_cap, sp.n_old,
                                                                               dst, sp.dest, i0, i1);
      c->launches++;
    }
  }
  set_view(sp, d.off_new);
  if (sp.n > 0) {
    k_fill_u32<<<(unsigned)((sp.n + 255) / 256), 256, 0, c->stream>>>(sp.key, sp.n, 0u);
    c->launches++;
  }
  sp.deferred = false;
  EB_CUDA(c, cudaGetLastError());
  return 0;
}
int32_t settle_particles(eb200_ctx* c) {
  for (SpeciesD& sp : c->species) if (sp.defined) EB_TRY(settle_species(c, sp));
  return 0;
}

// sort slots [0, n_slots) of a species; dead slots (key == KEY_INVALID) go last
// and are dropped.  keys_ready: the push kernel has already written keys and histogram.
// may_defer: leave the arrays in their present order and let the next push gather through
// sp.dest (= src) when the displacements fit the window (deferred permute).
int32_t sort_species(eb200_ctx* c, SpeciesD& sp, long long n_slots, bool keys_ready, bool may_defer) {
  if (sp.deferred) {
    if (keys_ready) return fail(c, EB200_E_STATE, "sort_species: keys handed over for a deferred species");
    EB_TRY(settle_species(c, sp));
  }
  ScopedPhase ph(c, "sort");
  if (n_slots == 0) { sp.n = 0; sp.sorted = true; sp.sort_dt = c->dt; return 0; }
  if (n_slots >= 0xFFFFFFF0LL) return fail(c, EB200_E_UNSUPPORTED, "more than 2^32 particles per species per GPU");
  EB_TRY(ensure_cells(c, sp));
  const uint32_t nbucket = (uint32_t)(c->g.ncell_ext + 1);   // last = invalid
  const int B = 256;
  const unsigned gp = (unsigned)((n_slots + B - 1) / B);
  PushP p = make_push_params(c, sp, n_slots);
  BcP3 bc;
  for (int d = 0; d < 3; ++d) {
    bc.on[d] = d < c->ndims && c->nprocdir[d] == 1;      // decomposed axes were handled by species_bcs (migration)
    bc.b[d] = make_bc(c, sp, d < c->ndims ? d : 0);
  }
  if (!keys_ready) {      // else the push kernel has already written keys and histogram
  EB_CUDA(c, cudaMemsetAsync(c->cell_count, 0, (size_t)(nbucket + 1) * sizeof(uint32_t), c->stream));
#define LK(ND, SH)                                                                               \
  if (c->ndims == ND && c->shape == SH)                                                          \
    k_keys<ND, SH><<<gp, B, 0, c->stream>>>(c->g, p, bc, sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], sp.a[5], sp.key, c->cell_count, nbucket);
  LK(1, 0) LK(1, 1) LK(1, 2) LK(2, 0) LK(2, 1) LK(2, 2) LK(3, 0) LK(3, 1) LK(3, 2)
#undef LK
  }
  size_t tmp = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp, c->cell_count, sp.cell_start, (int)(nbucket + 1), c->stream);
  if (tmp > c->cub_tmp_bytes) {
    if (c->cub_tmp) cudaFree(c->cub_tmp);
    c->cub_tmp = nullptr; c->cub_tmp_bytes = 0;
    EB_CUDA(c, cudaMalloc(&c->cub_tmp, tmp));
    c->cub_tmp_bytes = tmp;
  }
  cub::DeviceScan::ExclusiveSum(c->cub_tmp, tmp, c->cell_count, sp.cell_start, (int)(nbucket + 1), c->stream);
  EB_CUDA(c, cudaMemsetAsync(c->cell_count, 0, (size_t)(nbucket + 1) * sizeof(uint32_t), c->stream));
  const bool inverse = (may_defer && sp.win > 0) || c->deterministic;
  if (inverse) {
    if (!sp.far) {
      sp.far_cap = std::max<long long>(sp.cap / 256, 1 << 14);
      EB_CUDA(c, cudaMalloc(&sp.far, (size_t)7 * sp.far_cap * sizeof(double)));
    }
    if (n_slots + sp.far_cap >= 0xFFFFFFF0LL) return fail(c, EB200_E_UNSUPPORTED, "more than 2^32 particles per species per GPU");
    Attr7 at;
    for (int k = 0; k < 7; ++k) at.a[k] = (k < 3 && k >= c->ndims) ? nullptr : sp.a[k];
    const long long W = sp.off >= sp.gap + sp.win ? sp.off - sp.gap : sp.gap + sp.win - sp.off;
    // the window is shared between the displacements (<= limit) and the particles of one launch
    const bool deferrable = may_defer && sp.win > 0;      // else: deterministic mode, permute below
    const long long limit = deferrable ? W / 4 : (1LL << 62);
    sp.disp_fwd = sp.disp_bwd = deferrable ? limit : W;
    EB_CUDA(c, cudaMemsetAsync(c->d_counters + 12, 0, sizeof(unsigned long long), c->stream));
    k_src<<<(unsigned)((n_slots + (long long)B * SRC_U - 1) / ((long long)B * SRC_U)), B, 0, c->stream>>>(sp.key, n_slots, sp.cell_start, c->cell_count, sp.dest, nbucket, c->d_counters + 12,
                                   at, sp.far, sp.far_cap, limit);
    EB_CUDA(c, cudaMemcpyAsync(c->h_counters + 12, c->d_counters + 12, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    if (c->deterministic) {
      Canon cn; cn.at = at; cn.far = sp.far; cn.far_cap = sp.far_cap; cn.n_main = n_slots;
      const long long ncell = (long long)nbucket - 1;      // the dead bucket keeps its order: it is dropped
      k_canon<<<(unsigned)((ncell + 127) / 128), 128, 0, c->stream>>>(cn, sp.cell_start, ncell, sp.dest);
      c->launches++;
    }
  } else {
    k_dest<<<gp, B, 0, c->stream>>>(sp.key, n_slots, sp.cell_start, c->cell_count, sp.dest, nbucket);
  }
  c->launches += keys_ready ? 3 : 4;
  if (!inverse) {
    // scatter each attribute through one spare array; the spare rotates.  The result is the low state.
    EB_TRY(ensure_spare(c, sp.cap));
    for (int k = 0; k < 7; ++k) {
      if (k < 3 && k >= c->ndims) continue;
      k_permute<<<gp, B, 0, c->stream>>>(sp.a[k], c->attr_scratch + sp.gap, sp.dest, n_slots);
      std::swap(sp.base[k], c->attr_scratch);
      c->launches++;
    }
  }
  // first particle of every tile, for the host's slab partition of the split push/deposit kernels
  const long long nt1 = c->g.ncell_ext / c->g.cpt + 1;
  {
    if (sp.tile_cap < nt1) {
      if (sp.d_tile_start) cudaFree(sp.d_tile_start);
      if (sp.h_tile_start) cudaFreeHost(sp.h_tile_start);
      sp.d_tile_start = sp.h_tile_start = nullptr; sp.tile_cap = 0;
      EB_CUDA(c, cudaMalloc(&sp.d_tile_start, (size_t)nt1 * sizeof(uint32_t)));
      EB_CUDA(c, cudaMallocHost(&sp.h_tile_start, (size_t)nt1 * sizeof(uint32_t)));
      sp.tile_cap = nt1;
    }
    k_tile_start<<<(unsigned)((nt1 + 255) / 256), 256, 0, c->stream>>>(sp.cell_start, sp.d_tile_start, nt1, c->g.cpt);
    c->launches++;
    EB_CUDA(c, cudaMemcpyAsync(sp.h_tile_start, sp.d_tile_start, (size_t)nt1 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
  }
  // live count = start of the invalid bucket
  EB_CUDA(c, cudaMemcpyAsync(c->h_counters, sp.cell_start + (nbucket - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  sp.n = (long long)(*(uint32_t*)c->h_counters);
  bool view_changed = false;
  if (inverse) {
    const DeferPlan d = defer_plan(sp);
    long long biggest = 1;
    for (long long t = 0; t + 1 < nt1; ++t) biggest = std::max<long long>(biggest, (long long)(sp.h_tile_start[t + 1] - sp.h_tile_start[t]));
    const long long n_far = (long long)c->h_counters[12];
    if (may_defer && sp.win > 0 && d.room >= biggest && n_far < sp.far_cap && d.off_new + sp.n + sp.gap <= sp.cap) {
      sp.deferred = true; sp.n_old = n_slots;
      c->n_deferred++;
    } else if (sp.n > 0) {
      // displacements beyond the window (first sort of an unordered load, a re-sliced domain, ...):
      // permute now, gathering through the spare into the low state
      EB_TRY(ensure_spare(c, sp.cap));
      const unsigned gl = (unsigned)((sp.n + B - 1) / B);
      for (int k = 0; k < 7; ++k) {
        if (k < 3 && k >= c->ndims) continue;
        k_gather_copy<<<gl, B, 0, c->stream>>>(sp.a[k], sp.far + (size_t)k * sp.far_cap, n_slots, c->attr_scratch + sp.gap,
                                               sp.dest, 0, sp.n);
        std::swap(sp.base[k], c->attr_scratch);
        c->launches++;
      }
      view_changed = true;
      c->n_permuted++;
    }
  } else {
    view_changed = true;
    c->n_permuted++;
  }
  if (view_changed) {
    const bool moved = sp.off != sp.gap;
    set_view(sp, sp.gap);
    // keys are not permuted: stale KEY_INVALID marks must not survive in the slots now live
    if ((sp.n != n_slots || moved) && sp.n > 0) {
      k_fill_u32<<<(unsigned)((sp.n + B - 1) / B), B, 0, c->stream>>>(sp.key, sp.n, 0u);
      c->launches++;
    }
    // a spare of tens of GB is given back once the deferred permute carries the steady state
    if (sp.win > 0 && c->attr_scratch && (size_t)c->attr_scratch_cap * sizeof(double) > ((size_t)1 << 30)) {
      EB_CUDA(c, cudaStreamSynchronize(c->stream));
      cudaFree(c->attr_scratch);
      c->attr_scratch = nullptr; c->attr_scratch_cap = 0;
    }
  }
  sp.sorted = true;
  sp.was_ordered = true;
  sp.sort_dt = c->dt;
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------
// particle boundary conditions (one axis at a time: x, then y, then z)
// ---------------------------------------------------------------------------------
// flags: dest[i] = 0 stay, 1 send to low neighbour, 2 send to high neighbour
__global__ void k_particle_bc(BcP b, long long n, double* __restrict__ pos, double* __restrict__ mom,
                              uint32_t* key, uint32_t* flag, unsigned long long* counters) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (key[i] == KEY_INVALID) { if (b.multi) flag[i] = 0; return; }
  const double part_pos = pos[i];
  int xbd = 0;
  bool out = false;
  if (b.cpml_min) {
    if (b.min_bnd) { if (part_pos < b.min_outer) out = true; }
    else if (part_pos < b.min_local) xbd = -1;
  } else if (part_pos < b.min_local) {
    xbd = -1;
    if (b.min_bnd) {
      xbd = 0;
      if (b.bc_min == EB200_BC_REFLECT) { pos[i] = 2.0 * b.gmin - part_pos; mom[i] = -mom[i]; }
      else if (b.bc_min == EB200_BC_PERIODIC) { xbd = -1; pos[i] = part_pos + b.shift; }
      if (part_pos < b.min_outer && b.bc_min != EB200_BC_PERIODIC) out = true;
    }
  }
  if (b.cpml_max) {
    if (b.max_bnd) { if (part_pos >= b.max_outer) out = true; }
    else if (part_pos >= b.max_local) xbd = 1;
  } else if (part_pos >= b.max_local) {
    xbd = 1;
    if (b.max_bnd) {
      xbd = 0;
      if (b.bc_max == EB200_BC_REFLECT) { pos[i] = 2.0 * b.gmax - part_pos; mom[i] = -mom[i]; }
      else if (b.bc_max == EB200_BC_PERIODIC) { xbd = 1; pos[i] = part_pos - b.shift; }
      if (part_pos >= b.max_outer && b.bc_max != EB200_BC_PERIODIC) out = true;
    }
  }
  if (out) { key[i] = KEY_INVALID; atomicAdd(counters + 2, 1ULL); xbd = 0; }
  if (b.multi) {
    flag[i] = xbd < 0 ? 1u : (xbd > 0 ? 2u : 0u);
    if (xbd != 0) atomicAdd(counters + (xbd < 0 ? 0 : 1), 1ULL);
  }
}
// ---- thermal particle boundaries (boundary.F90:1405-1461 for x_min, the five other faces alike) ----------------
// A particle beyond the OUTER edge of a thermal face is re-emitted: temperature from ext_temp at its transverse
// position (triangle weighting, always), momenta drawn in the order x, y, z (normal component from the flux
// distribution heading inwards: SQRT(N1^2 + N2^2) of two normal deviates; the others normal), position mirrored at
// the outer edge.  Runs before the other particle boundaries of the step.  The reference handles a particle's axes
// in turn, so a particle that also wraps / reflects on an EARLIER axis in the same step is seen here with that
// position taken into account (a local copy; the wrap itself is stored by the regular pass).  The deviates come
// from a counter-based generator keyed by (seed, call, slot): the reference's is one sequential host stream whose
// draws depend on the list order, so the distributions agree, not the numbers.
struct ThermalP {
  int nd, ng, on[6], ext[3], wrap_min[3], wrap_max[3], refl_min[3], refl_max[3];
  double edge[6], gml[3], d[3], min_local[3], max_local[3], gmin[3], gmax[3], shift[3], sigma2;   // sigma2 = kB * mass
  const double* temp[6];
  unsigned long long seed, call;
};
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {      // splitmix64 finaliser
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
__device__ __forceinline__ double u01(unsigned long long key, unsigned long long ctr) {   // uniform in (0, 1)
  const unsigned long long r = mix64(key + ctr * 0x9E3779B97F4A7C15ULL);
  return ((double)(r >> 11) + 0.5) * (1.0 / 9007199254740992.0);
}
__global__ void k_thermal_bc(ThermalP t, long long n, double* x, double* y, double* z, double* px, double* py, double* pz,
                             const uint32_t* __restrict__ key, unsigned long long* counters) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || key[i] == KEY_INVALID) return;
  double q[3] = {x[i], t.nd >= 2 ? y[i] : 0.0, t.nd >= 3 ? z[i] : 0.0};
  double mom[3];
  bool hit = false;
  const unsigned long long kbase = mix64(t.seed ^ mix64(t.call * 0xD1B54A32D192ED03ULL + (unsigned long long)i));
  unsigned long long ctr = 0;
  for (int d = 0; d < t.nd; ++d) {
    const double part_pos = q[d];
    for (int side = 0; side < 2; ++side) {
      const int b = 2 * d + side;
      if (!t.on[b]) {
        // the position an earlier axis leaves behind for the later ones (periodic wrap / reflection; not stored here)
        if (side == 0 && part_pos < t.min_local[d]) { if (t.wrap_min[d]) q[d] = part_pos + t.shift[d]; else if (t.refl_min[d]) q[d] = 2.0 * t.gmin[d] - part_pos; }
        if (side == 1 && part_pos >= t.max_local[d]) { if (t.wrap_max[d]) q[d] = part_pos - t.shift[d]; else if (t.refl_max[d]) q[d] = 2.0 * t.gmax[d] - part_pos; }
        continue;
      }
      if (side == 0 ? !(part_pos < t.edge[b]) : !(part_pos >= t.edge[b])) continue;
      int ta[2], nt = 0;
      for (int a = 0; a < t.nd; ++a) if (a != d) ta[nt++] = a;
      int cell[2] = {1, 1};
      double g[2][3] = {{0.0, 1.0, 0.0}, {0.0, 1.0, 0.0}};
      for (int k = 0; k < nt; ++k) {
        const int a = ta[k];
        const double cell_r = __ddiv_rn(__dsub_rn(q[a], t.gml[a]), t.d[a]);
        const int c0 = (int)floor(__dadd_rn(cell_r, 0.5));
        const double cf = __dsub_rn((double)c0, cell_r);
        cell[k] = c0 + 1;
        const double cf2 = __dmul_rn(cf, cf);
        g[k][0] = 0.5 * (0.25 + cf2 + cf);
        g[k][1] = 0.75 - cf2;
        g[k][2] = 0.5 * (0.25 + cf2 - cf);
      }
      const long long e1 = nt >= 1 ? t.ext[ta[0]] : 1, e2 = nt >= 2 ? t.ext[ta[1]] : 1;
      const double* et = t.temp[b];
      double temp[3];
      for (int c = 0; c < 3; ++c) {
        temp[c] = 0.0;
        if (nt == 0) { temp[c] = et[c]; continue; }
        for (int i2 = (nt >= 2 ? -1 : 0); i2 <= (nt >= 2 ? 1 : 0); ++i2)
          for (int i1 = -1; i1 <= 1; ++i1) {
            long long j1 = cell[0] + i1 - (1 - t.ng), j2 = nt >= 2 ? cell[1] + i2 - (1 - t.ng) : 0;
            j1 = j1 < 0 ? 0 : (j1 >= e1 ? e1 - 1 : j1);        // (memory safety for a particle far outside the transverse range)
            j2 = j2 < 0 ? 0 : (j2 >= e2 ? e2 - 1 : j2);
            const double w2 = nt >= 2 ? g[1][i2 + 1] : 1.0;
            temp[c] = temp[c] + g[0][i1 + 1] * w2 * et[((long long)c * e2 + j2) * e1 + j1];
          }
      }
      // four normal deviates (two Box-Muller pairs): the normal component takes two, the others one each
      double nrm[4];
      for (int k = 0; k < 2; ++k) {
        const double u1 = u01(kbase, ctr++), u2 = u01(kbase, ctr++);
        const double r = sqrt(-2.0 * log(u1));
        double sn, cs;
        sincospi(2.0 * u2, &sn, &cs);
        nrm[2 * k] = r * cs; nrm[2 * k + 1] = r * sn;
      }
      int used = 0;
      for (int c = 0; c < 3; ++c) {
        const double stdev = sqrt(temp[c] * t.sigma2);
        if (c == d) {
          const double m1 = nrm[used] * stdev, m2 = nrm[used + 1] * stdev;
          used += 2;
          mom[c] = (side == 0 ? 1.0 : -1.0) * sqrt(m1 * m1 + m2 * m2);
        } else {
          mom[c] = nrm[used++] * stdev;
        }
      }
      q[d] = 2.0 * t.edge[b] - part_pos;
      hit = true;
      double* pa = d == 0 ? x : (d == 1 ? y : z);
      pa[i] = q[d];
      atomicAdd(counters + 9, 1ULL);
      break;      // (mirrored at the min edge: it cannot be beyond the max edge too)
    }
  }
  if (hit) { px[i] = mom[0]; py[i] = mom[1]; pz[i] = mom[2]; }
}

// re-emission at this rank's thermal faces, before the other particle boundaries of the step
static int32_t thermal_bcs(eb200_ctx* c, SpeciesD& sp, long long n_slots) {
  ThermalP t{};
  bool any = false;
  for (int d = 0; d < c->ndims; ++d) {
    const BcP b = make_bc(c, sp, d);
    for (int side = 0; side < 2; ++side) {
      const int f = 2 * d + side;
      const bool at_face = side == 0 ? b.min_bnd : b.max_bnd;
      const bool cpml = side == 0 ? b.cpml_min : b.cpml_max;
      t.on[f] = at_face && !cpml && sp.bc[f] == EB200_BC_THERMAL;
      if (t.on[f] && !sp.ext_temp[f]) return fail(c, EB200_E_STATE, "thermal particle boundary without set_thermal_boundary");
      t.edge[f] = side == 0 ? b.min_outer : b.max_outer;
      t.temp[f] = sp.ext_temp[f];
      any = any || t.on[f];
    }
    t.gml[d] = c->g.gml[d]; t.d[d] = c->g.d[d]; t.ext[d] = c->g.n[d] + 2 * c->ng;
    t.min_local[d] = b.min_local; t.max_local[d] = b.max_local; t.gmin[d] = b.gmin; t.gmax[d] = b.gmax; t.shift[d] = b.shift;
    t.wrap_min[d] = b.min_bnd && !b.cpml_min && b.bc_min == EB200_BC_PERIODIC; t.wrap_max[d] = b.max_bnd && !b.cpml_max && b.bc_max == EB200_BC_PERIODIC;
    t.refl_min[d] = b.min_bnd && !b.cpml_min && b.bc_min == EB200_BC_REFLECT; t.refl_max[d] = b.max_bnd && !b.cpml_max && b.bc_max == EB200_BC_REFLECT;
  }
  if (!any || n_slots <= 0) return 0;
  // (the rank enters the key: ranks handed the same seed must not draw the same deviates for the same slot and call)
  t.nd = c->ndims; t.ng = c->ng; t.sigma2 = KB * sp.mass;
  t.seed = sp.thermal_seed ^ ((unsigned long long)c->rank * 0xA0761D6478BD642FULL); t.call = sp.thermal_calls++;
  k_thermal_bc<<<(unsigned)((n_slots + 255) / 256), 256, 0, c->stream>>>(t, n_slots, sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], sp.a[5],
                                                                       sp.key, c->d_counters);
  c->launches++;
  EB_CUDA(c, cudaGetLastError());
  return 0;
}
static bool has_thermal(const eb200_ctx* c, const SpeciesD& sp) {
  for (int f = 0; f < 2 * c->ndims; ++f) if (sp.bc[f] == EB200_BC_THERMAL) return true;
  return false;
}

// Migration messages are fixed-capacity mailboxes: [packed count, wanted count, cap x nvar doubles].  The
// capacity of a directed link is agreed without talking: both ends grow it by the same rule from the count the
// sender WANTED to send (it travels in the header), read a step late at the sort's synchronise.  So particle_bcs
// needs no host round trip of its own, and a rank that sees more leavers than a mailbox holds keeps the rest one
// more step (they sit in its ghost cells: gather and deposit still reach valid cells) instead of failing.
constexpr int MIG_HDR = 2;

// pack flagged particles into the per-side mailboxes and free their slots; a leaver beyond the capacity stays
// (its periodic shift, applied by k_particle_bc, is taken back)
__global__ void k_pack_leavers(long long n, int nd, int ax, double* x, double* y, double* z,
                               const double* px, const double* py, const double* pz, const double* w,
                               uint32_t* key, const uint32_t* flag, unsigned long long* cursors,
                               double* send_lo, double* send_hi, long long cap_lo, long long cap_hi, double undo_lo,
                               double undo_hi, uint32_t* count = nullptr, uint32_t invalid_bucket = 0,
                               long long n_counted = 0, int hdr = MIG_HDR) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t f = flag[i];
  if (f == 0) return;
  const unsigned long long slot = atomicAdd(cursors + (f == 1 ? 3 : 4), 1ULL);
  if (slot >= (unsigned long long)(f == 1 ? cap_lo : cap_hi)) {
    double* pa = ax == 0 ? x : (ax == 1 ? y : z);
    pa[i] -= f == 1 ? undo_lo : undo_hi;
    return;
  }
  double* q = (f == 1 ? send_lo : send_hi) + hdr + slot * (nd + 4);
  q[0] = x[i];
  if (nd >= 2) q[1] = y[i];
  if (nd >= 3) q[2] = z[i];
  q[nd] = px[i]; q[nd + 1] = py[i]; q[nd + 2] = pz[i]; q[nd + 3] = w[i];
  // the push kernel has already counted this slot under its key (fused keys): move the count to the dead bucket
  if (count && i < n_counted) { atomicSub(count + key[i], 1u); atomicAdd(count + invalid_bucket, 1u); }
  key[i] = KEY_INVALID;
}

// the same from the slot lists the push kernel made (fused migration): blockIdx.y = side; the counts are read on
// the device (counters[0..1]); a leaver through a global periodic boundary is shifted to the other end
// (boundary.F90:1390-1461) as it is packed
__global__ void k_pack_list(const uint32_t* __restrict__ list_lo, const uint32_t* __restrict__ list_hi,
                            const unsigned long long* __restrict__ counters, long long list_cap, long long cap_lo,
                            long long cap_hi, int nd, int ax, double shift_lo, double shift_hi, const double* x,
                            const double* y, const double* z, const double* px, const double* py, const double* pz,
                            const double* w, uint32_t* key, double* send_lo, double* send_hi, uint32_t* count,
                            uint32_t invalid_bucket) {
  const int side = blockIdx.y;
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long wanted = (long long)counters[side];
  long long m = wanted < list_cap ? wanted : list_cap;
  m = m < (side == 0 ? cap_lo : cap_hi) ? m : (side == 0 ? cap_lo : cap_hi);
  double* buf = side == 0 ? send_lo : send_hi;
  if (j == 0) { buf[0] = (double)m; buf[1] = (double)wanted; }
  if (j >= m) return;
  const uint32_t i = (side == 0 ? list_lo : list_hi)[j];
  double* q = buf + MIG_HDR + j * (nd + 4);
  q[0] = x[i];
  if (nd >= 2) q[1] = y[i];
  if (nd >= 3) q[2] = z[i];
  q[ax] += side == 0 ? shift_lo : shift_hi;
  q[nd] = px[i]; q[nd + 1] = py[i]; q[nd + 2] = pz[i]; q[nd + 3] = w[i];
  atomicSub(count + key[i], 1u); atomicAdd(count + invalid_bucket, 1u);
  key[i] = KEY_INVALID;
}
// headers of the mailboxes filled by k_pack_leavers (counters[0..1] = wanted, cursors clamp to the capacity)
__global__ void k_mig_header(const unsigned long long* counters, double* send_lo, double* send_hi, long long cap_lo, long long cap_hi) {
  const long long w0 = (long long)counters[0], w1 = (long long)counters[1];
  send_lo[0] = (double)(w0 < cap_lo ? w0 : cap_lo); send_lo[1] = (double)w0;
  send_hi[0] = (double)(w1 < cap_hi ? w1 : cap_hi); send_hi[1] = (double)w1;
}
// arrivals of one mailbox -> slots [off, off + cap): the first `packed` are particles, the rest dead slots
__global__ void k_unpack_mailbox(const double* __restrict__ box, int nd, long long cap, long long off,
                                 double* x, double* y, double* z, double* px, double* py, double* pz, double* w,
                                 uint32_t* key) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cap) return;
  const long long n = (long long)box[0];
  if (i >= n) { key[off + i] = KEY_INVALID; return; }
  const double* q = box + MIG_HDR + i * (nd + 4);
  x[off + i] = q[0];
  if (nd >= 2) y[off + i] = q[1];
  if (nd >= 3) z[off + i] = q[2];
  px[off + i] = q[nd]; py[off + i] = q[nd + 1]; pz[off + i] = q[nd + 2];
  w[off + i] = q[nd + 3];
  key[off + i] = 0;
}
// what both ends of the four links learn this step: wanted counts out (lo, hi) and in (from hi, from lo)
__global__ void k_mig_info(const unsigned long long* counters, const double* recv_hi, const double* recv_lo, int has_hi,
                           int has_lo, unsigned long long* info) {
  info[0] = counters[0]; info[1] = counters[1];
  info[2] = has_hi ? (unsigned long long)recv_hi[1] : 0ULL;
  info[3] = has_lo ? (unsigned long long)recv_lo[1] : 0ULL;
}

static int32_t read_counters(eb200_ctx* c, int n) {
  EB_CUDA(c, cudaMemcpyAsync(c->h_counters, c->d_counters, n * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

static long long grow_cap(long long cap, unsigned long long wanted) {
  long long need = 2 * (long long)wanted;
  while (cap < need) cap *= 2;
  return cap;
}

// applies particle_bcs to slots [0, n_slots) of one species; returns the new slot
// count (arrivals are appended) and whether the live count may have changed
// n_counted > 0: the push kernel wrote keys + histogram for the slots [0, n_counted) (fused keys)
// list_axis >= 0: the push kernel has listed the leavers of that axis (c->mig_list, counts in d_counters[0..1])
// No host synchronisation: counts stay on the device, the mailboxes have agreed capacities (see above); the
// appended slots [n_slots, n_slots + capacities) hold the arrivals and dead slots, which the sort drops.
static int32_t species_bcs(eb200_ctx* c, SpeciesD& sp, long long& n_slots, bool& changed, long long n_counted = 0,
                           int list_axis = -1) {
  const int nvar = c->ndims + 4;
  for (int ax = 0; ax < c->ndims; ++ax) {
    const BcP b = make_bc(c, sp, ax);
    if (!b.multi) continue;        // axes that are not decomposed are handled inside the sort's key pass
    changed = true;
    const bool lists = ax == list_axis;
    const int lo = c->nbr_lo[ax], hi = c->nbr_hi[ax];
    // capacities of the four links: grown from what was wanted last time (read at the sort's synchronise since)
    if (!sp.h_mig) {
      EB_CUDA(c, cudaMallocHost(&sp.h_mig, 12 * sizeof(unsigned long long)));
      EB_CUDA(c, cudaMalloc(&sp.d_mig, 12 * sizeof(unsigned long long)));
      for (int k = 0; k < 12; ++k) sp.h_mig[k] = 0;
    }
    for (int k = 0; k < 4; ++k) {
      if (sp.mig_cap[ax][k] < c->mig_min_particles) sp.mig_cap[ax][k] = c->mig_min_particles;
      if (sp.mig_pending[ax]) sp.mig_cap[ax][k] = grow_cap(sp.mig_cap[ax][k], sp.h_mig[4 * ax + k]);
    }
    if (sp.mig_pending[ax] && lists) {
      const long long most = (long long)std::max(sp.h_mig[4 * ax], sp.h_mig[4 * ax + 1]);
      if (2 * most > c->mig_list_cap) c->mig_list_want = 4 * most;     // (grown before the next push)
    }
    sp.mig_pending[ax] = false;
    const long long cap_lo = sp.mig_cap[ax][0], cap_hi = sp.mig_cap[ax][1], cap_fhi = sp.mig_cap[ax][2], cap_flo = sp.mig_cap[ax][3];
    const long long cap_max = std::max(std::max(cap_lo, cap_hi), std::max(cap_fhi, cap_flo));
    EB_TRY(ensure_stage(c, std::max<long long>(MIG_HDR + cap_max * nvar, 1 << 16)));
    const double undo_lo = b.min_bnd && b.bc_min == EB200_BC_PERIODIC ? b.shift : 0.0;
    const double undo_hi = b.max_bnd && b.bc_max == EB200_BC_PERIODIC ? -b.shift : 0.0;
    if (lists) {
      const dim3 gl((unsigned)((std::max(cap_lo, cap_hi) + 255) / 256), 2);
      k_pack_list<<<gl, 256, 0, c->stream>>>(c->mig_list[0], c->mig_list[1], c->d_counters, c->mig_list_cap, cap_lo, cap_hi,
          c->ndims, ax, undo_lo, undo_hi, sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], sp.a[5], sp.a[6],
          sp.key, c->mig_send[0], c->mig_send[1], c->cell_count, (uint32_t)c->g.ncell_ext);
      c->launches++;
    } else {
      EB_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 8 * sizeof(unsigned long long), c->stream));
      if (n_slots > 0) {
        k_particle_bc<<<(unsigned)((n_slots + 255) / 256), 256, 0, c->stream>>>(b, n_slots, sp.a[ax], sp.a[3 + ax],
            sp.key, sp.dest, c->d_counters);
        k_pack_leavers<<<(unsigned)((n_slots + 255) / 256), 256, 0, c->stream>>>(n_slots, c->ndims, ax, sp.a[0], sp.a[1], sp.a[2],
            sp.a[3], sp.a[4], sp.a[5], sp.a[6], sp.key, sp.dest, c->d_counters, c->mig_send[0], c->mig_send[1], cap_lo, cap_hi,
            undo_lo, undo_hi, n_counted > 0 ? c->cell_count : nullptr, (uint32_t)c->g.ncell_ext, n_counted);
        c->launches += 2;
      }
      k_mig_header<<<1, 1, 0, c->stream>>>(c->d_counters, c->mig_send[0], c->mig_send[1], cap_lo, cap_hi);
      c->launches++;
    }
    // one exchange for both directions.  A particle leaves through a physical boundary only if that boundary is
    // periodic (xbd is reset to 0 otherwise), so the wrap-around neighbours are always valid.  lo-bound leavers
    // travel to `lo`; what arrives from `hi` are hi's lo-bound leavers.
    EB_TRY(comm_exchange2(c, c->mig_send[0], lo, (size_t)(MIG_HDR + cap_lo * nvar), c->mig_send[1], hi, (size_t)(MIG_HDR + cap_hi * nvar),
                          c->mig_recv[0], hi, (size_t)(MIG_HDR + cap_fhi * nvar), c->mig_recv[1], lo, (size_t)(MIG_HDR + cap_flo * nvar)));
    const long long add = (hi >= 0 ? cap_fhi : 0) + (lo >= 0 ? cap_flo : 0);
    if (add > 0) {
      const long long keep_n = sp.n;
      sp.n = n_slots;                       // regrow copies [0, sp.n)
      EB_TRY(species_reserve(c, sp, n_slots + add));
      sp.n = keep_n;
      // appended; the deferred permute takes slots this far from their sorted place through its side buffer
      long long at = n_slots;
      if (hi >= 0) {
        k_unpack_mailbox<<<(unsigned)((cap_fhi + 255) / 256), 256, 0, c->stream>>>(c->mig_recv[0], c->ndims, cap_fhi, at,
            sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], sp.a[5], sp.a[6], sp.key);
        at += cap_fhi; c->launches++;
      }
      if (lo >= 0) {
        k_unpack_mailbox<<<(unsigned)((cap_flo + 255) / 256), 256, 0, c->stream>>>(c->mig_recv[1], c->ndims, cap_flo, at,
            sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], sp.a[5], sp.a[6], sp.key);
        c->launches++;
      }
      n_slots += add;
    }
    k_mig_info<<<1, 1, 0, c->stream>>>(c->d_counters, c->mig_recv[0], c->mig_recv[1], hi >= 0, lo >= 0, sp.d_mig + 4 * ax);
    c->launches++;
    EB_CUDA(c, cudaMemcpyAsync(sp.h_mig + 4 * ax, sp.d_mig + 4 * ax, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    sp.mig_pending[ax] = true;             // (valid after the next synchronise: the sort's)
  }
  // open boundaries of the axes that are not decomposed are applied by the key pass of the sort
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------
// distribute_particles (balance.F90:2751-2809): after the domain was re-sliced every particle
// goes to the rank that owns its position.  A ring of nproc - 1 shifts; per shift the leavers
// for one destination are packed and exchanged in bounded chunks (the staging buffers stay small
// however much of the plasma changes hands).
// ---------------------------------------------------------------------------------
__global__ void k_flag_owner(OwnerP own, int nd, int npx, int npy, long long n, const double* __restrict__ x,
                             const double* __restrict__ y, const double* __restrict__ z, const uint32_t* __restrict__ key,
                             uint32_t* flag, unsigned long long* counters, int dst, unsigned long long cap) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  flag[i] = 0;
  if (key[i] == KEY_INVALID) return;
  const double pos[3] = {x[i], nd >= 2 ? y[i] : 0.0, nd >= 3 ? z[i] : 0.0};
  int co[3] = {0, 0, 0};
  for (int a = 0; a < nd; ++a)
    for (int p = own.np[a] - 1; p >= 1; --p)
      if (pos[a] >= own.lb[a][p]) { co[a] = p; break; }
  const int owner = (co[2] * npy + co[1]) * npx + co[0];
  if (owner != dst) return;
  if (atomicAdd(counters + 0, 1ULL) < cap) flag[i] = 1;
  else atomicAdd(counters + 1, 1ULL);
}

int32_t redistribute_particles(eb200_ctx* c, const OwnerP& own, std::vector<long long>& slots) {
  const int nvar = c->ndims + 4, P = c->nproc;
  slots.assign(c->species.size(), 0);
  for (size_t is = 0; is < c->species.size(); ++is) {
    SpeciesD& sp = c->species[is];
    if (!sp.defined) continue;
    long long n_slots = sp.n;
    if (n_slots > 0) {                     // every live slot is a candidate
      EB_TRY(species_reserve(c, sp, n_slots));
      k_fill_u32<<<(unsigned)((n_slots + 255) / 256), 256, 0, c->stream>>>(sp.key, n_slots, 0u);
      c->launches++;
    }
    const long long chunk = std::max<long long>(c->mig_min_particles, 1LL << 22);
    EB_TRY(ensure_stage(c, chunk * nvar));
    for (int s = 1; s < P; ++s) {
      const int dst = (c->rank + s) % P, src = (c->rank - s + P) % P;
      for (;;) {
        EB_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 8 * sizeof(unsigned long long), c->stream));
        if (n_slots > 0) {
          k_flag_owner<<<(unsigned)((n_slots + 255) / 256), 256, 0, c->stream>>>(own, c->ndims, c->nprocdir[0], c->nprocdir[1],
              n_slots, sp.a[0], sp.a[1], sp.a[2], sp.key, sp.dest, c->d_counters, dst, (unsigned long long)chunk);
          c->launches++;
        }
        EB_TRY(read_counters(c, 2));
        const long long total = (long long)c->h_counters[0], left = (long long)c->h_counters[1], nsend = total - left;
        if (nsend > 0) {
          k_pack_leavers<<<(unsigned)((n_slots + 255) / 256), 256, 0, c->stream>>>(n_slots, c->ndims, 0, sp.a[0], sp.a[1], sp.a[2],
              sp.a[3], sp.a[4], sp.a[5], sp.a[6], sp.key, sp.dest, c->d_counters, c->mig_send[0], c->mig_send[1], 1LL << 62, 1LL << 62,
              0.0, 0.0, nullptr, 0, 0, 0);
          c->launches++;
        }
        double* cnt = c->d_red;            // [0] sent now, [1] still to come -> [2], [3] from src
        const double hc[2] = {(double)nsend, (double)left};
        EB_CUDA(c, cudaMemcpyAsync(cnt, hc, 2 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        EB_TRY(comm_sendrecv(c, cnt, dst, cnt + 2, src, 2));
        EB_CUDA(c, cudaMemcpyAsync(c->h_red, cnt, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        EB_CUDA(c, cudaStreamSynchronize(c->stream));
        const long long nrecv = (long long)c->h_red[2], src_left = (long long)c->h_red[3];
        EB_TRY(comm_sendrecv_v(c, c->mig_send[0], nsend > 0 ? dst : -1, (size_t)nsend * nvar,
                               c->mig_recv[0], nrecv > 0 ? src : -1, (size_t)nrecv * nvar));
        if (nrecv > 0) {
          const long long keep_n = sp.n;
          sp.n = n_slots;
          EB_TRY(species_reserve(c, sp, n_slots + nrecv));
          sp.n = keep_n;
          k_unpack_aos<<<(unsigned)((nrecv + 255) / 256), 256, 0, c->stream>>>(c->mig_recv[0], c->ndims, nrecv, n_slots,
              sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], sp.a[5], sp.a[6], sp.key);
          n_slots += nrecv; c->launches++;
        }
        if (left == 0 && src_left == 0) break;
      }
    }
    sp.sorted = false;
    slots[is] = n_slots;
  }
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

// sort_species for the caller of redistribute_particles (after the new grid is set)
int32_t resort_after_redistribute(eb200_ctx* c, const std::vector<long long>& slots) {
  for (size_t is = 0; is < c->species.size(); ++is)
    if (c->species[is].defined) EB_TRY(sort_species(c, c->species[is], slots[is]));
  return 0;
}

int32_t particle_bcs(eb200_ctx* c) {
  for (SpeciesD& sp : c->species) {
    if (!sp.defined) continue;
    long long n_slots = sp.n;
    bool changed = false;
    {
      ScopedPhase ph(c, "bcs");
      if (has_thermal(c, sp)) { EB_TRY(settle_species(c, sp)); EB_TRY(thermal_bcs(c, sp, n_slots)); }
      EB_TRY(species_bcs(c, sp, n_slots, changed));
    }
    EB_TRY(sort_species(c, sp, n_slots, false, sp.was_ordered && (c->push_kernel == 0 || c->push_kernel == 3)));
  }
  return 0;
}

// ---------------------------------------------------------------------------------
// push_particles (particles.F90:28-614)
// ---------------------------------------------------------------------------------
int32_t push_particles(eb200_ctx* c) {
  // jx = jy = jz = 0 (particles.F90:147-149): one memset over the 3 contiguous arrays
  {
    ScopedPhase ph(c, "push");
    EB_CUDA(c, cudaMemsetAsync(c->f.f[JX], 0, (size_t)3 * c->g.total * sizeof(double), c->stream));
  }
  std::vector<char> done(c->species.size(), 0), tiled(c->species.size(), 0);
  const bool mixed = bc_mixed(c);
  for (size_t is = 0; is < c->species.size(); ++is) {
    SpeciesD& sp = c->species[is];
    if (!sp.defined) continue;
    if (sp.immobile || sp.species_type == 1) { EB_TRY(settle_species(c, sp)); continue; }
    // (a re-sort after a window shift or a dt change moves particles by about a tile: the permute can be
    //  left to the push like any other; an unordered load overflows the side buffer and is permuted at once)
    if (!sp.sorted || sp.sort_dt != c->dt)
      EB_TRY(sort_species(c, sp, sp.n, false, sp.was_ordered && (c->push_kernel == 0 || c->push_kernel == 3)));
    if (sp.n == 0) { sp.deferred = false; continue; }
    PushP p = make_push_params(c, sp, sp.n);
    bool handled = false, fused_keys = false;
    int mig_axis = -1;
    {
      ScopedPhase ph(c, "push");
      ScopedPhase pk(c, "push_kernel");     // the kernel(s) alone (roofline figure of bench.py)
      if (c->push_kernel == 0 || c->push_kernel == 3) {
        // single rank: no migration, so the push kernel can apply the particle BCs itself and emit
        // the next sort keys + histogram (saves the sort's pass over the particle arrays)
        std::vector<char> ksbuf(keysink_size());
        const void* ks = nullptr;
        // Several ranks: the same, when the decomposed axes are periodic for the particles (nothing but
        // migration happens there): leavers give their count back when they are packed, arrivals are keyed
        // and counted by a pass over the arrivals alone.
        bool fusable = c->fuse_keys != 0 && !has_thermal(c, sp);     // (thermal faces re-emit before the keys are made)
        for (int d = 0; d < c->ndims; ++d)
          if (c->nprocdir[d] > 1 && (sp.bc[2 * d] != EB200_BC_PERIODIC || sp.bc[2 * d + 1] != EB200_BC_PERIODIC)) fusable = false;
        if (fusable) {
          EB_TRY(ensure_cells(c, sp));
          const uint32_t nbucket = (uint32_t)(c->g.ncell_ext + 1);
          EB_CUDA(c, cudaMemsetAsync(c->cell_count, 0, (size_t)(nbucket + 1) * sizeof(uint32_t), c->stream));
          BcP3 bc;
          for (int d = 0; d < 3; ++d) { bc.on[d] = d < c->ndims && c->nprocdir[d] == 1; bc.b[d] = make_bc(c, sp, d < c->ndims ? d : 0); }
          keysink_fill(ksbuf.data(), bc, sp.key, c->cell_count, nbucket);
          ks = ksbuf.data();
          // one decomposed axis: the push lists its leavers as well (fused migration search)
          int ndec = 0, dax = -1;
          for (int d = 0; d < c->ndims; ++d) if (c->nprocdir[d] > 1) { ++ndec; dax = d; }
          mig_axis = -1;
          if (ndec == 1 && c->fuse_keys) {
            const long long want = std::max<long long>(std::max<long long>(c->mig_list_want, c->mig_min_particles), 1 << 16);
            if (c->mig_list_cap < want) {
              for (int s2 = 0; s2 < 2; ++s2) {
                if (c->mig_list[s2]) { EB_CUDA(c, cudaStreamSynchronize(c->stream)); cudaFree(c->mig_list[s2]); c->mig_list[s2] = nullptr; }
                EB_CUDA(c, cudaMalloc(&c->mig_list[s2], (size_t)want * sizeof(uint32_t)));
              }
              c->mig_list_cap = want;
            }
            EB_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 8 * sizeof(unsigned long long), c->stream));
            const BcP bm = make_bc(c, sp, dax);
            keysink_set_migration(ksbuf.data(), dax, bm.min_local, bm.max_local, c->mig_list[0], c->mig_list[1], c->d_counters,
                                  (uint32_t)std::min<long long>(c->mig_list_cap, 0xFFFFFFF0LL));
            mig_axis = dax;
          }
        }
        EB_TRY(launch_push_split(c, sp, p, &handled, ks));     // consumes a deferred permute
        fused_keys = handled && ks != nullptr;
        tiled[is] = handled;
      }
      if (!handled) {
        if (c->deterministic)
          return fail(c, EB200_E_UNSUPPORTED, "deterministic_deposit needs the tiled kernels (option push_kernel)");
        EB_TRY(settle_species(c, sp));
        EB_TRY(launch_push_baseline(c, sp, p));
      }
    }
    if (mixed) {      // current_bcs(species) (particles.F90:609): fold this species' current with ITS boundaries
      ScopedPhase ph(c, "halo");
      for (int k = 0; k < 3; ++k) EB_TRY(summation_bcs(c, c->f.f[JX + k], k, sp.bc));
    }
    if (fused_keys && c->nproc == 1) { EB_TRY(sort_species(c, sp, sp.n, true, true)); done[is] = 1; }
    else if (fused_keys) {
      const long long n_counted = sp.n;
      long long n_slots = sp.n;
      bool changed = false;
      {
        ScopedPhase ph(c, "bcs");
        EB_TRY(species_bcs(c, sp, n_slots, changed, n_counted, mig_axis));
        if (n_slots > n_counted) {       // arrivals: BCs of the other axes, key, histogram
          BcP3 bc;
          for (int d = 0; d < 3; ++d) { bc.on[d] = d < c->ndims && c->nprocdir[d] == 1; bc.b[d] = make_bc(c, sp, d < c->ndims ? d : 0); }
          PushP pk2 = make_push_params(c, sp, n_slots);
          const uint32_t nbucket = (uint32_t)(c->g.ncell_ext + 1);
          const unsigned ga = (unsigned)((n_slots - n_counted + 255) / 256);
#define LK(ND, SH)                                                                               \
          if (c->ndims == ND && c->shape == SH)                                                  \
            k_keys<ND, SH><<<ga, 256, 0, c->stream>>>(c->g, pk2, bc, sp.a[0], sp.a[1], sp.a[2], sp.a[3], sp.a[4], sp.a[5], sp.key, c->cell_count, nbucket, n_counted);
          LK(1, 0) LK(1, 1) LK(1, 2) LK(2, 0) LK(2, 1) LK(2, 2) LK(3, 0) LK(3, 1) LK(3, 2)
#undef LK
          c->launches++;
        }
      }
      EB_TRY(sort_species(c, sp, n_slots, true, true));
      done[is] = 1;
    }
  }
  // current_bcs(species) is a no-op without mixed per-species BCs (boundary.F90:1130-1136)
  // particle_bcs for the species the push did not already finish
  for (size_t is = 0; is < c->species.size(); ++is) {
    SpeciesD& sp = c->species[is];
    if (!sp.defined || done[is]) continue;
    long long n_slots = sp.n;
    bool changed = false;
    {
      ScopedPhase ph(c, "bcs");
      if (has_thermal(c, sp)) EB_TRY(thermal_bcs(c, sp, n_slots));
      EB_TRY(species_bcs(c, sp, n_slots, changed));
    }
    EB_TRY(sort_species(c, sp, n_slots, false, tiled[is] != 0));
  }
  return 0;
}

}  // namespace eb
