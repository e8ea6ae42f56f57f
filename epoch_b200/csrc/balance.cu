// balance.cu -- device side of the dynamic load balancer (housekeeping/balance.F90).
// The decision stays on the host (balance_workload :107-319, calculate_breaks :2526-2669: a few
// hundred integers); the device supplies what that decision needs and carries it out:
//   * get_load: the per-axis particle histograms of part_load_func (:2468-2495) for this rank,
//     before the MPI_ALLREDUCE of get_load (:2411-2464);
//   * redistribute: redistribute_fields (:476-1293 -> redistribute_field_3d :1652-1966, the
//     intersections of every old sub-domain with every new one, edge ghost cells travelling with
//     the edge cells, followed by do_field_mpi_with_lengths) for the nine field arrays and the
//     CPML psi arrays, and distribute_particles (:2751-2809) -- both over NCCL, rank to rank.
#include "ctx.cuh"
#include <vector>

namespace eb {

int32_t resort_after_redistribute(eb200_ctx* c, const std::vector<long long>& slots);   // particles.cu

namespace {

struct LoadP { int nd, tophat, len[3], ng; double gmin[3], d[3]; unsigned long long* load[3]; };
// part_load_func: cell = FLOOR((pos - x_grid_min) / dx + 1.5) (tophat: FLOOR(..) + 1), an index
// into load(1-ng : n_global+ng)
__global__ void k_load(LoadP p, long long n, const double* __restrict__ x, const double* __restrict__ y,
                       const double* __restrict__ z) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double pos[3] = {x[i], p.nd >= 2 ? y[i] : 0.0, p.nd >= 3 ? z[i] : 0.0};
  for (int a = 0; a < p.nd; ++a) {
    const double r = (pos[a] - p.gmin[a]) / p.d[a];
    int cell = p.tophat ? (int)floor(r) + 1 : (int)floor(r + 1.5);
    cell = min(max(cell + p.ng - 1, 0), p.len[a] - 1);      // position in the array (1-ng -> 0)
    atomicAdd(p.load[a] + cell, 1ULL);
  }
}

struct BoxP { long long st[3]; int org[3]; int dim[3]; };
__global__ void k_box_pack(const double* __restrict__ a, BoxP b, double* __restrict__ buf, long long count) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count) return;
  const int i = (int)(t % b.dim[0]);
  const long long r = t / b.dim[0];
  const int j = (int)(r % b.dim[1]), k = (int)(r / b.dim[1]);
  buf[t] = a[(b.org[0] + i) * b.st[0] + (b.org[1] + j) * b.st[1] + (b.org[2] + k) * b.st[2]];
}
__global__ void k_box_unpack(double* __restrict__ a, BoxP b, const double* __restrict__ buf, long long count) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count) return;
  const int i = (int)(t % b.dim[0]);
  const long long r = t / b.dim[0];
  const int j = (int)(r % b.dim[1]), k = (int)(r / b.dim[1]);
  a[(b.org[0] + i) * b.st[0] + (b.org[1] + j) * b.st[1] + (b.org[2] + k) * b.st[2]] = buf[t];
}

// [lo, hi] of a sub-domain along one axis, widened by the ghost cells where it touches the edge of
// the global grid (ng0 / ng1 of redistribute_field_3d :1741-1744)
inline void ext_range(int lo, int hi, int nglob, int ng, int& elo, int& ehi) {
  elo = lo == 1 ? lo - ng : lo;
  ehi = hi == nglob ? hi + ng : hi;
}

}  // namespace

int32_t get_load(eb200_ctx* c, const double grid_min_global[3], long long* const load[3]) {
  const GridP& g = c->g;
  LoadP p;
  p.nd = c->ndims; p.tophat = c->shape == SHAPE_TOPHAT; p.ng = c->ng;
  size_t total = 0;
  for (int a = 0; a < 3; ++a) {
    p.len[a] = a < c->ndims ? c->nglob[a] + 2 * c->ng : 0;
    p.gmin[a] = a < c->ndims ? grid_min_global[a] : 0.0;
    p.d[a] = g.d[a];
    total += (size_t)p.len[a];
  }
  unsigned long long* d_load = nullptr;
  EB_CUDA(c, cudaMalloc(&d_load, total * sizeof(unsigned long long)));
  EB_CUDA(c, cudaMemsetAsync(d_load, 0, total * sizeof(unsigned long long), c->stream));
  size_t off = 0;
  for (int a = 0; a < 3; ++a) { p.load[a] = d_load + off; off += (size_t)p.len[a]; }
  for (SpeciesD& sp : c->species) {
    if (!sp.defined || sp.n == 0) continue;
    k_load<<<(unsigned)((sp.n + 255) / 256), 256, 0, c->stream>>>(p, sp.n, sp.a[0], sp.a[1], sp.a[2]);
    c->launches++;
  }
  std::vector<unsigned long long> h(total);
  cudaError_t e = cudaMemcpyAsync(h.data(), d_load, total * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cudaFree(d_load);
  EB_CUDA(c, e);
  off = 0;
  for (int a = 0; a < c->ndims; ++a) {
    if (load[a]) for (int i = 0; i < p.len[a]; ++i) load[a][i] = (long long)h[off + i];
    off += (size_t)p.len[a];
  }
  return 0;
}

// tables: [x entries | y entries | z entries], nprocdir[a] entries per existing axis, 1-based
// inclusive global cell ranges per processor coordinate (cell_x_min, cell_x_max, ... of
// balance.F90 and their new_* counterparts)
int32_t redistribute(eb200_ctx* c, const int32_t* old_cell_min, const int32_t* old_cell_max, const int32_t* new_cell_min,
                     const int32_t* new_cell_max, const double grid_min_global[3], const double grid_min_local[3],
                     const double min_local[3], const double max_local[3]) {
  const int nd = c->ndims, P = c->nproc, ng = c->ng;
  const int32_t *omin[3], *omax[3], *nmin[3], *nmax[3];
  {
    int off = 0;
    for (int a = 0; a < 3; ++a) {
      omin[a] = old_cell_min + off; omax[a] = old_cell_max + off; nmin[a] = new_cell_min + off; nmax[a] = new_cell_max + off;
      if (a < nd) off += c->nprocdir[a];
    }
  }
  OwnerP own;
  for (int a = 0; a < 3; ++a) {
    own.np[a] = a < nd ? c->nprocdir[a] : 1;
    if (own.np[a] > MAXP_AXIS) return fail(c, EB200_E_UNSUPPORTED, "more than 32 processors along one axis");
    for (int p = 0; p < own.np[a]; ++p) {
      if (a >= nd) { own.lb[a][p] = 0.0; continue; }
      if (omin[a][p] != c->nglob_min[a] && p == c->coords[a])
        return fail(c, EB200_E_ARG, "redistribute: old_cell_min does not match the current grid of this rank");
      if (nmax[a][p] < nmin[a][p] || (p > 0 && nmin[a][p] != nmax[a][p - 1] + 1) || (p == 0 && nmin[a][p] != 1) ||
          (p == own.np[a] - 1 && nmax[a][p] != c->nglob[a]))
        return fail(c, EB200_E_ARG, "redistribute: new cell ranges do not tile the global grid");
      // x_grid_mins(p) - dx / 2 with x_grid_mins(p) = x_grid_min + (cell_min - 1) * dx
      own.lb[a][p] = (grid_min_global[a] + (double)(nmin[a][p] - 1) * c->g.d[a]) - c->g.d[a] * 0.5;
    }
  }
  // ---- particles (independent of the grid arrays) ----
  std::vector<long long> slots;
  EB_TRY(redistribute_particles(c, own, slots));

  // ---- fields: keep the old arrays, set the new grid, move the intersections ----
  const GridP og = c->g;
  double* old_f = c->fbase;
  double* old_psi = c->cpml_psi;
  c->fbase = nullptr; c->cpml_psi = nullptr;
  double* old_bnd[6];        // (set_grid frees every array of the old field shape: the boundary planes are decided on below)
  for (int b = 0; b < 6; ++b) { old_bnd[b] = c->bnd_planes[b]; c->bnd_planes[b] = nullptr; }
  int32_t nl[3] = {1, 1, 1}, ngm[3] = {1, 1, 1}, ngl[3] = {1, 1, 1};
  double dd[3], gml[3] = {0, 0, 0}, mnl[3] = {0, 0, 0}, mxl[3] = {0, 0, 0}, gmn[3], gmx[3];
  for (int a = 0; a < 3; ++a) {
    dd[a] = og.d[a]; gmn[a] = og.gmin[a]; gmx[a] = og.gmax[a];
    ngl[a] = c->nglob[a];
    if (a >= nd) continue;
    nl[a] = nmax[a][c->coords[a]] - nmin[a][c->coords[a]] + 1;
    ngm[a] = nmin[a][c->coords[a]];
    gml[a] = grid_min_local[a]; mnl[a] = min_local[a]; mxl[a] = max_local[a];
  }
  int32_t rc = eb200_set_grid(c, nl, ngl, ngm, dd, gml, mnl, mxl, gmn, gmx, c->cpml_thickness);
  if (rc != 0) { cudaFree(old_f); if (old_psi) cudaFree(old_psi); for (int b = 0; b < 6; ++b) if (old_bnd[b]) cudaFree(old_bnd[b]); return rc; }
  const GridP& g = c->g;
  const int npsi = old_psi ? 12 : 0;
  if (npsi) {
    EB_CUDA(c, cudaMalloc(&c->cpml_psi, (size_t)12 * g.total * sizeof(double)));
    EB_CUDA(c, cudaMemsetAsync(c->cpml_psi, 0, (size_t)12 * g.total * sizeof(double), c->stream));
  }
  // boxes: send to r = old_ext(me) ^ new_ext(r); receive from r = old_ext(r) ^ new_ext(me)
  std::vector<BoxP> sbox(P), rbox(P);
  std::vector<size_t> scnt(P, 0), rcnt(P, 0), soff(P, 0), roff(P, 0);
  size_t stot = 0, rtot = 0;
  for (int r = 0; r < P; ++r) {
    const int rc3[3] = {r % c->nprocdir[0], (r / c->nprocdir[0]) % c->nprocdir[1], r / (c->nprocdir[0] * c->nprocdir[1])};
    size_t sv = 1, rv = 1;
    for (int a = 0; a < 3; ++a) {
      sbox[r].st[a] = og.st[a]; rbox[r].st[a] = g.st[a];
      if (a >= nd) { sbox[r].org[a] = rbox[r].org[a] = 0; sbox[r].dim[a] = rbox[r].dim[a] = 1; continue; }
      int olo, ohi, nlo, nhi, lo, hi;
      // what I held, what r will hold
      ext_range(omin[a][c->coords[a]], omax[a][c->coords[a]], c->nglob[a], ng, olo, ohi);
      ext_range(nmin[a][rc3[a]], nmax[a][rc3[a]], c->nglob[a], ng, nlo, nhi);
      lo = std::max(olo, nlo); hi = std::min(ohi, nhi);
      sbox[r].dim[a] = std::max(0, hi - lo + 1);
      sbox[r].org[a] = lo - omin[a][c->coords[a]] + ng;          // array position of global cell lo in my old array
      // what r held, what I will hold
      ext_range(omin[a][rc3[a]], omax[a][rc3[a]], c->nglob[a], ng, olo, ohi);
      ext_range(nmin[a][c->coords[a]], nmax[a][c->coords[a]], c->nglob[a], ng, nlo, nhi);
      lo = std::max(olo, nlo); hi = std::min(ohi, nhi);
      rbox[r].dim[a] = std::max(0, hi - lo + 1);
      rbox[r].org[a] = lo - nmin[a][c->coords[a]] + ng;
      sv *= (size_t)sbox[r].dim[a]; rv *= (size_t)rbox[r].dim[a];
    }
    for (int a = 0; a < nd; ++a) { if (sbox[r].dim[a] == 0) sv = 0; if (rbox[r].dim[a] == 0) rv = 0; }
    scnt[r] = sv; rcnt[r] = rv; soff[r] = stot; roff[r] = rtot;
    stot += sv; rtot += rv;
  }
  double *sbuf = nullptr, *rbuf = nullptr;
  EB_CUDA(c, cudaMalloc(&sbuf, std::max<size_t>(stot, 1) * sizeof(double)));
  EB_CUDA(c, cudaMalloc(&rbuf, std::max<size_t>(rtot, 1) * sizeof(double)));
  int32_t err = 0;
  for (int k = 0; k < NFIELD + npsi && err == 0; ++k) {
    const double* src = k < NFIELD ? old_f + (size_t)k * og.total : old_psi + (size_t)(k - NFIELD) * og.total;
    double* dst = k < NFIELD ? c->f.f[k] : c->cpml_psi + (size_t)(k - NFIELD) * g.total;
    for (int r = 0; r < P; ++r)
      if (scnt[r]) {
        k_box_pack<<<(unsigned)((scnt[r] + 255) / 256), 256, 0, c->stream>>>(src, sbox[r], sbuf + soff[r], (long long)scnt[r]);
        c->launches++;
      }
    err = comm_alltoallv(c, sbuf, soff.data(), scnt.data(), rbuf, roff.data(), rcnt.data());
    for (int r = 0; r < P && err == 0; ++r)
      if (rcnt[r]) {
        k_box_unpack<<<(unsigned)((rcnt[r] + 255) / 256), 256, 0, c->stream>>>(dst, rbox[r], rbuf + roff[r], (long long)rcnt[r]);
        c->launches++;
      }
  }
  cudaStreamSynchronize(c->stream);
  cudaFree(sbuf); cudaFree(rbuf); cudaFree(old_f);
  if (old_psi) cudaFree(old_psi);
  // the boundary planes of setup_field_boundaries (the reference's host remaps its own copies, balance.F90:847-1010):
  // a plane whose extents changed belongs to the old slicing and is dropped -- outflow_bcs then refuses to run
  // until the host has handed it over again (set_boundary_fields)
  for (int b = 0; b < 6; ++b) {
    if (!old_bnd[b]) continue;
    const int a = b / 2;
    if (og.n[(a + 1) % 3] == g.n[(a + 1) % 3] && og.n[(a + 2) % 3] == g.n[(a + 2) % 3]) { c->bnd_planes[b] = old_bnd[b]; continue; }
    cudaFree(old_bnd[b]); c->bnd_dropped[b] = true;
  }
  if (err) return err;
  EB_CUDA(c, cudaGetLastError());
  // do_field_mpi_with_lengths on every remapped array (remap_field :1371-1374)
  EB_TRY(field_bc(c, EX, 3));
  EB_TRY(field_bc(c, BX, 3));
  EB_TRY(field_bc(c, JX, 3));
  for (int k = 0; k < npsi; ++k) EB_TRY(field_bc_array(c, c->cpml_psi + (size_t)k * g.total));
  // particles: keys and cell lists of the new grid
  EB_TRY(resort_after_redistribute(c, slots));
  return 0;
}

}  // namespace eb
