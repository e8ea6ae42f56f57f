// capi.cu -- the C ABI of include/epoch_b200.h (extern "C", plain pointers and
// sizes).  One context per rank == per GPU; all work is enqueued on the context's
// private stream.  There is no CPU fallback: without a CUDA device every entry
// that touches data fails with EB200_E_CUDA.
#include "ctx.cuh"
#include <cstring>
#include <cstdlib>
#include <new>

static thread_local std::string g_last_error;

namespace eb {

int32_t fail(eb200_ctx* c, int32_t code, const std::string& msg) {
  if (c) c->err = msg;
  g_last_error = msg;
  return code;
}

ScopedPhase::ScopedPhase(eb200_ctx* c_, const char* n) : c(c_), name(n) {
  if (!c->timers_on) return;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, c->stream);
  launches0 = c->launches;
}
ScopedPhase::~ScopedPhase() {
  if (!e0) return;
  cudaEventRecord(e1, c->stream);
  PhaseTimer& t = c->timers[name];
  t.ev.emplace_back(e0, e1);
  t.launches += c->launches - launches0;
}

static void free_species(SpeciesD& sp) {
  for (int k = 0; k < 7; ++k) { if (sp.base[k]) cudaFree(sp.base[k]); sp.base[k] = sp.a[k] = nullptr; }
  if (sp.key_base) cudaFree(sp.key_base);
  sp.key_base = nullptr; sp.off = 0; sp.deferred = false;
  if (sp.probes.buf) cudaFree(sp.probes.buf);
  if (sp.probes.count) cudaFree(sp.probes.count);
  sp.probes = ProbeSet();
  if (sp.h_mig) cudaFreeHost(sp.h_mig);
  if (sp.d_mig) cudaFree(sp.d_mig);
  sp.h_mig = sp.d_mig = nullptr;
  if (sp.far) cudaFree(sp.far);
  sp.far = nullptr; sp.far_cap = 0;
  for (int b = 0; b < 6; ++b) { if (sp.ext_temp[b]) cudaFree(sp.ext_temp[b]); sp.ext_temp[b] = nullptr; }
  if (sp.dest) cudaFree(sp.dest);
  if (sp.d_tile_start) cudaFree(sp.d_tile_start);
  if (sp.h_tile_start) cudaFreeHost(sp.h_tile_start);
  sp.d_tile_start = sp.h_tile_start = nullptr; sp.tile_cap = 0;
  if (sp.cell_start) cudaFree(sp.cell_start);
  sp.cell_start = nullptr; sp.cell_cap = 0;
  sp.key = sp.dest = nullptr; sp.n = sp.cap = 0;
}

}  // namespace eb
using namespace eb;

#define CHECK_CTX(c) do { if (!(c)) return fail(nullptr, EB200_E_ARG, "null context"); cudaSetDevice((c)->device); } while (0)
#define NEED_GRID(c) do { if (!(c)->grid_set) return fail((c), EB200_E_STATE, "set_grid has not been called"); } while (0)
#define NEED_DT(c) do { if (!((c)->dt > 0)) return fail((c), EB200_E_STATE, "set_time has not been called (dt <= 0)"); } while (0)

extern "C" {

int32_t eb200_create(eb200_ctx** out, int32_t ndims, int32_t shape, int32_t device, int32_t rank,
                     int32_t nproc, const int32_t nprocdir[3], const int32_t coords[3], const void* nccl_id) {
  if (!out) return fail(nullptr, EB200_E_ARG, "out is null");
  *out = nullptr;
  if (ndims < 1 || ndims > 3) return fail(nullptr, EB200_E_ARG, "ndims must be 1, 2 or 3");
  if (shape < 0 || shape > 2) return fail(nullptr, EB200_E_ARG, "shape must be 0 (tophat), 1 (triangle) or 2 (bspline3)");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(nullptr, EB200_E_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) + " (this library has no CPU path)");
  if (device < 0) device = (-(device + 1)) % ndev;      // -(local_rank + 1): the library wraps the node-local rank onto the devices present
  if (device >= ndev) return fail(nullptr, EB200_E_ARG, "bad device ordinal");
  eb200_ctx* c = new (std::nothrow) eb200_ctx();
  if (!c) return fail(nullptr, EB200_E_NOMEM, "host allocation failed");
  c->ndims = ndims; c->shape = shape; c->device = device; c->rank = rank; c->nproc = nproc < 1 ? 1 : nproc;
  int np = 1;
  for (int d = 0; d < 3; ++d) {
    c->nprocdir[d] = (nprocdir && d < ndims) ? nprocdir[d] : 1;
    c->coords[d] = (coords && d < ndims) ? coords[d] : 0;
    if (c->nprocdir[d] < 1 || c->coords[d] < 0 || c->coords[d] >= c->nprocdir[d]) { delete c; return fail(nullptr, EB200_E_ARG, "bad nprocdir/coords"); }
    np *= c->nprocdir[d];
    c->min_bnd[d] = c->coords[d] == 0;
    c->max_bnd[d] = c->coords[d] == c->nprocdir[d] - 1;
  }
  if (np != c->nproc) { delete c; return fail(nullptr, EB200_E_ARG, "product of nprocdir != nproc"); }
  // neighbours of the (nprocz,nprocy,nprocx) Cartesian communicator, always with
  // wrap-around; the call sites decide whether a wrap link is used
  for (int d = 0; d < 3; ++d) {
    int cl[3] = {c->coords[0], c->coords[1], c->coords[2]}, ch[3] = {c->coords[0], c->coords[1], c->coords[2]};
    cl[d] = (cl[d] - 1 + c->nprocdir[d]) % c->nprocdir[d];
    ch[d] = (ch[d] + 1) % c->nprocdir[d];
    c->nbr_lo[d] = (cl[2] * c->nprocdir[1] + cl[1]) * c->nprocdir[0] + cl[0];
    c->nbr_hi[d] = (ch[2] * c->nprocdir[1] + ch[1]) * c->nprocdir[0] + ch[0];
  }
  // constants.F90:472-492
  if (shape == 2) { c->sf_min = -2; c->sf_max = 2; c->png = 4; }
  else if (shape == 0) { c->sf_min = 0; c->sf_max = 1; c->png = 2; }
  else { c->sf_min = -1; c->sf_max = 1; c->png = 3; }
  c->ng = c->png + 2;
  cudaSetDevice(device);
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaMalloc(&c->d_counters, 64 * sizeof(unsigned long long)) != cudaSuccess ||
      cudaMemset(c->d_counters, 0, 64 * sizeof(unsigned long long)) != cudaSuccess ||
      cudaMallocHost(&c->h_counters, 64 * sizeof(unsigned long long)) != cudaSuccess ||
      cudaMalloc(&c->d_red, 64 * sizeof(double)) != cudaSuccess ||
      cudaMallocHost(&c->h_red, 64 * sizeof(double)) != cudaSuccess) {
    std::string m = cudaGetErrorString(cudaGetLastError());
    delete c;
    return fail(nullptr, EB200_E_CUDA, "context setup failed: " + m);
  }
  const char* t = getenv("EB200_TIMERS");
  c->timers_on = t && t[0] == '1';
  int32_t rc = comm_init(c, nccl_id);
  if (rc != 0) { std::string m = c->err; eb200_destroy(c); return fail(nullptr, rc, m); }
  *out = c;
  return 0;
}

int32_t eb200_nccl_unique_id(void* out128) {
  if (!out128) return fail(nullptr, EB200_E_ARG, "null");
  return nccl_unique_id(out128);
}

int32_t eb200_host_alloc(void** out, int64_t bytes) {
  if (!out || bytes < 0) return EB200_E_ARG;
  *out = nullptr;
  if (bytes == 0) return 0;
  const cudaError_t e = cudaHostAlloc(out, (size_t)bytes, cudaHostAllocPortable);
  if (e != cudaSuccess) { cudaGetLastError(); *out = nullptr; return e == cudaErrorMemoryAllocation ? EB200_E_NOMEM : EB200_E_CUDA; }
  return 0;
}
int32_t eb200_host_free(void* p) {
  if (p && cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return EB200_E_CUDA; }
  return 0;
}

int32_t eb200_destroy(eb200_ctx* c) {
  if (!c) return 0;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  comm_destroy(c);
  for (SpeciesD& sp : c->species) free_species(sp);
  for (int b = 0; b < 6; ++b) for (Laser& L : c->lasers[b]) { if (L.profile) cudaFree(L.profile); if (L.phase) cudaFree(L.phase); }
  void* ptrs[] = {c->fbase, c->attr_scratch, c->cell_count, c->rec_scratch, c->det_tiles, c->mig_list[0], c->mig_list[1], c->cub_tmp, c->halo_send, c->halo_recv,
                  c->mig_send[0], c->mig_send[1], c->mig_recv[0], c->mig_recv[1], c->d_counters, c->d_red, c->res_ref, c->rho, c->smooth_wk, c->win_planes, c->cpml_psi, c->diag_wk,
                  c->cpml_prof[0], c->cpml_prof[1], c->cpml_prof[2]};
  for (void* p : ptrs) if (p) cudaFree(p);
  for (int b = 0; b < 6; ++b) if (c->bnd_planes[b]) cudaFree(c->bnd_planes[b]);
  if (c->d_probeset) cudaFree(c->d_probeset);
  free_pinned(c);
  if (c->h_counters) cudaFreeHost(c->h_counters);
  if (c->h_red) cudaFreeHost(c->h_red);
  for (auto& kv : c->timers) for (auto& ev : kv.second.ev) { cudaEventDestroy(ev.first); cudaEventDestroy(ev.second); }
  for (int k = 0; k < 2; ++k) { if (c->ev_push[k]) cudaEventDestroy(c->ev_push[k]); if (c->ev_dep[k]) cudaEventDestroy(c->ev_dep[k]); }
  if (c->stream2) cudaStreamDestroy(c->stream2);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
  return 0;
}

int32_t eb200_set_grid(eb200_ctx* c, const int32_t n_local[3], const int32_t n_global[3], const int32_t n_global_min[3],
                       const double d[3], const double grid_min_local[3], const double min_local[3],
                       const double max_local[3], const double min_global[3], const double max_global[3],
                       int32_t cpml_thickness) {
  CHECK_CTX(c);
  if (!n_local || !d || !grid_min_local || !min_local || !max_local || !min_global || !max_global)
    return fail(c, EB200_E_ARG, "null argument");
  if (cpml_thickness < 0) return fail(c, EB200_E_ARG, "cpml_thickness < 0");
  c->cpml_thickness = cpml_thickness;
  GridP& g = c->g;
  g.ndims = c->ndims; g.shape = c->shape; g.ng = c->ng;
  long long stride = 1, ncell = 1;
  for (int k = 0; k < 3; ++k) {
    const bool ex = k < c->ndims;
    g.n[k] = ex ? n_local[k] : 1;
    if (g.n[k] < 1) return fail(c, EB200_E_ARG, "n_local < 1");
    if (ex && d[k] <= 0) return fail(c, EB200_E_ARG, "cell size <= 0");
    c->nglob[k] = ex && n_global ? n_global[k] : g.n[k];
    c->nglob_min[k] = ex && n_global_min ? n_global_min[k] : 1;
    g.ngd[k] = ex ? c->ng : 0;
    g.d[k] = ex ? d[k] : 1.0; g.id[k] = 1.0 / g.d[k];
    g.gml[k] = ex ? grid_min_local[k] : 0.0;
    g.min_local[k] = ex ? min_local[k] : 0.0; g.max_local[k] = ex ? max_local[k] : 0.0;
    g.gmin[k] = ex ? min_global[k] : 0.0; g.gmax[k] = ex ? max_global[k] : 0.0;
    g.st[k] = stride;
    stride *= (long long)(g.n[k] + 2 * g.ngd[k]);
    g.kn[k] = ex ? g.n[k] + 2 * KM : 1;
    g.tdim[k] = tile_dim(c->ndims, k);
    g.ntile[k] = (g.kn[k] + g.tdim[k] - 1) / g.tdim[k];
    ncell *= (long long)g.ntile[k] * g.tdim[k];
  }
  g.cpt = g.tdim[0] * g.tdim[1] * g.tdim[2];
  if (ncell >= 0xFFFFFFF0LL) return fail(c, EB200_E_UNSUPPORTED, "local grid too large for 32-bit cell keys");
  g.total = stride; g.ncell_ext = ncell;
  if (c->fbase) { cudaStreamSynchronize(c->stream); cudaFree(c->fbase); c->fbase = nullptr; }
  // every work array of the field shape belongs to the old grid (set_grid again = rebalance / restart)
  double** shaped[] = {&c->rho, &c->res_ref, &c->smooth_wk, &c->win_planes, &c->cpml_psi, &c->diag_wk,
                       &c->cpml_prof[0], &c->cpml_prof[1], &c->cpml_prof[2], &c->bnd_planes[0], &c->bnd_planes[1],
                       &c->bnd_planes[2], &c->bnd_planes[3], &c->bnd_planes[4], &c->bnd_planes[5]};
  for (double** p : shaped)
    if (*p) { cudaFree(*p); *p = nullptr; }
  EB_CUDA(c, cudaMalloc(&c->fbase, (size_t)NFIELD * g.total * sizeof(double)));
  EB_CUDA(c, cudaMemsetAsync(c->fbase, 0, (size_t)NFIELD * g.total * sizeof(double), c->stream));
  for (int f = 0; f < NFIELD; ++f) c->f.f[f] = c->fbase + (size_t)f * g.total;
  for (SpeciesD& sp : c->species) { sp.sorted = false; sp.was_ordered = false; }
  c->grid_set = true;
  return 0;
}

int32_t eb200_set_time(eb200_ctx* c, double dt, double time, int32_t step) {
  CHECK_CTX(c);
  if (!(dt > 0)) return fail(c, EB200_E_ARG, "dt must be > 0");
  // (the sort keys depend on dt: push_particles re-sorts a species whose keys were made for another dt; the
  //  dt/2 excursion of the initial bfield_final_bcs, epoch3d.F90:153-159, costs nothing)
  c->dt = dt; c->time = time; c->step = step;
  return 0;
}

int32_t eb200_set_field_bcs(eb200_ctx* c, const int32_t bc_field[6]) {
  CHECK_CTX(c);
  if (!bc_field) return fail(c, EB200_E_ARG, "null");
  // setup_boundaries, boundary.F90:30-59
  for (int i = 0; i < 6; ++i) {
    int bc = i < 2 * c->ndims ? bc_field[i] : EB200_BC_PERIODIC;
    c->add_laser[i] = false;
    if (bc == EB200_BC_OTHER) bc = EB200_BC_CLAMP;
    if (bc == EB200_BC_SIMPLE_LASER) c->add_laser[i] = true;       // cpml_laser: set by set_cpml (one rank only)
    if (bc == EB200_BC_REFLECT) bc = EB200_BC_CLAMP;
    if (bc == EB200_BC_OPEN) bc = EB200_BC_SIMPLE_OUTFLOW;
    if (bc != EB200_BC_PERIODIC && bc != EB200_BC_CLAMP && bc != EB200_BC_SIMPLE_LASER && bc != EB200_BC_SIMPLE_OUTFLOW &&
        bc != EB200_BC_ZERO_GRADIENT && bc != EB200_BC_CONDUCT && bc != EB200_BC_CPML_LASER && bc != EB200_BC_CPML_OUTFLOW)
      return fail(c, EB200_E_ARG, "unrecognised field boundary code");
    c->bc_field[i] = bc;
  }
  for (int d = 0; d < c->ndims; ++d)
    if ((c->bc_field[2 * d] == EB200_BC_PERIODIC) != (c->bc_field[2 * d + 1] == EB200_BC_PERIODIC))
      return fail(c, EB200_E_ARG, "periodic field boundaries must come in min/max pairs");
  c->cpml_on = false;
  for (int i = 0; i < 2 * c->ndims; ++i)
    if (c->bc_field[i] == EB200_BC_CPML_LASER || c->bc_field[i] == EB200_BC_CPML_OUTFLOW) c->cpml_on = true;
  return 0;
}

int32_t eb200_set_cpml(eb200_ctx* c, int32_t axis, const double* kappa_e, const double* kappa_b, const double* sigma_e,
                       const double* sigma_b, const double* a_e, const double* a_b, int32_t min_start, int32_t min_end,
                       int32_t max_start, int32_t max_end, int32_t laser_idx_min, int32_t laser_idx_max) {
  CHECK_CTX(c); NEED_GRID(c);
  if (axis < 0 || axis >= c->ndims) return fail(c, EB200_E_ARG, "axis");
  if (!kappa_e || !kappa_b || !sigma_e || !sigma_b || !a_e || !a_b) return fail(c, EB200_E_ARG, "null profile");
  const size_t len = (size_t)(c->g.n[axis] + 2 * c->ng);
  if (!c->cpml_prof[axis]) EB_CUDA(c, cudaMalloc(&c->cpml_prof[axis], 6 * len * sizeof(double)));
  const double* src[6] = {kappa_e, kappa_b, sigma_e, sigma_b, a_e, a_b};
  for (int k = 0; k < 6; ++k)
    EB_CUDA(c, cudaMemcpyAsync(c->cpml_prof[axis] + k * len, src[k], len * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  c->cpml_start[2 * axis] = min_start; c->cpml_end[2 * axis] = min_end;
  c->cpml_start[2 * axis + 1] = max_start; c->cpml_end[2 * axis + 1] = max_end;
  // a negative index = the laser plane of that side is not on this rank (add_laser stays false)
  for (int side = 0; side < 2; ++side) {
    const int b = 2 * axis + side, idx = side ? laser_idx_max : laser_idx_min;
    if (c->bc_field[b] == EB200_BC_CPML_LASER) { c->add_laser[b] = idx >= 0; c->cpml_laser_idx[b] = idx >= 0 ? idx : 1; }
  }
  if (!c->cpml_psi) {
    EB_CUDA(c, cudaMalloc(&c->cpml_psi, (size_t)12 * c->g.total * sizeof(double)));
    EB_CUDA(c, cudaMemsetAsync(c->cpml_psi, 0, (size_t)12 * c->g.total * sizeof(double), c->stream));
  }
  return 0;
}

int32_t eb200_set_solver(eb200_ctx* c, int32_t field_order, int32_t maxwell_solver, const double coeffs[15]) {
  CHECK_CTX(c);
  if (field_order != 2 && field_order != 4 && field_order != 6) return fail(c, EB200_E_ARG, "field_order must be 2, 4 or 6");
  // constants.F90:125-132: custom -1, yee 0, lehe_x/y/z 2/3/4, cowan 5, pukhov 6 (1 = "lehe" is an
  // alias the deck reader resolves); the coefficients come from set_maxwell_solver on the host
  if (maxwell_solver < -1 || maxwell_solver > 6 || maxwell_solver == 1) return fail(c, EB200_E_ARG, "unknown maxwell_solver");
  if (maxwell_solver != 0 && !coeffs) return fail(c, EB200_E_ARG, "an extended-stencil solver needs its 15 coefficients");
  if (field_order / 2 > c->ng) return fail(c, EB200_E_ARG, "field_order needs more ghost cells");
  c->field_order = field_order; c->maxwell_solver = maxwell_solver;
  if (coeffs) std::memcpy(c->coeffs, coeffs, 15 * sizeof(double));
  return 0;
}

int32_t eb200_set_current_options(eb200_ctx* c, int32_t smooth, int32_t its, int32_t comp_its, int32_t nstrides,
                                  const int32_t* strides, int32_t use_corr, const double initial_j[3]) {
  CHECK_CTX(c);
  if (its < 0 || comp_its < 0 || nstrides < 0 || (nstrides > 0 && !strides)) return fail(c, EB200_E_ARG, "bad smoothing parameters");
  for (int k = 0; k < nstrides; ++k)
    if (strides[k] < 1 || strides[k] > c->ng) return fail(c, EB200_E_ARG, "smooth_strides must lie in 1..ng");
  if (use_corr && !initial_j) return fail(c, EB200_E_ARG, "current correction needs initial_j");
  c->smooth = smooth != 0; c->smooth_its = its; c->smooth_comp_its = comp_its;
  c->smooth_strides.assign(strides, strides + nstrides);
  if (c->smooth_strides.empty()) c->smooth_strides.push_back(1);
  c->use_current_correction = use_corr != 0;
  for (int k = 0; k < 3; ++k) c->initial_j[k] = use_corr ? initial_j[k] : 0.0;
  return 0;
}

int32_t eb200_set_window_fields(eb200_ctx* c, const double* ex, const double* ey, const double* ez, const double* bx,
                                const double* by, const double* bz) {
  CHECK_CTX(c); NEED_GRID(c);
  const double* pl[6] = {ex, ey, ez, bx, by, bz};
  return set_window_planes(c, pl);
}
int32_t eb200_setup_field_boundaries(eb200_ctx* c) {
  CHECK_CTX(c); NEED_GRID(c);
  return setup_field_boundaries(c);
}
int32_t eb200_set_boundary_fields(eb200_ctx* c, int32_t boundary, const double* ex, const double* ey, const double* ez,
                                  const double* bx, const double* by, const double* bz) {
  CHECK_CTX(c); NEED_GRID(c);
  if (boundary < 1 || boundary > 2 * c->ndims) return fail(c, EB200_E_ARG, "boundary out of range");
  const double* pl[6] = {ex, ey, ez, bx, by, bz};
  return set_boundary_fields(c, boundary - 1, pl);
}
int32_t eb200_shift_window(eb200_ctx* c, double x_grid_min_local, double x_min_local, double x_max_local, double x_min,
                           double x_max) {
  CHECK_CTX(c); NEED_GRID(c);
  if (c->g.n[0] < 3) return fail(c, EB200_E_ARG, "moving window needs nx >= 3");
  EB_TRY(settle_particles(c));
  return shift_window(c, x_grid_min_local, x_min_local, x_max_local, x_min, x_max);
}

int32_t eb200_set_species(eb200_ctx* c, int32_t isp, double charge, double mass, double weight, int32_t immobile,
                          int32_t zero_current, int32_t species_type, const int32_t bc_particle[6]) {
  CHECK_CTX(c);
  if (isp < 0 || isp > 62) return fail(c, EB200_E_ARG, "species index out of range (0..62)");
  if (!(mass > 0) && species_type != 1) return fail(c, EB200_E_ARG, "mass must be > 0");
  if ((size_t)isp >= c->species.size()) c->species.resize(isp + 1);
  SpeciesD& sp = c->species[isp];
  int bcs[6];
  for (int i = 0; i < 6; ++i) {
    int bc = (bc_particle && i < 2 * c->ndims) ? bc_particle[i] : EB200_BC_PERIODIC;
    // setup_particle_boundary, boundary.F90:98-140
    if (bc == EB200_BC_OTHER || bc == EB200_BC_CONDUCT) bc = EB200_BC_REFLECT;
    if (bc == EB200_BC_SIMPLE_LASER || bc == EB200_BC_SIMPLE_OUTFLOW || bc == EB200_BC_CPML_LASER || bc == EB200_BC_CPML_OUTFLOW) bc = EB200_BC_OPEN;
    // heat_bath = open boundary + a boundary injector (setup_domain_dependent_boundaries, boundary.F90:75-93); injectors stay on the host
    if (bc == EB200_BC_HEAT_BATH) bc = EB200_BC_OPEN;
    if (bc != EB200_BC_PERIODIC && bc != EB200_BC_REFLECT && bc != EB200_BC_OPEN && bc != EB200_BC_THERMAL)
      return fail(c, EB200_E_ARG, "unrecognised particle boundary code");
    bcs[i] = bc;
  }
  for (int d = 0; d < c->ndims; ++d)
    if ((bcs[2 * d] == EB200_BC_PERIODIC) != (bcs[2 * d + 1] == EB200_BC_PERIODIC))
      return fail(c, EB200_E_ARG, "periodic particle boundaries must come in min/max pairs");
  // (species may differ in their boundaries: bc_allspecies = c_bc_mixed, current_bcs(species) per species)
  sp.defined = true; sp.charge = charge; sp.mass = mass; sp.weight = weight;
  sp.immobile = immobile; sp.zero_current = zero_current; sp.species_type = species_type;
  for (int i = 0; i < 6; ++i) sp.bc[i] = bcs[i];
  sp.sorted = false;
  return 0;
}

static int32_t check_species(eb200_ctx* c, int32_t isp, int32_t nvar) {
  if (isp < 0 || (size_t)isp >= c->species.size() || !c->species[isp].defined) return fail(c, EB200_E_ARG, "species not defined");
  if (nvar != c->ndims + 4) return fail(c, EB200_E_ARG, "nvar must be ndims+4 (pos, p(3), weight): optional per-particle fields are not on this path");
  return 0;
}
int32_t eb200_upload_particles(eb200_ctx* c, int32_t isp, int64_t n, const double* packed, int32_t nvar) {
  CHECK_CTX(c); NEED_GRID(c);
  EB_TRY(check_species(c, isp, nvar));
  if (n < 0 || (n > 0 && !packed)) return fail(c, EB200_E_ARG, "bad particle buffer");
  c->species[isp].deferred = false;      // the old particles are dropped
  return add_particles(c, isp, n, packed, nvar, true);
}
int32_t eb200_append_particles(eb200_ctx* c, int32_t isp, int64_t n, const double* packed, int32_t nvar) {
  CHECK_CTX(c); NEED_GRID(c);
  EB_TRY(check_species(c, isp, nvar));
  if (n < 0 || (n > 0 && !packed)) return fail(c, EB200_E_ARG, "bad particle buffer");
  EB_TRY(settle_species(c, c->species[isp]));
  return add_particles(c, isp, n, packed, nvar, false);
}
int32_t eb200_particle_count(eb200_ctx* c, int32_t isp, int64_t* n) {
  CHECK_CTX(c);
  if (!n || isp < 0 || (size_t)isp >= c->species.size()) return fail(c, EB200_E_ARG, "bad argument");
  *n = c->species[isp].n;
  return 0;
}
int32_t eb200_download_particles(eb200_ctx* c, int32_t isp, int64_t first, int64_t capacity, double* packed,
                                 int32_t nvar, int64_t* n_out) {
  CHECK_CTX(c);
  EB_TRY(check_species(c, isp, nvar));
  if (first < 0 || capacity < 0 || (capacity > 0 && !packed)) return fail(c, EB200_E_ARG, "bad particle buffer");
  long long got = 0;
  EB_TRY(settle_species(c, c->species[isp]));
  EB_TRY(get_particles(c, isp, first, capacity, packed, nvar, &got));
  if (n_out) *n_out = got;
  return 0;
}

int32_t eb200_upload_fields(eb200_ctx* c, const double* ex, const double* ey, const double* ez, const double* bx,
                            const double* by, const double* bz, const double* jx, const double* jy, const double* jz) {
  CHECK_CTX(c); NEED_GRID(c);
  const double* src[NFIELD] = {ex, ey, ez, bx, by, bz, jx, jy, jz};
  for (int f = 0; f < NFIELD; ++f)
    if (src[f]) {
      char* dev = (char*)c->f.f[f];
      EB_TRY(staged_h2d(c, src[f], (size_t)c->g.total * sizeof(double), sizeof(double),
                        [dev](size_t o, size_t) { return (void*)(dev + o); }, nullptr));
    }
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}
int32_t eb200_download_fields(eb200_ctx* c, double* ex, double* ey, double* ez, double* bx, double* by, double* bz,
                              double* jx, double* jy, double* jz) {
  CHECK_CTX(c); NEED_GRID(c);
  double* dst[NFIELD] = {ex, ey, ez, bx, by, bz, jx, jy, jz};
  for (int f = 0; f < NFIELD; ++f)
    if (dst[f]) {
      const char* dev = (const char*)c->f.f[f];
      EB_TRY(staged_d2h(c, dst[f], (size_t)c->g.total * sizeof(double), sizeof(double),
                        [dev](size_t o, size_t) { return (const void*)(dev + o); }));
    }
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int32_t eb200_get_load(eb200_ctx* c, const double grid_min_global[3], int64_t* load_x, int64_t* load_y, int64_t* load_z) {
  CHECK_CTX(c); NEED_GRID(c);
  if (!grid_min_global) return fail(c, EB200_E_ARG, "get_load: null grid_min_global");
  long long* const load[3] = {(long long*)load_x, (long long*)load_y, (long long*)load_z};
  EB_TRY(settle_particles(c));
  return get_load(c, grid_min_global, load);
}

int32_t eb200_redistribute(eb200_ctx* c, const int32_t* old_cell_min, const int32_t* old_cell_max,
                           const int32_t* new_cell_min, const int32_t* new_cell_max, const double grid_min_global[3],
                           const double grid_min_local[3], const double min_local[3], const double max_local[3]) {
  CHECK_CTX(c); NEED_GRID(c);
  if (!old_cell_min || !old_cell_max || !new_cell_min || !new_cell_max || !grid_min_global || !grid_min_local ||
      !min_local || !max_local)
    return fail(c, EB200_E_ARG, "redistribute: null argument");
  EB_TRY(settle_particles(c));
  return redistribute(c, old_cell_min, old_cell_max, new_cell_min, new_cell_max, grid_min_global, grid_min_local,
                      min_local, max_local);
}

int32_t eb200_calc_derived(eb200_ctx* c, int32_t kind, int32_t ispecies, int32_t direction, double* out) {
  CHECK_CTX(c); NEED_GRID(c);
  if (!out) return fail(c, EB200_E_ARG, "calc_derived: null output array");
  if (kind < 0 || kind > EB200_DF_AVERAGE_MOMENTUM) return fail(c, EB200_E_ARG, "calc_derived: unknown kind");
  if (ispecies >= (int32_t)c->species.size() || (ispecies >= 0 && !c->species[ispecies].defined))
    return fail(c, EB200_E_ARG, "calc_derived: species not defined");
  const int ad = direction < 0 ? -direction : direction;
  const bool need_dir = kind == EB200_DF_EKFLUX || kind == EB200_DF_POYNT_FLUX || kind == EB200_DF_SPECIES_CURRENT ||
                        kind == EB200_DF_AVERAGE_MOMENTUM;
  if (ad > 3 || (need_dir && ad == 0) || (direction < 0 && kind != EB200_DF_EKFLUX))
    return fail(c, EB200_E_ARG, "calc_derived: direction must be 1..3 (negative only for ekflux; 0 only where optional)");
  if (bc_mixed(c)) return fail(c, EB200_E_UNSUPPORTED, "calc_derived with mixed per-species particle boundaries");
  EB_TRY(settle_particles(c));
  return calc_derived(c, kind, ispecies, direction, out);
}

int32_t eb200_set_thermal_boundary(eb200_ctx* c, int32_t isp, int32_t boundary, const double* ext_temp, int64_t seed) {
  CHECK_CTX(c); NEED_GRID(c);
  if (isp < 0 || (size_t)isp >= c->species.size() || !c->species[isp].defined) return fail(c, EB200_E_ARG, "species not set");
  if (boundary < 1 || boundary > 2 * c->ndims || !ext_temp) return fail(c, EB200_E_ARG, "set_thermal_boundary: bad boundary / null table");
  SpeciesD& sp = c->species[isp];
  const int d = (boundary - 1) / 2;
  size_t plane = 1;
  for (int a = 0; a < c->ndims; ++a) if (a != d) plane *= (size_t)(c->g.n[a] + 2 * c->ng);
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (sp.ext_temp[boundary - 1]) { cudaFree(sp.ext_temp[boundary - 1]); sp.ext_temp[boundary - 1] = nullptr; }
  EB_CUDA(c, cudaMalloc(&sp.ext_temp[boundary - 1], 3 * plane * sizeof(double)));
  EB_CUDA(c, cudaMemcpy(sp.ext_temp[boundary - 1], ext_temp, 3 * plane * sizeof(double), cudaMemcpyHostToDevice));
  sp.thermal_seed = (unsigned long long)seed;
  return 0;
}
int32_t eb200_set_probe(eb200_ctx* c, int32_t isp, int32_t iprobe, const double* point, const double* normal, double ek_min,
                        double ek_max) {
  CHECK_CTX(c);
  if (isp < 0 || (size_t)isp >= c->species.size() || !c->species[isp].defined) return fail(c, EB200_E_ARG, "set_probe: species not defined");
  if (iprobe < 0 || iprobe >= MAX_PROBES || !point || !normal) return fail(c, EB200_E_ARG, "set_probe: iprobe 0..7, point and normal required");
  ProbeSet& ps = c->species[isp].probes;
  if (!ps.buf) {
    ps.cap = c->probe_capacity;
    EB_CUDA(c, cudaMalloc(&ps.buf, (size_t)MAX_PROBES * ps.cap * (c->ndims + 4) * sizeof(double)));
    EB_CUDA(c, cudaMalloc(&ps.count, MAX_PROBES * sizeof(unsigned long long)));
    EB_CUDA(c, cudaMemsetAsync(ps.count, 0, MAX_PROBES * sizeof(unsigned long long), c->stream));
  }
  ps.on[iprobe] = 1; ps.ek_min[iprobe] = ek_min; ps.ek_max[iprobe] = ek_max;
  for (int d = 0; d < 3; ++d) { ps.point[iprobe][d] = d < c->ndims ? point[d] : 0.0; ps.normal[iprobe][d] = d < c->ndims ? normal[d] : 0.0; }
  if (iprobe + 1 > ps.n) ps.n = iprobe + 1;
  EB_CUDA(c, cudaMemsetAsync(ps.count + iprobe, 0, sizeof(unsigned long long), c->stream));
  return 0;
}
int32_t eb200_download_probe(eb200_ctx* c, int32_t isp, int32_t iprobe, int64_t capacity, double* packed, int32_t nvar,
                             int64_t* n_out, int32_t clear) {
  CHECK_CTX(c);
  EB_TRY(check_species(c, isp, nvar));
  ProbeSet& ps = c->species[isp].probes;
  if (iprobe < 0 || iprobe >= MAX_PROBES || !ps.on[iprobe] || capacity < 0) return fail(c, EB200_E_ARG, "download_probe: probe not defined");
  unsigned long long n = 0;
  EB_CUDA(c, cudaMemcpyAsync(&n, ps.count + iprobe, sizeof(n), cudaMemcpyDeviceToHost, c->stream));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  const long long kept = std::min<long long>((long long)n, ps.cap), m = std::min<long long>(kept, capacity);
  if (packed && m > 0) {
    const char* src = (const char*)(ps.buf + (size_t)iprobe * ps.cap * nvar);
    EB_TRY(staged_d2h(c, packed, (size_t)m * nvar * sizeof(double), (size_t)nvar * sizeof(double),
                      [src](size_t o, size_t) { return (const void*)(src + o); }));
    EB_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  if (n_out) *n_out = (int64_t)kept;          // rows a caller can fetch: never more than the device list holds
  if (clear) EB_CUDA(c, cudaMemsetAsync(ps.count + iprobe, 0, sizeof(unsigned long long), c->stream));
  if ((long long)n > ps.cap)
    return fail(c, EB200_E_NOMEM, "probe list overflowed: " + std::to_string(n) + " particles sampled, " + std::to_string(kept) +
                                  " kept (raise option probe_capacity)");
  return 0;
}
int32_t eb200_dist_fn(eb200_ctx* c, int32_t ispecies, int32_t curdims, const int32_t direction[3], const double ranges[6],
                      const int32_t resolution[3], const double dgrid[3], const double* restrictions,
                      const int32_t* use_restrictions, double* out) {
  CHECK_CTX(c); NEED_GRID(c);
  if (ispecies < 0 || (size_t)ispecies >= c->species.size() || !c->species[ispecies].defined)
    return fail(c, EB200_E_ARG, "dist_fn: species not defined");
  if (curdims < 1 || curdims > 3 || !direction || !ranges || !resolution || !dgrid || !out)
    return fail(c, EB200_E_ARG, "dist_fn: bad argument");
  for (int i = 0; i < curdims; ++i) {
    if (direction[i] < 1 || direction[i] > c->ndims + 9) return fail(c, EB200_E_ARG, "dist_fn: unknown direction (dist_fn.F90:378-385)");
    if (resolution[i] < 1 || !(dgrid[i] > 0.0)) return fail(c, EB200_E_ARG, "dist_fn: resolution and dgrid must be positive");
  }
  if ((restrictions == nullptr) != (use_restrictions == nullptr)) return fail(c, EB200_E_ARG, "dist_fn: restrictions and use_restrictions go together");
  EB_TRY(settle_species(c, c->species[ispecies]));
  return dist_fn(c, ispecies, curdims, direction, ranges, resolution, dgrid, restrictions, use_restrictions, out);
}

int32_t eb200_set_laser(eb200_ctx* c, int32_t boundary, int32_t id, double amp, double pol_angle, double t_start,
                        double t_end, const double* profile, const double* phase) {
  CHECK_CTX(c); NEED_GRID(c);
  if (boundary < 1 || boundary > 2 * c->ndims) return fail(c, EB200_E_ARG, "boundary must be 1..2*ndims");
  const int b = boundary - 1, a = b / 2;
  size_t plane = 1;
  for (int d = 0; d < c->ndims; ++d) if (d != a) plane *= (size_t)(c->g.n[d] + 1);
  Laser* L = nullptr;
  for (Laser& l : c->lasers[b]) if (l.id == id) L = &l;
  if (!L) { c->lasers[b].emplace_back(); L = &c->lasers[b].back(); L->id = id; }
  L->amp = amp; L->pol = pol_angle; L->t_start = t_start; L->t_end = t_end; L->plane = plane;
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (L->profile) { cudaFree(L->profile); L->profile = nullptr; }
  if (L->phase) { cudaFree(L->phase); L->phase = nullptr; }
  if (profile) {
    EB_CUDA(c, cudaMalloc(&L->profile, plane * sizeof(double)));
    EB_CUDA(c, cudaMemcpy(L->profile, profile, plane * sizeof(double), cudaMemcpyHostToDevice));
  }
  if (phase) {
    EB_CUDA(c, cudaMalloc(&L->phase, plane * sizeof(double)));
    EB_CUDA(c, cudaMemcpy(L->phase, phase, plane * sizeof(double), cudaMemcpyHostToDevice));
  }
  return 0;
}
int32_t eb200_set_laser_step(eb200_ctx* c, int32_t boundary, int32_t id, double t_env, double integral_phase) {
  CHECK_CTX(c);
  if (boundary < 1 || boundary > 6) return fail(c, EB200_E_ARG, "bad boundary");
  for (Laser& l : c->lasers[boundary - 1])
    if (l.id == id) { l.t_env = t_env; l.integral_phase = integral_phase; return 0; }
  return fail(c, EB200_E_ARG, "laser id has not been set on this boundary");
}

int32_t eb200_update_eb_fields_half(eb200_ctx* c) {
  CHECK_CTX(c); NEED_GRID(c); NEED_DT(c);
  // fields.f90:819-847
  EB_TRY(update_e_field(c));
  EB_TRY(efield_bcs(c));
  EB_TRY(update_b_field(c));
  return bfield_bcs(c, true);
}
int32_t eb200_update_eb_fields_final(eb200_ctx* c) {
  CHECK_CTX(c); NEED_GRID(c); NEED_DT(c);
  // fields.f90:851-872
  EB_TRY(update_b_field(c));
  EB_TRY(bfield_final_bcs(c));
  EB_TRY(update_e_field(c));
  return efield_bcs(c);
}
int32_t eb200_push_particles(eb200_ctx* c) { CHECK_CTX(c); NEED_GRID(c); NEED_DT(c); return push_particles(c); }
int32_t eb200_current_finish(eb200_ctx* c) { CHECK_CTX(c); NEED_GRID(c); return current_finish(c); }
int32_t eb200_particle_bcs(eb200_ctx* c) { CHECK_CTX(c); NEED_GRID(c); NEED_DT(c); EB_TRY(settle_particles(c)); return particle_bcs(c); }
int32_t eb200_efield_bcs(eb200_ctx* c) { CHECK_CTX(c); NEED_GRID(c); return efield_bcs(c); }
int32_t eb200_bfield_final_bcs(eb200_ctx* c) { CHECK_CTX(c); NEED_GRID(c); NEED_DT(c); return bfield_final_bcs(c); }

int32_t eb200_energy(eb200_ctx* c, double* fe, double* per_species) { CHECK_CTX(c); NEED_GRID(c); EB_TRY(settle_particles(c)); return energy(c, fe, per_species); }
int32_t eb200_charge_residual(eb200_ctx* c, double* max_res, double* max_rho) {
  CHECK_CTX(c); NEED_GRID(c); NEED_DT(c);
  if (bc_mixed(c)) return fail(c, EB200_E_UNSUPPORTED, "charge_residual with mixed per-species particle boundaries");
  EB_TRY(settle_particles(c));
  return charge_residual(c, max_res, max_rho, false);
}

// no counterpart in the reference (see include/epoch_b200.h)
int32_t eb200_charge_residual_static(eb200_ctx* c, double* max_res, double* max_rho) {
  CHECK_CTX(c); NEED_GRID(c); NEED_DT(c);
  if (bc_mixed(c)) return fail(c, EB200_E_UNSUPPORTED, "charge_residual with mixed per-species particle boundaries");
  return charge_residual_static(c, max_res, max_rho);
}

int32_t eb200_set_option(eb200_ctx* c, const char* key, const char* value) {
  CHECK_CTX(c);
  if (!key || !value) return fail(c, EB200_E_ARG, "null");
  const std::string k(key), v(value);
  if (k == "capture_charge_reference") { NEED_GRID(c); NEED_DT(c); double a, b; EB_TRY(settle_particles(c)); return charge_residual(c, &a, &b, true); }
  if (k == "push_kernel") {
    if (v == "auto") c->push_kernel = 0; else if (v == "baseline") c->push_kernel = 1;
    else if (v == "tiled" || v == "split") c->push_kernel = 3;
    else return fail(c, EB200_E_ARG, "push_kernel: auto|baseline|tiled");
    return 0;
  }
  if (k == "fuse_keys") { c->fuse_keys = atoi(value) != 0; return 0; }
  if (k == "tma") { c->use_tma = atoi(value) != 0; c->tmap_valid = false; return 0; }
  // deferred permute: head room (slots) between the two states of the particle arrays; takes effect when
  // the arrays are next (re)allocated.  "auto" = need / 8, "0" = off (permute inside the sort)
  if (k == "defer_window") { c->defer_window = v == "auto" ? -1 : atoll(value); return 0; }
  if (k == "deterministic_deposit") { c->deterministic = atoi(value) != 0; return 0; }
  if (k == "timers") { c->timers_on = atoi(value) != 0; return 0; }
  if (k == "reserve") {   // "<species>:<particles>": allocate particle storage once, exactly (no regrow copies)
    const size_t colon = v.find(':');
    if (colon == std::string::npos) return fail(c, EB200_E_ARG, "reserve: <species>:<particles>");
    const int isp = atoi(v.substr(0, colon).c_str());
    const long long n = atoll(v.substr(colon + 1).c_str());
    if (isp < 0 || (size_t)isp >= c->species.size() || !c->species[isp].defined || n < 0) return fail(c, EB200_E_ARG, "reserve: bad species or count");
    return species_reserve(c, c->species[isp], n, true);
  }
  if (k == "push_overlap") { c->push_overlap = atoi(value) != 0; return 0; }
  if (k == "deposit_scratch") { c->rec_scratch_max = atoll(value) < 1024 ? 1024 : atoll(value); return 0; }
  if (k == "probe_capacity") { c->probe_capacity = std::max<long long>(atoll(value), 16); return 0; }
  if (k == "host_threads") { c->host_threads = atoi(value); return 0; }
  if (k == "pin_bytes") {      // chunk size of the host transfers (before the first transfer)
    if (c->pin[0]) return fail(c, EB200_E_STATE, "pin_bytes: the staging ring is already allocated");
    c->pin_bytes = (size_t)std::max<long long>(atoll(value), 1024);
    return 0;
  }
  if (k == "mig_capacity") {      // particles per migration mailbox to start with (the same on every rank)
    if (atoll(value) < 1) return fail(c, EB200_E_ARG, "mig_capacity must be >= 1");
    c->mig_min_particles = atoll(value); return 0;
  }
  return fail(c, EB200_E_ARG, "unknown option " + k);
}

int32_t eb200_sync(eb200_ctx* c) {
  CHECK_CTX(c);
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int32_t eb200_timer(eb200_ctx* c, const char* phase, int32_t reset, double* ms, int64_t* launches) {
  CHECK_CTX(c);
  if (!phase) return fail(c, EB200_E_ARG, "null");
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  const std::string k(phase);
  if (k == "launches") {   // total kernels launched by this context
    if (ms) *ms = 0;
    if (launches) *launches = c->launches;
    if (reset) c->launches = 0;
    return 0;
  }
  if (k == "deferred_sorts" || k == "permuted_sorts") {   // sorts whose permute was left to the next push / done in the sort
    if (ms) *ms = 0;
    if (launches) *launches = k == "deferred_sorts" ? c->n_deferred : c->n_permuted;
    if (reset) { c->n_deferred = 0; c->n_permuted = 0; }
    return 0;
  }
  if (k == "cell_mismatch" || k == "thermal_reemitted") {
    // particles whose sort key and push disagreed on the cell (must stay 0) / particles re-emitted by thermal walls so far
    unsigned long long v = 0;
    EB_CUDA(c, cudaStreamSynchronize(c->stream));
    EB_CUDA(c, cudaMemcpy(&v, c->d_counters + (k == "cell_mismatch" ? 8 : 9), sizeof(v), cudaMemcpyDeviceToHost));
    if (ms) *ms = 0;
    if (launches) *launches = (int64_t)v;
    return 0;
  }
  PhaseTimer& t = c->timers[k];
  for (auto& ev : t.ev) {
    float f = 0;
    cudaEventElapsedTime(&f, ev.first, ev.second);
    t.ms += f;
    cudaEventDestroy(ev.first); cudaEventDestroy(ev.second);
  }
  t.ev.clear();
  if (ms) *ms = t.ms;
  if (launches) *launches = t.launches;
  if (reset) { t.ms = 0; t.launches = 0; }
  return 0;
}

int32_t eb200_mark(eb200_ctx* c, const char* phase, int32_t begin) {
  CHECK_CTX(c);
  if (!phase) return fail(c, EB200_E_ARG, "null");
  PhaseTimer& t = c->timers[phase];
  cudaEvent_t e = nullptr;
  EB_CUDA(c, cudaEventCreate(&e));
  EB_CUDA(c, cudaEventRecord(e, c->stream));
  if (begin) {
    if (c->open_marks.count(phase)) { cudaEventDestroy(c->open_marks[phase].first); }
    c->open_marks[phase] = std::make_pair(e, c->launches);
  } else {
    auto it = c->open_marks.find(phase);
    if (it == c->open_marks.end()) { cudaEventDestroy(e); return fail(c, EB200_E_STATE, "mark(end) without mark(begin)"); }
    t.ev.emplace_back(it->second.first, e);
    t.launches += c->launches - it->second.second;
    c->open_marks.erase(it);
  }
  return 0;
}

const char* eb200_last_error(eb200_ctx* c) { return c ? c->err.c_str() : g_last_error.c_str(); }
const char* eb200_version(void) { return "epoch_b200 0.1 (sm_100a)"; }

}  // extern "C"
