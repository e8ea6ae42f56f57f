// oracle/orc_api.cpp -- CPU ORACLE (test infrastructure only): the same C ABI as
// include/epoch_b200.h with prefix orc_, so the parity tests drive oracle and
// product through one harness.  Single rank only (nproc must be 1 per context;
// multi-rank tests exchange halos themselves through orc_field_ptr).
#define EB200_NAME(x) orc_##x
#define EB200_CTX orc_ctx
#include <cstdlib>
#include "../include/epoch_b200.h"
#include "orc_state.hpp"
#include <algorithm>
#include <chrono>
#include <map>
#include <new>

struct orc_ctx { orc::State s; };
using namespace orc;

static thread_local std::string g_err;
static int32_t fail(orc_ctx* c, int32_t code, const std::string& m) {
  if (c) c->s.err = m;
  g_err = m;
  return code;
}

extern "C" {

int32_t orc_create(orc_ctx** out, int32_t ndims, int32_t shape, int32_t, int32_t rank,
                   int32_t nproc, const int32_t nprocdir[3], const int32_t coords[3], const void*) {
  if (!out || ndims < 1 || ndims > 3 || shape < 0 || shape > 2) return fail(nullptr, EB200_E_ARG, "bad ndims/shape");
  orc_ctx* c = new (std::nothrow) orc_ctx();
  if (!c) return EB200_E_NOMEM;
  State& s = c->s;
  s.ndims = ndims; s.shape = shape; s.rank = rank; s.nproc = nproc;
  for (int d = 0; d < 3; ++d) {
    s.nprocdir[d] = nprocdir ? nprocdir[d] : 1;
    s.coords[d] = coords ? coords[d] : 0;
    s.min_bnd[d] = s.coords[d] == 0;
    s.max_bnd[d] = s.coords[d] == s.nprocdir[d] - 1;
  }
  // constants.F90:472-492
  if (shape == 2) { s.sf_min = -2; s.sf_max = 2; s.png = 4; }
  else if (shape == 0) { s.sf_min = 0; s.sf_max = 1; s.png = 2; }
  else { s.sf_min = -1; s.sf_max = 1; s.png = 3; }
  s.ng = s.png + 2;
  *out = c;
  return 0;
}
int32_t orc_host_alloc(void** out, int64_t bytes) {
  if (!out || bytes < 0) return EB200_E_ARG;
  *out = bytes ? std::malloc((size_t)bytes) : nullptr;
  return (bytes && !*out) ? EB200_E_NOMEM : 0;
}
int32_t orc_host_free(void* p) { std::free(p); return 0; }
int32_t orc_nccl_unique_id(void* out128) { if (out128) std::memset(out128, 0, 128); return 0; }
int32_t orc_destroy(orc_ctx* c) { delete c; return 0; }

int32_t orc_set_grid(orc_ctx* c, const int32_t n_local[3], const int32_t n_global[3],
                     const int32_t n_global_min[3], const double d[3], const double grid_min_local[3],
                     const double min_local[3], const double max_local[3], const double min_global[3],
                     const double max_global[3], int32_t cpml_thickness) {
  if (!c) return EB200_E_ARG;
  State& s = c->s;
  if (cpml_thickness < 0) return fail(c, EB200_E_ARG, "cpml_thickness < 0");
  int64_t stride = 1;
  for (int k = 0; k < 3; ++k) {
    const bool ex = k < s.ndims;
    s.n[k] = ex ? n_local[k] : 1;
    s.nglob[k] = ex ? n_global[k] : 1;
    s.nglob_min[k] = ex ? n_global_min[k] : 1;
    s.ngd[k] = ex ? s.ng : 0;
    s.d[k] = ex ? d[k] : 1.0;
    s.grid_min_local[k] = ex ? grid_min_local[k] : 0.0;
    s.min_local[k] = ex ? min_local[k] : 0.0;
    s.max_local[k] = ex ? max_local[k] : 0.0;
    s.gmin[k] = ex ? min_global[k] : 0.0;
    s.gmax[k] = ex ? max_global[k] : 0.0;
    if (s.n[k] < 1) return fail(c, EB200_E_ARG, "n_local < 1");
    s.st[k] = stride;
    stride *= (int64_t)(s.n[k] + 2 * s.ngd[k]);
  }
  s.total = stride;
  s.cpml_thickness = cpml_thickness;
  for (int f = 0; f < NFIELD; ++f) s.f[f].assign((size_t)s.total, 0.0);
  s.grid_set = true;
  return 0;
}

int32_t orc_set_time(orc_ctx* c, double dt, double time, int32_t step) {
  if (!c) return EB200_E_ARG;
  c->s.dt = dt; c->s.time = time; c->s.step = step;
  return 0;
}

int32_t orc_set_field_bcs(orc_ctx* c, const int32_t bc_field[6]) {
  if (!c) return EB200_E_ARG;
  State& s = c->s;
  // setup_boundaries, boundary.F90:30-59
  for (int i = 0; i < 6; ++i) {
    int bc = i < 2 * s.ndims ? bc_field[i] : BC_PERIODIC;
    s.add_laser[i] = false;
    if (bc == BC_OTHER) bc = BC_CLAMP;
    if (bc == BC_SIMPLE_LASER) s.add_laser[i] = true;       // cpml_laser: set by set_cpml
    if (bc == BC_REFLECT) bc = BC_CLAMP;
    if (bc == BC_OPEN) bc = BC_SIMPLE_OUTFLOW;
    s.bc_field[i] = bc;
  }
  s.cpml_on = false;
  for (int i = 0; i < 2 * s.ndims; ++i)
    if (s.bc_field[i] == BC_CPML_LASER || s.bc_field[i] == BC_CPML_OUTFLOW) s.cpml_on = true;
  return 0;
}

int32_t orc_set_cpml(orc_ctx* c, int32_t axis, const double* kappa_e, const double* kappa_b, const double* sigma_e,
                     const double* sigma_b, const double* a_e, const double* a_b, int32_t min_start, int32_t min_end,
                     int32_t max_start, int32_t max_end, int32_t laser_idx_min, int32_t laser_idx_max) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  State& s = c->s;
  if (axis < 0 || axis >= s.ndims) return fail(c, EB200_E_ARG, "axis");
  if (!kappa_e || !kappa_b || !sigma_e || !sigma_b || !a_e || !a_b) return fail(c, EB200_E_ARG, "null profile");
  const size_t len = (size_t)(s.n[axis] + 2 * s.ng);
  const double* src[6] = {kappa_e, kappa_b, sigma_e, sigma_b, a_e, a_b};
  for (int k = 0; k < 6; ++k) s.cpml_prof[axis][k].assign(src[k], src[k] + len);
  s.cpml_start[2 * axis] = min_start; s.cpml_end[2 * axis] = min_end;
  s.cpml_start[2 * axis + 1] = max_start; s.cpml_end[2 * axis + 1] = max_end;
  for (int side = 0; side < 2; ++side) {
    const int b = 2 * axis + side, idx = side ? laser_idx_max : laser_idx_min;
    if (s.bc_field[b] == BC_CPML_LASER) { s.add_laser[b] = idx >= 0; s.cpml_laser_idx[b] = idx >= 0 ? idx : 1; }
  }
  for (int k = 0; k < 4; ++k) if (s.cpml_psi[axis][k].empty()) s.cpml_psi[axis][k].assign((size_t)s.total, 0.0);
  return 0;
}

int32_t orc_set_window_fields(orc_ctx* c, const double* ex, const double* ey, const double* ez, const double* bx,
                              const double* by, const double* bz) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  const double* pl[6] = {ex, ey, ez, bx, by, bz};
  const size_t cnt = (size_t)(c->s.total / (c->s.n[0] + 2 * c->s.ngd[0]));
  for (int k = 0; k < 6; ++k) {
    c->s.win_planes[k].assign(cnt, 0.0);
    if (pl[k]) std::copy(pl[k], pl[k] + cnt, c->s.win_planes[k].begin());
  }
  return 0;
}
int32_t orc_setup_field_boundaries(orc_ctx* c) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  orc::setup_field_boundaries(c->s);
  return 0;
}
int32_t orc_set_boundary_fields(orc_ctx* c, int32_t boundary, const double* ex, const double* ey, const double* ez,
                                const double* bx, const double* by, const double* bz) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  if (boundary < 1 || boundary > 2 * c->s.ndims) return fail(c, EB200_E_ARG, "boundary out of range");
  const int b = boundary - 1, a = b / 2;
  const double* pl[6] = {ex, ey, ez, bx, by, bz};
  const size_t cnt = (size_t)(c->s.total / (c->s.n[a] + 2 * c->s.ngd[a]));
  for (int k = 0; k < 6; ++k) {
    c->s.bnd_planes[b][k].assign(cnt, 0.0);
    if (pl[k]) std::copy(pl[k], pl[k] + cnt, c->s.bnd_planes[b][k].begin());
  }
  return 0;
}
int32_t orc_shift_window(orc_ctx* c, double x_grid_min_local, double x_min_local, double x_max_local, double x_min,
                         double x_max) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  if (c->s.n[0] < 3) return fail(c, EB200_E_ARG, "moving window needs nx >= 3");
  orc::shift_window(c->s, x_grid_min_local, x_min_local, x_max_local, x_min, x_max);
  return 0;
}

int32_t orc_set_current_options(orc_ctx* c, int32_t smooth, int32_t its, int32_t comp_its, int32_t nstrides,
                                const int32_t* strides, int32_t use_corr, const double initial_j[3]) {
  if (!c) return EB200_E_ARG;
  if (its < 0 || comp_its < 0 || nstrides < 0 || (nstrides > 0 && !strides)) return fail(c, EB200_E_ARG, "bad smoothing parameters");
  for (int k = 0; k < nstrides; ++k)
    if (strides[k] < 1 || strides[k] > c->s.ng) return fail(c, EB200_E_ARG, "smooth_strides must lie in 1..ng");
  if (use_corr && !initial_j) return fail(c, EB200_E_ARG, "current correction needs initial_j");
  c->s.smooth = smooth != 0; c->s.smooth_its = its; c->s.smooth_comp_its = comp_its;
  c->s.smooth_strides.assign(strides, strides + nstrides);
  c->s.use_current_correction = use_corr != 0;
  for (int k = 0; k < 3; ++k) c->s.initial_j[k] = use_corr ? initial_j[k] : 0.0;
  return 0;
}

int32_t orc_set_solver(orc_ctx* c, int32_t field_order, int32_t maxwell_solver, const double coeffs[15]) {
  if (!c) return EB200_E_ARG;
  if (field_order != 2 && field_order != 4 && field_order != 6) return fail(c, EB200_E_ARG, "field_order");
  if (maxwell_solver < -1 || maxwell_solver > 6 || maxwell_solver == 1) return fail(c, EB200_E_ARG, "unknown maxwell_solver");
  if (maxwell_solver != 0 && !coeffs) return fail(c, EB200_E_ARG, "an extended-stencil solver needs its 15 coefficients");
  c->s.field_order = field_order; c->s.maxwell_solver = maxwell_solver;
  if (coeffs) std::copy(coeffs, coeffs + 15, c->s.coeffs);
  return 0;
}

int32_t orc_set_species(orc_ctx* c, int32_t isp, double charge, double mass, double weight,
                        int32_t immobile, int32_t zero_current, int32_t species_type,
                        const int32_t bc_particle[6]) {
  if (!c || isp < 0 || isp > 1023) return EB200_E_ARG;
  State& s = c->s;
  if ((size_t)isp >= s.species.size()) s.species.resize(isp + 1);
  Species& sp = s.species[isp];
  sp.defined = true; sp.charge = charge; sp.mass = mass; sp.weight = weight;
  sp.immobile = immobile; sp.zero_current = zero_current; sp.species_type = species_type;
  for (int i = 0; i < 6; ++i) {
    int bc = (bc_particle && i < 2 * s.ndims) ? bc_particle[i] : BC_PERIODIC;
    // setup_particle_boundary, boundary.F90:98-140
    if (bc == BC_OTHER || bc == BC_CONDUCT) bc = BC_REFLECT;
    if (bc == BC_SIMPLE_LASER || bc == BC_SIMPLE_OUTFLOW || bc == BC_CPML_LASER || bc == BC_CPML_OUTFLOW) bc = BC_OPEN;
    if (bc == BC_HEAT_BATH) bc = BC_OPEN;      // setup_domain_dependent_boundaries, boundary.F90:75-93 (the injector stays on the host)
    if (bc != BC_PERIODIC && bc != BC_REFLECT && bc != BC_OPEN && bc != BC_THERMAL) return fail(c, EB200_E_ARG, "bad particle BC");
    sp.bc[i] = bc;
  }
  return 0;
}

static int32_t add_particles(orc_ctx* c, int32_t isp, int64_t n, const double* packed, int32_t nvar, bool replace) {
  if (!c || isp < 0 || (size_t)isp >= c->s.species.size() || !c->s.species[isp].defined) return fail(c, EB200_E_ARG, "species not set");
  State& s = c->s;
  if (nvar != s.ndims + 4) return fail(c, EB200_E_ARG, "nvar must be ndims+4");
  if (n < 0 || (n > 0 && !packed)) return EB200_E_ARG;
  Species& sp = s.species[isp];
  if (replace) { for (int d = 0; d < 3; ++d) { sp.pos[d].clear(); sp.p[d].clear(); } sp.w.clear(); }
  for (int64_t i = 0; i < n; ++i) {
    const double* q = packed + i * nvar;
    for (int d = 0; d < s.ndims; ++d) sp.pos[d].push_back(q[d]);
    for (int d = 0; d < 3; ++d) sp.p[d].push_back(q[s.ndims + d]);
    sp.w.push_back(q[s.ndims + 3]);
  }
  return 0;
}
int32_t orc_upload_particles(orc_ctx* c, int32_t isp, int64_t n, const double* packed, int32_t nvar) {
  return add_particles(c, isp, n, packed, nvar, true);
}
int32_t orc_append_particles(orc_ctx* c, int32_t isp, int64_t n, const double* packed, int32_t nvar) {
  return add_particles(c, isp, n, packed, nvar, false);
}
int32_t orc_particle_count(orc_ctx* c, int32_t isp, int64_t* n) {
  if (!c || !n || isp < 0 || (size_t)isp >= c->s.species.size()) return EB200_E_ARG;
  *n = (int64_t)c->s.species[isp].count();
  return 0;
}
int32_t orc_download_particles(orc_ctx* c, int32_t isp, int64_t first, int64_t capacity, double* packed,
                               int32_t nvar, int64_t* n_out) {
  if (!c || isp < 0 || (size_t)isp >= c->s.species.size()) return EB200_E_ARG;
  State& s = c->s;
  if (nvar != s.ndims + 4) return fail(c, EB200_E_ARG, "nvar must be ndims+4");
  const Species& sp = s.species[isp];
  int64_t n = std::max<int64_t>(0, std::min<int64_t>((int64_t)sp.count() - first, capacity));
  for (int64_t i = 0; i < n; ++i) {
    double* q = packed + i * nvar;
    for (int d = 0; d < s.ndims; ++d) q[d] = sp.pos[d][first + i];
    for (int d = 0; d < 3; ++d) q[s.ndims + d] = sp.p[d][first + i];
    q[s.ndims + 3] = sp.w[first + i];
  }
  if (n_out) *n_out = n;
  return 0;
}

int32_t orc_upload_fields(orc_ctx* c, const double* ex, const double* ey, const double* ez, const double* bx,
                          const double* by, const double* bz, const double* jx, const double* jy, const double* jz) {
  if (!c || !c->s.grid_set) return fail(c, EB200_E_STATE, "grid not set");
  const double* src[NFIELD] = {ex, ey, ez, bx, by, bz, jx, jy, jz};
  for (int f = 0; f < NFIELD; ++f)
    if (src[f]) std::copy(src[f], src[f] + c->s.total, c->s.f[f].begin());
  return 0;
}
int32_t orc_download_fields(orc_ctx* c, double* ex, double* ey, double* ez, double* bx, double* by, double* bz,
                            double* jx, double* jy, double* jz) {
  if (!c || !c->s.grid_set) return fail(c, EB200_E_STATE, "grid not set");
  double* dst[NFIELD] = {ex, ey, ez, bx, by, bz, jx, jy, jz};
  for (int f = 0; f < NFIELD; ++f)
    if (dst[f]) std::copy(c->s.f[f].begin(), c->s.f[f].end(), dst[f]);
  return 0;
}
// oracle-only: direct pointer to a field array (multi-rank CPU tests)
double* orc_field_ptr(orc_ctx* c, int32_t comp) {
  if (!c || comp < 0 || comp >= NFIELD) return nullptr;
  return c->s.f[comp].data();
}

static size_t plane_size(const State& s, int boundary) {
  const int a = boundary / 2;
  size_t n = 1;
  for (int d = 0; d < s.ndims; ++d) if (d != a) n *= (size_t)(s.n[d] + 1);
  return n;
}
int32_t orc_set_laser(orc_ctx* c, int32_t boundary, int32_t id, double amp, double pol, double t_start,
                      double t_end, const double* profile, const double* phase) {
  if (!c || !c->s.grid_set) return fail(c, EB200_E_STATE, "grid not set");
  State& s = c->s;
  if (boundary < 1 || boundary > 2 * s.ndims) return fail(c, EB200_E_ARG, "bad boundary");
  const int b = boundary - 1;
  Laser* L = nullptr;
  for (Laser& l : s.lasers[b]) if (l.id == id) L = &l;
  if (!L) { s.lasers[b].emplace_back(); L = &s.lasers[b].back(); L->id = id; }
  L->amp = amp; L->pol = pol; L->t_start = t_start; L->t_end = t_end;
  const size_t n = plane_size(s, b);
  if (profile) L->profile.assign(profile, profile + n); else L->profile.clear();
  if (phase) L->phase.assign(phase, phase + n); else L->phase.clear();
  return 0;
}
int32_t orc_set_laser_step(orc_ctx* c, int32_t boundary, int32_t id, double t_env, double integral_phase) {
  if (!c || boundary < 1 || boundary > 6) return EB200_E_ARG;
  for (Laser& l : c->s.lasers[boundary - 1])
    if (l.id == id) { l.t_env = t_env; l.integral_phase = integral_phase; return 0; }
  return fail(c, EB200_E_ARG, "laser id not set");
}

static bool thermal_tables_missing(const State& s) {
  for (const Species& sp : s.species)
    for (int b = 0; sp.defined && b < 2 * s.ndims; ++b)
      if (sp.bc[b] == BC_THERMAL && sp.ext_temp[b].empty()) return true;
  return false;
}
#define HOT(name, call)                                                         \
  int32_t orc_##name(orc_ctx* c) {                                              \
    if (!c || !c->s.grid_set) return fail(c, EB200_E_STATE, "grid not set");    \
    if (c->s.dt <= 0) return fail(c, EB200_E_STATE, "dt not set");              \
    if (thermal_tables_missing(c->s)) return fail(c, EB200_E_STATE, "thermal boundary without set_thermal_boundary"); \
    call(c->s);                                                                 \
    return 0;                                                                   \
  }
HOT(update_eb_fields_half, update_eb_fields_half)
HOT(push_particles, push_particles)
HOT(current_finish, current_finish)
HOT(update_eb_fields_final, update_eb_fields_final)
HOT(particle_bcs, particle_bcs)
HOT(efield_bcs, efield_bcs)
HOT(bfield_final_bcs, bfield_final_bcs)

int32_t orc_energy(orc_ctx* c, double* fe, double* per_species) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  energy(c->s, fe, per_species);
  return 0;
}
int32_t orc_charge_residual(orc_ctx* c, double* max_res, double* max_rho) {
  if (c && bc_mixed(c->s)) return fail(c, EB200_E_UNSUPPORTED, "charge_residual with mixed per-species particle BCs");
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  charge_residual(c->s, max_res, max_rho, false);
  return 0;
}
// (the CUDA library's no-allocation variant has nothing to vary on the CPU: the same routine under the second name,
//  so that one binding table serves both libraries)
int32_t orc_charge_residual_static(orc_ctx* c, double* max_res, double* max_rho) { return orc_charge_residual(c, max_res, max_rho); }
// part_load_func (balance.F90:2468-2495) for this (single) rank
int32_t orc_get_load(orc_ctx* c, const double grid_min_global[3], int64_t* load_x, int64_t* load_y, int64_t* load_z) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  if (!grid_min_global) return EB200_E_ARG;
  State& s = c->s;
  int64_t* load[3] = {load_x, load_y, load_z};
  for (int a = 0; a < s.ndims; ++a)
    if (load[a]) std::fill(load[a], load[a] + s.nglob[a] + 2 * s.ng, (int64_t)0);
  for (const Species& sp : s.species) {
    if (!sp.defined) continue;
    for (size_t ip = 0; ip < sp.count(); ++ip)
      for (int a = 0; a < s.ndims; ++a) {
        if (!load[a]) continue;
        const double r = (sp.pos[a][ip] - grid_min_global[a]) / s.d[a];
        const int cell = s.shape == 0 ? (int)std::floor(r) + 1 : (int)std::floor(r + 1.5);
        const int pos = std::min(std::max(cell + s.ng - 1, 0), s.nglob[a] + 2 * s.ng - 1);
        load[a][pos] += 1;
      }
  }
  return 0;
}
// the oracle is a single rank: a redistribution can only map the whole grid onto itself
int32_t orc_redistribute(orc_ctx* c, const int32_t* old_cell_min, const int32_t* old_cell_max, const int32_t* new_cell_min,
                         const int32_t* new_cell_max, const double*, const double*, const double*, const double*) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  if (!old_cell_min || !old_cell_max || !new_cell_min || !new_cell_max) return EB200_E_ARG;
  for (int a = 0; a < c->s.ndims; ++a)
    if (new_cell_min[a] != 1 || new_cell_max[a] != c->s.nglob[a] || old_cell_min[a] != 1 || old_cell_max[a] != c->s.nglob[a])
      return EB200_E_ARG;
  return 0;
}
int32_t orc_calc_derived(orc_ctx* c, int32_t kind, int32_t ispecies, int32_t direction, double* out) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  if (bc_mixed(c->s)) return fail(c, EB200_E_UNSUPPORTED, "calc_derived with mixed per-species particle BCs");
  if (!out || kind < 0 || kind > 10 || ispecies >= (int32_t)c->s.species.size()) return EB200_E_ARG;
  const int ad = direction < 0 ? -direction : direction;
  const bool need_dir = kind == 2 || kind == 3 || kind == 9 || kind == 10;
  if (ad > 3 || (need_dir && ad == 0) || (direction < 0 && kind != 2)) return EB200_E_ARG;
  calc_derived(c->s, kind, ispecies, direction, out);
  return 0;
}
int32_t orc_set_thermal_boundary(orc_ctx* c, int32_t isp, int32_t boundary, const double* ext_temp, int64_t seed) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  State& s = c->s;
  if (isp < 0 || isp >= (int32_t)s.species.size() || !s.species[isp].defined || !ext_temp) return EB200_E_ARG;
  if (boundary < 1 || boundary > 2 * s.ndims) return EB200_E_ARG;
  const int d = (boundary - 1) / 2;
  size_t plane = 1;
  for (int a = 0; a < s.ndims; ++a) if (a != d) plane *= (size_t)(s.n[a] + 2 * s.ng);
  s.species[isp].ext_temp[boundary - 1].assign(ext_temp, ext_temp + 3 * plane);
  (void)seed;      // the oracle draws from the reference's own sequential generator (orc_random_init)
  return 0;
}
int32_t orc_set_probe(orc_ctx* c, int32_t isp, int32_t iprobe, const double* point, const double* normal, double ek_min,
                      double ek_max) {
  if (!c || isp < 0 || isp >= (int32_t)c->s.species.size() || iprobe < 0 || iprobe > 7 || !point || !normal) return EB200_E_ARG;
  orc::Species::Probe& pr = c->s.species[isp].probes[iprobe];
  pr.on = true; pr.ek_min = ek_min; pr.ek_max = ek_max; pr.sampled.clear();
  for (int d = 0; d < c->s.ndims; ++d) { pr.point[d] = point[d]; pr.normal[d] = normal[d]; }
  return 0;
}
int32_t orc_download_probe(orc_ctx* c, int32_t isp, int32_t iprobe, int64_t capacity, double* packed, int32_t nvar,
                           int64_t* n_out, int32_t clear) {
  if (!c || isp < 0 || isp >= (int32_t)c->s.species.size() || iprobe < 0 || iprobe > 7 || nvar != c->s.ndims + 4) return EB200_E_ARG;
  orc::Species::Probe& pr = c->s.species[isp].probes[iprobe];
  if (!pr.on) return EB200_E_ARG;
  const int64_t n = (int64_t)(pr.sampled.size() / nvar);
  if (packed) std::copy(pr.sampled.begin(), pr.sampled.begin() + (size_t)std::min(n, capacity) * nvar, packed);
  if (n_out) *n_out = n;
  if (clear) pr.sampled.clear();
  return 0;
}
int32_t orc_dist_fn(orc_ctx* c, int32_t ispecies, int32_t curdims, const int32_t direction[3], const double ranges[6],
                    const int32_t resolution[3], const double dgrid[3], const double* restrictions,
                    const int32_t* use_restrictions, double* out) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  if (ispecies < 0 || ispecies >= (int32_t)c->s.species.size() || curdims < 1 || curdims > 3 || !direction || !ranges ||
      !resolution || !dgrid || !out || ((restrictions == nullptr) != (use_restrictions == nullptr)))
    return EB200_E_ARG;
  for (int i = 0; i < curdims; ++i)
    if (direction[i] < 1 || direction[i] > c->s.ndims + 9 || resolution[i] < 1 || !(dgrid[i] > 0.0)) return EB200_E_ARG;
  dist_fn(c->s, ispecies, curdims, direction, ranges, resolution, dgrid, restrictions, use_restrictions, out);
  return 0;
}
int32_t orc_set_option(orc_ctx* c, const char* key, const char*) {
  if (c && key && std::string(key) == "capture_charge_reference") {
    double a, b;
    charge_residual(c->s, &a, &b, true);
  }
  return 0;
}
int32_t orc_sync(orc_ctx*) { return 0; }
// wall-clock brackets (housekeeping/timer.f90:27-113)
static std::map<std::string, double> g_t_open, g_t_acc;
static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
int32_t orc_mark(orc_ctx*, const char* phase, int32_t begin) {
  if (!phase) return EB200_E_ARG;
  if (begin) g_t_open[phase] = now_ms();
  else if (g_t_open.count(phase)) { g_t_acc[phase] += now_ms() - g_t_open[phase]; g_t_open.erase(phase); }
  return 0;
}
int32_t orc_timer(orc_ctx*, const char* phase, int32_t reset, double* ms, int64_t* launches) {
  const std::string k(phase ? phase : "");
  if (ms) *ms = g_t_acc.count(k) ? g_t_acc[k] : 0.0;
  if (launches) *launches = 0;
  if (reset) g_t_acc.erase(k);
  return 0;
}
const char* orc_last_error(orc_ctx* c) { return c ? c->s.err.c_str() : g_err.c_str(); }
const char* orc_version(void) { return "epoch-oracle-cpu 0.1 (test infrastructure)"; }

}  // extern "C"
