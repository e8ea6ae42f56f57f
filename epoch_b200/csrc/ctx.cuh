// ctx.cuh -- context, device parameter blocks and helpers shared by all kernels.
// B200-native (sm_100a) implementation of EPOCH's PIC hot path; see DESIGN.md.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include <map>
#include <functional>
#include "../../include/epoch_b200.h"

namespace eb {

// constants.F90:144-151 (CODATA-2010 era literals, as the reference)
constexpr double PI = 3.141592653589793238462643383279503;
constexpr double C0 = 2.99792458e8;
constexpr double EPS0 = 8.854187817620389850536563031710750e-12;
constexpr double KB = 1.3806488e-23;          // constants.F90 (kb)

enum { EX = 0, EY, EZ, BX, BY, BZ, JX, JY, JZ, NFIELD };
enum { SHAPE_TOPHAT = 0, SHAPE_TRIANGLE = 1, SHAPE_BSPLINE3 = 2 };

// Grid description passed by value to kernels.  Arrays are Fortran shaped with
// ghosts, x fastest; dims that do not exist have n=1, ngd=0.
struct GridP {
  int ndims, shape, ng;
  int n[3];
  int ngd[3];
  long long st[3];
  long long total;
  double d[3], id[3];
  double gml[3];          // x_grid_min_local (centre of local cell 1)
  double min_local[3], max_local[3];
  double gmin[3], gmax[3];
  // extended cell grid used for sort keys: cells (1-KM .. n+KM) per existing dim,
  // cut into tiles of tdim cells; a key is tile-major: tile * cpt + cell-in-tile
  // (x fastest in both), so that the particles of a tile are contiguous
  int kn[3];              // extended extents
  int tdim[3], ntile[3], cpt;
  long long ncell_ext;    // number of key buckets = ntile[0]*ntile[1]*ntile[2]*cpt
  __host__ __device__ inline long long idx(int i, int j, int k) const {
    return (long long)(i + ngd[0] - 1) * st[0] + (long long)(j + ngd[1] - 1) * st[1] +
           (long long)(k + ngd[2] - 1) * st[2];
  }
};
constexpr int KM = 3;                    // key-grid margin in cells
// tile shape of the cell sort / tiled push kernel per dimensionality
#ifndef EB_T3X
#define EB_T3X 8
#define EB_T3Y 8
#define EB_T3Z 4
#endif
__host__ __device__ constexpr int tile_dim(int nd, int d) {
  return nd == 3 ? (d == 0 ? EB_T3X : d == 1 ? EB_T3Y : EB_T3Z) : nd == 2 ? (d == 2 ? 1 : 16) : (d == 0 ? 256 : 1);
}
constexpr uint32_t KEY_INVALID = 0xFFFFFFFFu;

struct FieldP { double* f[NFIELD]; };

// per-species device storage: structure of arrays, fp64
// particle probes of a species (shared_data.F90:392-404), passed to the general push kernel by value
constexpr int MAX_PROBES = 8;
struct ProbeSet {
  int n = 0;                  // highest defined probe + 1
  int on[MAX_PROBES] = {0, 0, 0, 0, 0, 0, 0, 0};
  double point[MAX_PROBES][3], normal[MAX_PROBES][3], ek_min[MAX_PROBES], ek_max[MAX_PROBES];
  double* buf = nullptr;                  // [MAX_PROBES][cap][nvar] sampled particles (wire format)
  unsigned long long* count = nullptr;    // [MAX_PROBES]
  long long cap = 0;
};

struct SpeciesD {
  bool defined = false;
  ProbeSet probes;
  double charge = 0, mass = 0, weight = 0;
  int immobile = 0, zero_current = 0, species_type = 0;
  int bc[6] = {1, 1, 1, 1, 1, 1};
  long long n = 0;          // live particles (valid after the last compaction)
  long long cap = 0;        // allocated slots per attribute
  double* a[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // x,y,z,px,py,pz,w: VIEW = base + off
  uint32_t* key = nullptr;  // sort key per slot (view = key_base + off)
  uint32_t* dest = nullptr; // destination index per slot; while `deferred`: src[i] = old slot of sorted index i
  // Deferred permute (particles.cu): the allocations are `cap` = need + win + 2 gap slots long and the
  // live slots start at `off` (low state: gap, high state: gap + win).  After a sort with `deferred`
  // set the arrays still hold the PRE-sort order; the next push gathers through src and writes the
  // sorted order into the other state, so the sort's permute passes never run.
  double* base[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  uint32_t* key_base = nullptr;
  long long off = 0, win = 0, gap = 0;
  bool deferred = false;
  double* far = nullptr;    // side buffer [7][far_cap] of the deferred permute: particles displaced beyond the window
  long long far_cap = 0;
  long long n_old = 0;      // slots of the pre-sort order spanned by src (deferred only)
  long long disp_fwd = 0, disp_bwd = 0;   // bounds of dest - j and j - dest over the slots j that src points at
  uint32_t* cell_start = nullptr;   // per species: first particle of every key bucket (ncell_ext + 2)
  long long cell_cap = 0;
  uint32_t* d_tile_start = nullptr; // first particle of every tile (+1 entry), device and pinned host copy
  uint32_t* h_tile_start = nullptr;
  long long tile_cap = 0;
  // migration mailboxes (particles.cu): capacity in particles of the links [axis][to lo, to hi, from hi, from lo],
  // the wanted counts of the last exchange (pinned copy, valid after the next synchronise) and whether they are unread
  long long mig_cap[3][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}};
  unsigned long long* h_mig = nullptr;
  unsigned long long* d_mig = nullptr;
  bool mig_pending[3] = {false, false, false};
  // thermal particle boundaries: ext_temp_<face> on the device (set_thermal_boundary), generator key and call counter
  double* ext_temp[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  unsigned long long thermal_seed = 0, thermal_calls = 0;
  bool sorted = false;      // keys/cell_start valid for the current positions ...
  bool was_ordered = false; // the slots are in (an older) cell order: a re-sort moves particles by about a tile
  double sort_dt = 0.0;     // ... and for this time step (the key is the cell after the half drift)
};

struct Laser {
  int id = 0;
  double amp = 0, pol = 0, t_start = 0, t_end = 0, t_env = 0, integral_phase = 0;
  double* profile = nullptr;   // device planes (or null)
  double* phase = nullptr;
  size_t plane = 0;
};

struct PhaseTimer {
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev;
  double ms = 0;
  long long launches = 0;
};

struct Nccl;  // comm.cu

}  // namespace eb

struct eb200_ctx {
  int ndims = 3, shape = 1, device = 0;
  int rank = 0, nproc = 1, nprocdir[3] = {1, 1, 1}, coords[3] = {0, 0, 0};
  bool min_bnd[3] = {true, true, true}, max_bnd[3] = {true, true, true};
  int nbr_lo[3] = {-1, -1, -1}, nbr_hi[3] = {-1, -1, -1};  // cartesian neighbours (rank or -1)
  int sf_min = -1, sf_max = 1, png = 3, ng = 5;
  bool grid_set = false;
  eb::GridP g;
  int nglob[3] = {1, 1, 1}, nglob_min[3] = {1, 1, 1};
  double dt = 0, time = 0;
  int step = 0;
  int bc_field[6] = {1, 1, 1, 1, 1, 1};
  bool add_laser[6] = {false, false, false, false, false, false};
  int field_order = 2, maxwell_solver = 0;
  double coeffs[15] = {0};
  // current_finish options (current_smooth.F90)
  int smooth = 0, smooth_its = 1, smooth_comp_its = 0;
  std::vector<int> smooth_strides;
  int use_current_correction = 0;
  double initial_j[3] = {0, 0, 0};
  double* smooth_wk = nullptr;      // 3 work arrays (allocated on first use)
  // CPML (boundary.F90 set_cpml_helpers / cpml_advance_*_currents): per axis the six 1-D profiles
  // kappa_e, kappa_b, sigma_e, sigma_b, a_e, a_b over (1-ng : n+ng) (device, contiguous), the slab
  // index ranges of both sides, the laser plane of a cpml_laser side, and the psi arrays
  // [axis][0: E(next axis), 1: E(next-next axis), 2: B(next axis), 3: B(next-next axis)]
  int cpml_thickness = 0;
  bool cpml_on = false;
  double* cpml_prof[3] = {nullptr, nullptr, nullptr};
  int cpml_start[6] = {0, 0, 0, 0, 0, 0}, cpml_end[6] = {-1, -1, -1, -1, -1, -1};
  int cpml_laser_idx[6] = {1, 1, 1, 1, 1, 1};
  double* diag_wk = nullptr;        // 5 arrays of g.total doubles (calc_derived)
  double* cpml_psi = nullptr;       // 12 arrays of g.total doubles (allocated when a CPML side exists)
  double* win_planes = nullptr;     // moving window: e*_x_max, b*_x_max planes (6 x y-z extent with ghosts)
  double* bnd_planes[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // setup_field_boundaries: per boundary 6 planes (ex..bz)
  bool bnd_dropped[6] = {false, false, false, false, false, false};   // ... dropped by redistribute, not handed over again yet
  double* fbase = nullptr;     // 9 * total doubles
  eb::FieldP f;
  std::vector<eb::SpeciesD> species;
  std::vector<eb::Laser> lasers[6];
  cudaStream_t stream = nullptr;
  std::string err;
  // scratch
  double* attr_scratch = nullptr; long long attr_scratch_cap = 0;   // one attribute array for the sort ping-pong
  uint32_t* cell_count = nullptr;   // ncell_ext + 2 (last bucket = invalid)
  long long cell_cap = 0;
  void* cub_tmp = nullptr; size_t cub_tmp_bytes = 0;
  double* halo_send = nullptr; double* halo_recv = nullptr; size_t halo_bytes = 0;
  double* mig_send[2] = {nullptr, nullptr}; double* mig_recv[2] = {nullptr, nullptr};
  long long mig_cap = 0;
  long long mig_min_particles = 1 << 16;   // option mig_capacity
  uint32_t* mig_list[2] = {nullptr, nullptr};   // slots leaving through the low / high side (listed by the push kernel)
  long long mig_list_cap = 0, mig_list_want = 0;
  // pinned staging ring of the host <-> device transfers (hostio.cu)
  static constexpr int NPIN = 3;
  void* pin[NPIN] = {nullptr, nullptr, nullptr};
  cudaEvent_t pin_ev[NPIN];
  bool pin_busy[NPIN] = {false, false, false};
  int pin_next = 0;
  size_t pin_bytes = (size_t)64 << 20;
  int host_threads = 0;         // option host_threads (0: min(12, hardware threads - 2))
  unsigned long long* d_counters = nullptr;   // small device scalar block (64 x u64)
  unsigned long long* h_counters = nullptr;   // pinned mirror
  double* d_red = nullptr;                    // reductions (64 doubles)
  double* h_red = nullptr;
  double* res_ref = nullptr;                  // captured charge residual
  double* rho = nullptr;
  double* rec_scratch = nullptr;              // deposit records of the split push/deposit kernels
  long long rec_scratch_cap = 0;              // ... capacity in particles
  // TMA descriptor of the E,B arrays for the push kernel's tile load (push_tiled.cu)
  alignas(64) unsigned char tmap[128] = {0};
  long long tmap_sig[6] = {0, 0, 0, 0, 0, 0};
  bool tmap_valid = false, tmap_ok = false;
  void* d_probeset = nullptr;           // device copy of the probe set of the species being pushed (tiled kernels)
  long long probe_capacity = 1 << 18;   // option probe_capacity: sampled particles kept per probe
  int use_tma = 1;              // option tma: 0 = stage the field tile with plain loads
  double* det_tiles = nullptr;                // deterministic deposit: one J tile per deposit CTA of a launch
  size_t det_tiles_cap = 0;                   // ... capacity in doubles
  long long rec_scratch_max = 1LL << 27;      // option deposit_scratch (particles per slab, 64 B each)
  // option push_overlap: the deposit of slab s runs on a second stream beside the push of slab s + 1 (the record
  // scratch is used as two halves); measured, see DESIGN.md section 8
  bool push_overlap = false;
  cudaStream_t stream2 = nullptr;
  cudaEvent_t ev_push[2] = {nullptr, nullptr}, ev_dep[2] = {nullptr, nullptr};
  // options
  int push_kernel = 0;          // 0 auto, 1 baseline (global atomics), 3 tiled (push + deposit kernels)
  int deterministic = 0;
  int fuse_keys = 1;            // option fuse_keys: push kernel emits the next sort keys (single rank)
  long long defer_window = -1;  // option defer_window: head room in slots of the deferred permute (-1 auto, 0 off)
  long long n_deferred = 0, n_permuted = 0;   // sorts that deferred / physically permuted (timer "deferred_sorts")
  bool timers_on = false;
  std::map<std::string, eb::PhaseTimer> timers;
  std::map<std::string, std::pair<cudaEvent_t, long long>> open_marks;
  long long launches = 0;       // kernels launched (for gpu_launches accounting)
  eb::Nccl* nccl = nullptr;
};

namespace eb {

int32_t fail(eb200_ctx* c, int32_t code, const std::string& msg);
// hostio.cu
int32_t staged_h2d(eb200_ctx* c, const void* host, size_t bytes, size_t granule,
                   const std::function<void*(size_t, size_t)>& dst, const std::function<int32_t(size_t, size_t)>& after);
int32_t staged_d2h(eb200_ctx* c, void* host, size_t bytes, size_t granule,
                   const std::function<const void*(size_t, size_t)>& src);
void free_pinned(eb200_ctx* c);
// window.cu
int32_t set_window_planes(eb200_ctx* c, const double* const planes[6]);
int32_t shift_window(eb200_ctx* c, double grid_min_local, double min_local, double max_local, double gmin, double gmax);
#define EB_CUDA(c, call)                                                                 \
  do {                                                                                   \
    cudaError_t e__ = (call);                                                            \
    if (e__ != cudaSuccess)                                                              \
      return eb::fail((c), e__ == cudaErrorMemoryAllocation ? EB200_E_NOMEM : EB200_E_CUDA, \
                      std::string(#call) + ": " + cudaGetErrorString(e__));              \
  } while (0)
#define EB_TRY(call)                 \
  do {                               \
    int32_t rc__ = (call);           \
    if (rc__ != 0) return rc__;      \
  } while (0)

struct ScopedPhase {
  eb200_ctx* c; const char* name; cudaEvent_t e0 = nullptr, e1 = nullptr; long long launches0 = 0;
  ScopedPhase(eb200_ctx* c_, const char* n);
  ~ScopedPhase();
};

// fdtd.cu
int32_t update_e_field(eb200_ctx* c);
int32_t update_b_field(eb200_ctx* c);
int32_t cpml_advance_currents(eb200_ctx* c, bool efield, double tstep);
// bcs.cu
int32_t field_bc(eb200_ctx* c, int first_comp, int ncomp);       // halo exchange of ncomp consecutive components
int32_t field_bc_array(eb200_ctx* c, double* a);                  // the same for one array of the field shape
int32_t zero_gradient_centre(eb200_ctx* c, double* a);
// diag.cu
int32_t calc_derived(eb200_ctx* c, int kind, int ispecies, int direction, double* out_host);
int32_t dist_fn(eb200_ctx* c, int ispecies, int curdims, const int32_t* direction, const double* ranges,
                const int32_t* resolution, const double* dgrid, const double* restrictions,
                const int32_t* use_restrictions, double* out_host);
int32_t efield_bcs(eb200_ctx* c);
int32_t bfield_bcs(eb200_ctx* c, bool mpi_only);
int32_t bfield_final_bcs(eb200_ctx* c);
int32_t setup_field_boundaries(eb200_ctx* c);
int32_t set_boundary_fields(eb200_ctx* c, int b, const double* const planes[6]);
int32_t current_finish(eb200_ctx* c);
int32_t summation_bcs(eb200_ctx* c, double* arr, int flip_dir, const int* bc6 = nullptr);  // processor_summation_bcs on one array
bool bc_mixed(const eb200_ctx* c);                               // species with different particle boundaries
// comm.cu
int32_t nccl_unique_id(void* out128);
int32_t comm_init(eb200_ctx* c, const void* id);
void comm_destroy(eb200_ctx* c);
// exchange `count` doubles with the low/high neighbours along an axis: send_lo goes to nbr_lo and
// recv_hi comes from nbr_hi (first call), etc.  Null neighbours are skipped.
int32_t comm_sendrecv(eb200_ctx* c, const double* send, int dst, double* recv, int src, size_t count);
int32_t comm_sendrecv_v(eb200_ctx* c, const double* send, int dst, size_t nsend, double* recv, int src, size_t nrecv);
// both directions of an axis in ONE group: send a to rank da and b to rank db, receive ra from rank sa and rb from sb
int32_t comm_exchange2(eb200_ctx* c, const double* a, int da, size_t na, const double* b, int db, size_t nb,
                       double* ra, int sa, size_t nra, double* rb, int sb, size_t nrb);
int32_t comm_alltoallv(eb200_ctx* c, const double* send, const size_t* off_s, const size_t* cnt_s, double* recv,
                       const size_t* off_r, const size_t* cnt_r);
// balance.cu / particles.cu: the new owner of a position (get_particle_processor, balance.F90:2673-2746):
// lb[a][i] = lower edge x_grid_mins(i) - dx/2 of processor coordinate i along axis a (lb[a][0] unused)
constexpr int MAXP_AXIS = 32;
struct OwnerP { int np[3]; double lb[3][MAXP_AXIS]; };
int32_t redistribute_particles(eb200_ctx* c, const OwnerP& own, std::vector<long long>& n_slots);
int32_t get_load(eb200_ctx* c, const double grid_min_global[3], long long* const load[3]);
int32_t redistribute(eb200_ctx* c, const int32_t* old_cell_min, const int32_t* old_cell_max, const int32_t* new_cell_min,
                     const int32_t* new_cell_max, const double grid_min_global[3], const double grid_min_local[3],
                     const double min_local[3], const double max_local[3]);
// particles.cu
int32_t push_particles(eb200_ctx* c);
int32_t particle_bcs(eb200_ctx* c);
int32_t species_reserve(eb200_ctx* c, SpeciesD& sp, long long need, bool exact = false);
int32_t add_particles(eb200_ctx* c, int isp, long long n, const double* packed, int nvar, bool replace);
int32_t get_particles(eb200_ctx* c, int isp, long long first, long long cap, double* packed, int nvar, long long* n_out);
int32_t sort_species(eb200_ctx* c, SpeciesD& sp, long long n_slots, bool keys_ready = false, bool may_defer = false);
void set_view(SpeciesD& sp, long long off);             // a[k] = base[k] + off, key = key_base + off
int32_t settle_species(eb200_ctx* c, SpeciesD& sp);    // finish a deferred permute (no-op otherwise)
int32_t settle_particles(eb200_ctx* c);                // ... for every species
// reduce.cu
int32_t energy(eb200_ctx* c, double* fe, double* per_species);
int32_t charge_residual(eb200_ctx* c, double* max_res, double* max_rho, bool capture);
int32_t charge_residual_static(eb200_ctx* c, double* max_res, double* max_rho);   // no allocation: file-scope device scratch
long long charge_residual_static_max();

}  // namespace eb
