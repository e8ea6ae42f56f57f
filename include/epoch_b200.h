/*
 * epoch_b200.h -- C ABI of the B200-native EPOCH PIC hot path.
 *
 * This is the drop-in boundary.  EPOCH (Fortran 90 + MPI) has no FFI today: the
 * "interface" of its per-timestep core is four argument-less module procedures
 * acting on MODULE shared_data globals, called from PROGRAM pic:
 *
 *     update_eb_fields_half    epoch3d/src/fields.f90:819    (call epoch3d.F90:208)
 *     push_particles           epoch3d/src/particles.F90:28  (call epoch3d.F90:213)
 *     current_finish           epoch3d/src/housekeeping/current_smooth.F90:29 (call epoch3d.F90:235)
 *     update_eb_fields_final   epoch3d/src/fields.f90:851    (call epoch3d.F90:246)
 *
 * Each entry below names the reference item it replaces.  The Fortran side binds
 * these with ISO_C_BINDING (see fortran/epoch_b200_iface.F90).
 *
 * Conventions
 *  - every function returns 0 on success, <0 on error (EB200_E_*); the message
 *    is available from eb200_last_error().  Nothing throws across the ABI.
 *  - all reals are IEEE fp64 (EPOCH's REAL(num), constants.F90:24), SI units.
 *  - grid arrays are Fortran shaped *including ghosts*: (1-ng:nx+ng [,1-ng:ny+ng
 *    [,1-ng:nz+ng]]), column major (x fastest); ng = 4/5/6 for tophat/triangle/
 *    bspline3 (constants.F90:472-492).  Dimensions that do not exist in a 1D/2D
 *    build are simply absent from the array (as in epoch1d / epoch2d).
 *  - packed particles use the partlist wire format (housekeeping/partlist.F90:
 *    411-553): nvar consecutive doubles per particle = pos(ndims), p(3), weight.
 *  - one context per MPI rank == one GPU.  A context is thread-compatible, not
 *    thread-safe.  Hot entries enqueue work on a context-private CUDA stream and
 *    may return before completion; eb200_sync()/download/energy synchronise.
 *
 * The same declarations are re-used (with prefix orc_) by the CPU oracle in
 * oracle/, which is TEST INFRASTRUCTURE ONLY.
 */
#ifndef EPOCH_B200_H
#define EPOCH_B200_H

#include <stdint.h>

#ifndef EB200_NAME
#define EB200_NAME(x) eb200_##x
#endif
#ifndef EB200_CTX
#define EB200_CTX eb200_ctx
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef struct EB200_CTX EB200_CTX;

/* error classes */
enum {
  EB200_OK = 0,
  EB200_E_ARG = -1,      /* bad argument / unsupported option          */
  EB200_E_STATE = -2,    /* call sequence error (grid not set, ...)     */
  EB200_E_CUDA = -3,     /* CUDA runtime failure                        */
  EB200_E_NCCL = -4,     /* NCCL failure / NCCL not loadable            */
  EB200_E_NOMEM = -5,    /* out of device (or host) memory              */
  EB200_E_UNSUPPORTED = -6 /* reference feature outside the hot path    */
};

/* particle shapes: compile-time #ifdefs in the reference (Makefile:186-207) */
enum { EB200_SHAPE_TOPHAT = 0, EB200_SHAPE_TRIANGLE = 1, EB200_SHAPE_BSPLINE3 = 2 };

/* boundary codes, identical to constants.F90:56-71 */
enum {
  EB200_BC_PERIODIC = 1, EB200_BC_OTHER = 2, EB200_BC_SIMPLE_LASER = 3,
  EB200_BC_SIMPLE_OUTFLOW = 4, EB200_BC_OPEN = 5, EB200_BC_DUMP = 6,
  EB200_BC_ZERO_GRADIENT = 7, EB200_BC_CLAMP = 8, EB200_BC_REFLECT = 9,
  EB200_BC_CONDUCT = 10, EB200_BC_THERMAL = 11, EB200_BC_CPML_LASER = 12,
  EB200_BC_CPML_OUTFLOW = 13, EB200_BC_MIXED = 14, EB200_BC_HEAT_BATH = 15
};
/* boundary location codes 1..6 = x_min,x_max,y_min,y_max,z_min,z_max (constants.F90:74-79) */

/* ---- lifecycle -------------------------------------------------------------
 * Replaces the allocation half of mpi_initialise (housekeeping/mpi_routines.F90:
 * 393-593) and setup_communicator (:274-389).  nprocdir/coords describe the
 * Cartesian process grid (x,y,z); nccl_unique_id is 128 bytes obtained from
 * eb200_nccl_unique_id() on rank 0 and broadcast by the host (MPI_BCAST); it may
 * be NULL when nproc == 1.  device_ordinal >= 0 names the CUDA device; a negative value -(r + 1) hands over the
 * rank's position r among the ranks of its node (MPI_COMM_TYPE_SHARED) and the library takes device
 * r mod cudaGetDeviceCount, so the host needs no knowledge of the node's GPU count. */
int32_t EB200_NAME(create)(EB200_CTX** out, int32_t ndims, int32_t shape,
                           int32_t device_ordinal, int32_t rank, int32_t nproc,
                           const int32_t nprocdir[3], const int32_t coords[3],
                           const void* nccl_unique_id);
int32_t EB200_NAME(nccl_unique_id)(void* out128);
int32_t EB200_NAME(destroy)(EB200_CTX*);
/* Page-locked host memory for the caller's particle / field buffers (in Fortran: C_F_POINTER on the result
 * instead of ALLOCATE, cf. the packed buffers of partlist.F90:808-862).  Uploads and downloads recognise such
 * buffers and let the copy engine read / write them directly at the PCIe rate; any other (pageable) buffer is
 * staged through the library's own pinned ring.  No context needed. */
int32_t EB200_NAME(host_alloc)(void** out, int64_t bytes);
int32_t EB200_NAME(host_free)(void* p);

/* ---- grid / time -----------------------------------------------------------
 * Mirrors the scalars of shared_data.F90:453,508-523 set by setup_grid
 * (housekeeping/setup.F90:174-231) and setup_grid_x (utilities.f90:343-373).
 * n_global_min is the 1-based global index of local cell 1.  grid_min_local is
 * the centre of local cell 1 (x_grid_min_local); min_local/max_local are the
 * particle-ownership bounds (x_min_local, x_max_local); min_global/max_global
 * are x_min/x_max of the deck.  Call again after a re-decomposition. */
int32_t EB200_NAME(set_grid)(EB200_CTX*, const int32_t n_local[3], const int32_t n_global[3],
                             const int32_t n_global_min[3], const double d[3],
                             const double grid_min_local[3], const double min_local[3],
                             const double max_local[3], const double min_global[3],
                             const double max_global[3], int32_t cpml_thickness);
/* dt, time and step as the reference's globals hold them at the next hot call
 * (epoch3d.F90:153-159,241-246). */
int32_t EB200_NAME(set_time)(EB200_CTX*, double dt, double time, int32_t step);
/* bc_field(1:6) after setup_boundaries normalisation is applied internally
 * (boundary.F90:30-75). */
int32_t EB200_NAME(set_field_bcs)(EB200_CTX*, const int32_t bc_field[6]);
/* set_field_order / set_maxwell_solver (fields.f90:34-179).  maxwell_solver uses
 * the reference's codes (constants.F90:125-132: custom -1, yee 0, lehe_x/y/z 2/3/4,
 * cowan 5, pukhov 6).  coeffs = alphax,alphay,alphaz, betaxy,betaxz,betayx,betayz,
 * betazx,betazy, gammax,gammay,gammaz, deltax,deltay,deltaz as set_maxwell_solver
 * leaves them on the host (terms of absent dimensions zero); ignored for Yee.  The
 * extended stencils act in the second-order update_b_field only (fields.f90:655-732). */
int32_t EB200_NAME(set_solver)(EB200_CTX*, int32_t field_order, int32_t maxwell_solver,
                               const double coeffs[15]);

/* CPML absorbing boundaries (bc_field = cpml_laser / cpml_outflow).  set_grid's cpml_thickness
 * is the deck's value (the host has already widened the grid, setup.F90:180-189,400-414).
 * set_cpml hands over, per axis, what set_cpml_helpers (boundary.F90:1989-2479; 1D :815-966)
 * leaves on the host: the six profiles cpml_kappa_e*, cpml_kappa_b*, cpml_sigma_e*,
 * cpml_sigma_b*, cpml_a_e*, cpml_a_b* over (1-ng : n+ng) (kappa = 1, others 0 outside the
 * layer), the local slab ranges cpml_*_min_start/end and cpml_*_max_start/end, and the local
 * laser-plane indices cpml_*_min/max_laser_idx (pass -1 when that plane is on another rank).
 * It must be called for EVERY axis of the simulation once a CPML boundary is set.  The library
 * then uses the kappa-scaled branches of update_e_field / update_b_field (fields.f90:192-310,
 * :437-630) followed by cpml_advance_e_currents / cpml_advance_b_currents (boundary.F90:
 * 2483-2815), zero-gradient ghosts on CPML sides, the laser plane at laser_idx, and removes
 * particles beyond x_min_outer / x_max_outer on a CPML side (boundary.F90:1386-1402). */
int32_t EB200_NAME(set_cpml)(EB200_CTX*, int32_t axis, const double* kappa_e, const double* kappa_b,
                             const double* sigma_e, const double* sigma_b, const double* a_e,
                             const double* a_b, int32_t min_start, int32_t min_end, int32_t max_start,
                             int32_t max_end, int32_t laser_idx_min, int32_t laser_idx_max);

/* current_finish options (housekeeping/current_smooth.F90:29-151): the strided compensated
 * binomial filter of smooth_current (smooth_its passes, smooth_comp_its = 0|1 extra pass, the
 * deck's smooth_strides; strides <= ng) and use_current_correction, which subtracts the three
 * scalars initial_jx,jy,jz that calc_initial_current (io/calc_df.F90:1292-1389) leaves on the
 * host.  nstrides = 0 means the default stride list {1}; smooth = 0 switches the filter off. */
int32_t EB200_NAME(set_current_options)(EB200_CTX*, int32_t smooth, int32_t smooth_its, int32_t smooth_comp_its,
                                        int32_t nstrides, const int32_t* strides, int32_t use_current_correction,
                                        const double initial_j[3]);

/* Moving window (housekeeping/window.F90).  shift_window moves the window by ONE cell, i.e. one
 * iteration of the loop in shift_window (window.F90:62-94): it takes this rank's new
 * x_grid_min_local, x_min_local, x_max_local (setup_grid_x, utilities.f90:343-373) and the new
 * global x_min, x_max, drops the particles with pos_x < x_min on an x_min boundary rank
 * (remove_particles, :362-384) and shifts all nine grid arrays one cell towards -x, refreshes
 * their halos and fixes the incoming cells on an x_max boundary rank (shift_fields, :98-194).
 * The new plasma of insert_particles (:198-358; deck expressions + host RNG) is appended by the
 * host with append_particles BEFORE the call; particle_bcs follows it (moving_window, :425).
 * set_window_fields uploads the planes ex_x_max .. bz_x_max of setup_field_boundaries
 * (setup.F90:425-470), each (1-ng:ny+ng [, 1-ng:nz+ng]), NULL = zero. */
int32_t EB200_NAME(set_window_fields)(EB200_CTX*, const double* ex_x_max, const double* ey_x_max,
                                      const double* ez_x_max, const double* bx_x_max,
                                      const double* by_x_max, const double* bz_x_max);
/* setup_field_boundaries (housekeeping/setup.F90:426-510, called from after_deck_last :241): capture, from
 * the fields NOW on the device, the 36 planes ex_x_min .. bz_z_max -- the field at the boundary plane
 * (tangential E, normal B) or the mean of the boundary plane and the one below it (normal E, tangential B);
 * the plane beyond the laser plane on a cpml_laser side.  outflow_bcs_* add them back at open / laser
 * boundaries (laser.f90:536,559-572 and the five cyclic copies): an external static field, or the fields of a
 * restart, pass through such a boundary unchanged.  Without the call the planes are zero.
 * set_boundary_fields uploads the six planes of ONE boundary (1..6 = x_min, x_max, y_min, ...; each over the
 * other axes incl. ghosts, lower axis fastest; NULL = zero) -- for a host that keeps its own copies, and after
 * redistribute, which drops them. */
int32_t EB200_NAME(setup_field_boundaries)(EB200_CTX*);
int32_t EB200_NAME(set_boundary_fields)(EB200_CTX*, int32_t boundary, const double* ex, const double* ey,
                                        const double* ez, const double* bx, const double* by, const double* bz);
int32_t EB200_NAME(shift_window)(EB200_CTX*, double x_grid_min_local, double x_min_local, double x_max_local,
                                 double x_min, double x_max);

/* ---- species / particles ---------------------------------------------------
 * TYPE particle_species scalars used by the push (shared_data.F90:184-263). */
int32_t EB200_NAME(set_species)(EB200_CTX*, int32_t ispecies, double charge, double mass,
                                double species_weight, int32_t immobile, int32_t zero_current,
                                int32_t species_type, const int32_t bc_particle[6]);
/* list -> device import (replaces walking attached_list; wire format of
 * pack_particle, partlist.F90:411-473).  upload replaces the species' particles,
 * append adds to them (chunked uploads, injectors, window insert_particles). */
int32_t EB200_NAME(upload_particles)(EB200_CTX*, int32_t ispecies, int64_t n,
                                     const double* packed, int32_t nvar);
int32_t EB200_NAME(append_particles)(EB200_CTX*, int32_t ispecies, int64_t n,
                                     const double* packed, int32_t nvar);
int32_t EB200_NAME(particle_count)(EB200_CTX*, int32_t ispecies, int64_t* n);
/* device -> list export (unpack_particle, partlist.F90:477-553); first = offset
 * of the first particle to copy so the host can stream in chunks. */
int32_t EB200_NAME(download_particles)(EB200_CTX*, int32_t ispecies, int64_t first,
                                       int64_t capacity, double* packed, int32_t nvar,
                                       int64_t* n_out);

/* ---- fields ----------------------------------------------------------------
 * ex..jz of shared_data.F90:461 (allocated mpi_routines.F90:531-540).  NULL
 * pointers are skipped. */
int32_t EB200_NAME(upload_fields)(EB200_CTX*, const double* ex, const double* ey, const double* ez,
                                  const double* bx, const double* by, const double* bz,
                                  const double* jx, const double* jy, const double* jz);
int32_t EB200_NAME(download_fields)(EB200_CTX*, double* ex, double* ey, double* ez,
                                    double* bx, double* by, double* bz,
                                    double* jx, double* jy, double* jz);

/* ---- laser / outflow boundaries ----------------------------------------------
 * TYPE laser_block (shared_data.F90) as used by outflow_bcs_* (laser.f90:509-973).
 * profile/phase are the boundary planes over indices 0..n of the tangential
 * dimensions, x-fastest, i.e. (ny+1)*(nz+1) doubles for an x boundary in 3D,
 * (ny+1) in 2D, 1 in 1D; NULL = 1.0 / 0.0.  The host evaluates the deck's time
 * profile and phase integral each step (laser.f90:207-226,334-346) and passes
 * them through set_laser_step. */
int32_t EB200_NAME(set_laser)(EB200_CTX*, int32_t boundary, int32_t id, double amp,
                              double pol_angle, double t_start, double t_end,
                              const double* profile, const double* phase);
int32_t EB200_NAME(set_laser_step)(EB200_CTX*, int32_t boundary, int32_t id, double t_env,
                                   double integral_phase);

/* ---- the four hot entries --------------------------------------------------*/
int32_t EB200_NAME(update_eb_fields_half)(EB200_CTX*);   /* fields.f90:819-847   */
int32_t EB200_NAME(push_particles)(EB200_CTX*);          /* particles.F90:28-614 incl. particle_bcs */
int32_t EB200_NAME(current_finish)(EB200_CTX*);          /* current_smooth.F90:29-45 */
int32_t EB200_NAME(update_eb_fields_final)(EB200_CTX*);  /* fields.f90:851-872   */
/* initial boundary calls of PROGRAM pic (epoch3d.F90:141-142,157) */
int32_t EB200_NAME(particle_bcs)(EB200_CTX*);            /* boundary.F90:1316    */
int32_t EB200_NAME(efield_bcs)(EB200_CTX*);              /* boundary.F90:1148    */
int32_t EB200_NAME(bfield_final_bcs)(EB200_CTX*);        /* boundary.F90:1267    */

/* ---- diagnostics -----------------------------------------------------------
 * calc_total_energy_sum (io/calc_df.F90:1190-1288): local (this rank) sums; the
 * host MPI_REDUCEs them as the reference does. */
int32_t EB200_NAME(energy)(EB200_CTX*, double* field_energy,
                           double* particle_energy_per_species);
/* max |div E - rho/eps0| and max |rho/eps0| over interior cells, rho deposited
 * with the normalised weights of calc_charge_density (calc_df.F90:439-519). */
int32_t EB200_NAME(charge_residual)(EB200_CTX*, double* max_abs_residual,
                                    double* max_abs_rho_over_eps0);
/* charge_residual_static: NO COUNTERPART IN THE REFERENCE -- added for callers
 * that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs
 * (the context's fields and particles) and outputs as charge_residual, but the charge-density scratch is a
 * file-scope static __device__ array of the library instead of a lazily cudaMalloc'ed one, and the call allocates
 * and frees nothing.  Limits: grids of at most 2 097 152 points including ghost cells (16 MiB, held by every GPU
 * that loads the library until exit; e.g. 118^3 cells with the triangle shape) -- larger grids are rejected with
 * EB200_E_ARG; single-rank contexts only (EB200_E_UNSUPPORTED otherwise).  Every call on a device shares the
 * array: it is cleared at the start of each call, and calls that overlap on the GPU (two contexts) race. */
int32_t EB200_NAME(charge_residual_static)(EB200_CTX*, double* max_abs_residual, double* max_abs_rho_over_eps0);
/* Load balancing (housekeeping/balance.F90).  The decision -- balance_workload :107-319 and
 * calculate_breaks :2526-2669, a few hundred integers -- stays on the host; the device supplies
 * its input and carries the result out.
 * get_load: this rank's part of part_load_func (:2468-2495), i.e. the arrays load_x, load_y,
 * load_z (1-ng : n_global+ng, any of them may be NULL) BEFORE the MPI_ALLREDUCE and the
 * push_per_field weighting of get_load (:2411-2464); grid_min_global = x_grid_min, y_grid_min,
 * z_grid_min (centre of global cell 1).
 * redistribute: redistribute_domain (:411-472) for the state the library holds: the nine field
 * arrays and the CPML psi arrays move as redistribute_field_3d does (:1652-1966: intersections of
 * old and new sub-domains, edge ghost cells travel with the edge cells, then
 * do_field_mpi_with_lengths), the particles as distribute_particles does (:2751-2809,
 * get_particle_processor :2673-2746).  The four tables are cell_{x,y,z}_min/max and
 * new_cell_{x,y,z}_min/max concatenated [x | y | z] with nprocdir[a] entries per existing axis;
 * the three scalar triples are the rank's NEW *_grid_min_local, *_min_local, *_max_local.
 * Afterwards the grid is the new one (as after set_grid); the host re-sends what it redistributes
 * itself: laser profiles (set_laser), window planes (set_window_fields), CPML profiles (set_cpml). */
int32_t EB200_NAME(get_load)(EB200_CTX*, const double grid_min_global[3], int64_t* load_x, int64_t* load_y,
                             int64_t* load_z);
int32_t EB200_NAME(redistribute)(EB200_CTX*, const int32_t* old_cell_min, const int32_t* old_cell_max,
                                 const int32_t* new_cell_min, const int32_t* new_cell_max,
                                 const double grid_min_global[3], const double grid_min_local[3],
                                 const double min_local[3], const double max_local[3]);

/* Derived grid variables of io/calc_df.F90, computed where the particles live so that
 * output_routines pulls one reduced grid instead of the particle lists.
 * kind: EB200_DF_* below; ispecies: 0-based species, or -1 for the sum over species
 * (current_species <= 0; zero_current species are then skipped as the reference does);
 * direction: 1, 2, 3 = c_dir_x, c_dir_y, c_dir_z, negative for the backward ekflux, 0 = absent
 * (temperature: all three degrees of freedom).  `out` is a host array of the field shape
 * (1-ng : n+ng per axis); cells the reference leaves undefined (poynt_flux ghosts) are zero.
 *   MASS_DENSITY calc_mass_density :35      EKBAR calc_ekbar :119         EKFLUX calc_ekflux :232
 *   POYNT_FLUX calc_poynt_flux :382         CHARGE_DENSITY calc_charge_density :439
 *   NUMBER_DENSITY calc_number_density :523 PPC calc_ppc :598             AVERAGE_WEIGHT :652
 *   TEMPERATURE calc_temperature :723       SPECIES_CURRENT calc_per_species_current :994
 *   AVERAGE_MOMENTUM calc_average_momentum :1104 */
#define EB200_DF_MASS_DENSITY 0
#define EB200_DF_EKBAR 1
#define EB200_DF_EKFLUX 2
#define EB200_DF_POYNT_FLUX 3
#define EB200_DF_CHARGE_DENSITY 4
#define EB200_DF_NUMBER_DENSITY 5
#define EB200_DF_PPC 6
#define EB200_DF_AVERAGE_WEIGHT 7
#define EB200_DF_TEMPERATURE 8
#define EB200_DF_SPECIES_CURRENT 9
#define EB200_DF_AVERAGE_MOMENTUM 10
int32_t EB200_NAME(calc_derived)(EB200_CTX*, int32_t kind, int32_t ispecies, int32_t direction, double* out);
/* The particle loop of general_dist_fn (io/dist_fn.F90:532-628): weighted histogram of one species over
 * curdims (1..3) of the directions c_dir_* of constants.F90:202-213 -- 1..ndims = x,y,z; then, numbered from
 * ndims+1: px, py, pz, en, gamma_m1, xy_angle, yz_angle, zx_angle, mod_p (EB200_DIR_* below are those offsets).
 * The host keeps what precedes the loop in the reference (:204-335,490-518: snapping spatial ranges to cells,
 * the -1..1 fall-back of an empty range, dgrid) and hands over ranges[2*idim + {0,1}], resolution, dgrid;
 * restrictions[2*(dir-1) + {0,1}] / use_restrictions[dir-1] for dir = 1..ndims+9, or NULL for none.  out has
 * resolution[0]*resolution[1]*resolution[2] doubles (unused dimensions: 1), first dimension fastest, and
 * receives this rank's sums (the reference's MPI_REDUCE over the sub-communicator stays on the host). */
#define EB200_DIR_PX 1
#define EB200_DIR_PY 2
#define EB200_DIR_PZ 3
#define EB200_DIR_EN 4
#define EB200_DIR_GAMMA_M1 5
#define EB200_DIR_XY_ANGLE 6
#define EB200_DIR_YZ_ANGLE 7
#define EB200_DIR_ZX_ANGLE 8
#define EB200_DIR_MOD_P 9
int32_t EB200_NAME(dist_fn)(EB200_CTX*, int32_t ispecies, int32_t curdims, const int32_t direction[3],
                            const double ranges[6], const int32_t resolution[3], const double dgrid[3],
                            const double* restrictions, const int32_t* use_restrictions, double* out);
/* Particle probes (particles.F90:567-606, TYPE particle_probe shared_data.F90:392-404): a plane (point, normal)
 * and an energy window per (species, probe).  During push_particles every particle of the species whose kinetic
 * energy (gamma - 1) m c^2 lies in (ek_min, ek_max) and that crosses the plane in the direction of the normal
 * (d_init >= 0 > d_final, :587-592) is copied -- position and momentum at the end of the step, weight -- to the
 * probe's list (the reference's sampled_particles), which the host reads and clears when it writes a dump.
 * iprobe = 0..7; point/normal have ndims components.  A species with probes is pushed by the general kernel. */
int32_t EB200_NAME(set_probe)(EB200_CTX*, int32_t ispecies, int32_t iprobe, const double* point, const double* normal,
                              double ek_min, double ek_max);
/* n_out: particles held in the probe's device list since the last clear (at most option "probe_capacity", default
 * 262144); copies min(n_out, capacity) of them (partlist wire format) when packed != NULL; clear != 0 empties the
 * list.  If more particles crossed than the list holds, the kept ones are still copied and the call returns
 * EB200_E_NOMEM (last_error says how many were sampled). */
int32_t EB200_NAME(download_probe)(EB200_CTX*, int32_t ispecies, int32_t iprobe, int64_t capacity, double* packed,
                                   int32_t nvar, int64_t* n_out, int32_t clear);
/* Thermal particle boundaries (c_bc_thermal; particle_bcs, boundary.F90:1405-1461 and the five other faces): a
 * particle of the species that passes the OUTER edge of boundary `boundary` (1..6 = x_min .. z_max) is re-emitted --
 * temperature interpolated from ext_temp (triangle weighting at its transverse position), momenta drawn in the order
 * x, y, z (normal component: flux_momentum_from_temperature heading inwards, the others momentum_from_temperature,
 * particle_temperature.F90:396-469, drift 0), position mirrored at the outer edge.  ext_temp is the species'
 * ext_temp_<face> array as the reference allocates it (mpi_routines.F90:561-584): transverse extents 1-ng:n+ng,
 * first transverse axis fastest, then the 3 temperature components (1D: 3 numbers).  Must be called (again after a
 * re-slice) before the first push of a species whose set_species lists EB200_BC_THERMAL on that face.  `seed` keys
 * the device's counter-based generator (one independent stream per re-emitted particle and call; the reference's
 * generator is one sequential host stream whose draws depend on the list order, so only the distributions agree).
 * EB200_BC_HEAT_BATH is accepted by set_species as EB200_BC_OPEN: setup_domain_dependent_boundaries (boundary.F90:
 * 75-93) turns it into an open boundary plus a boundary injector, and injectors stay on the host. */
int32_t EB200_NAME(set_thermal_boundary)(EB200_CTX*, int32_t ispecies, int32_t boundary, const double* ext_temp,
                                         int64_t seed);
/* options (key, value as strings):
 *   "deterministic_deposit" "0|1"   run-to-run bitwise reproducible currents (tiled kernels: 1D/2D/3D)
 *   "push_kernel"   "auto|baseline|tiled"   auto / tiled: the tiled push + deposit kernels (1D/2D/3D); baseline: the
 *                                   general kernel with one global RED per deposited value (debugging)
 *   "defer_window"  "auto|<slots>"  head room of the deferred permute (0: permute inside the sort); takes effect
 *                                   when the particle arrays are next allocated ("reserve", upload); auto = an eighth of
 *                                   the species and at least 4.5 layers of tiles along the slowest axis
 *   "push_overlap"  "0|1"           deposit of a slab on a second stream beside the next slab's push (measured: no gain)
 *   "reserve"       "<species>:<particles>"  allocate a species' storage once
 *   "deposit_scratch" "<particles>" particles per push/deposit launch (64 B of scratch each)
 *   "fuse_keys" "0|1", "tma" "0|1", "timers" "0|1", "host_threads" "<n>", "pin_bytes" "<bytes>",
 *   "mig_capacity" "<particles>"   first capacity of a migration mailbox (per neighbour and direction; the SAME on every
 *                                   rank); it doubles by itself when more particles want to cross
 *   "capture_charge_reference" "1" */
int32_t EB200_NAME(set_option)(EB200_CTX*, const char* key, const char* value);
int32_t EB200_NAME(sync)(EB200_CTX*);
/* milliseconds of device time spent in the named phase since the last call with
 * reset != 0 ("push", "push_kernel", "push_A", "push_B", "sort", "fields", "halo", "bcs", "window", any name given to
 * eb200_mark); CUDA-event based.  Counters read through the same call (in *launches): "launches", "deferred_sorts",
 * "permuted_sorts", "cell_mismatch" (must stay 0), "thermal_reemitted". */
int32_t EB200_NAME(timer)(EB200_CTX*, const char* phase, int32_t reset, double* ms,
                          int64_t* launches);
/* user bracket on the context's stream: mark(phase, 1) records a begin event,
 * mark(phase, 0) the matching end event; the device time between them is added
 * to `phase` and read back with timer().  Replaces timer_start/timer_stop of
 * housekeeping/timer.f90:27-113 (wall clock there, CUDA events here).  Always on,
 * independent of the "timers" option. */
int32_t EB200_NAME(mark)(EB200_CTX*, const char* phase, int32_t begin);
const char* EB200_NAME(last_error)(EB200_CTX*);
const char* EB200_NAME(version)(void);

#ifdef __cplusplus
}
#endif
#endif /* EPOCH_B200_H */
