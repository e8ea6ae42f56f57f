! epoch_b200_hooks.F90 -- drop-in replacements, under the reference's own names,
! for the four hot entries of PROGRAM pic (epoch3d/src/epoch3d.F90:208,213,235,246)
! plus the lazy host <-> device synchronisation the rest of EPOCH needs.
!
! Usage in the reference tree (see INTEGRATION.md): add this file and
! epoch_b200_iface.F90 to SRCFILES, link -lepoch_b200, and in epoch3d.F90 replace
!     USE fields    -> USE epoch_b200_hooks, ONLY: update_eb_fields_half, update_eb_fields_final
!     USE particles -> USE epoch_b200_hooks, ONLY: push_particles
!     USE current_smooth -> USE epoch_b200_hooks, ONLY: current_finish
! and call gpu_attach after mpi_initialise / set_dt, gpu_sync_to_host before anything
! that reads particles or fields on the host (output_routines when a dump fires,
! balance_workload, moving_window, restart), gpu_sync_to_device after anything that
! wrote them.  A step with no dump / balance / window does no transfer at all.
!
! This file cannot be compiled in the authoring image (no Fortran compiler); it is
! written against the reference's module variables by name and is reviewed, not built
! (UNVERIFIED: wherever a gfortran exists, `gfortran -cpp -DEPOCH_NDIMS=3 -fsyntax-only -I<epoch3d/obj>`
! against the reference's module files is the first check to run).
!
! One source for epoch1d / epoch2d / epoch3d: compile with -DEPOCH_NDIMS=1|2|3 (default 3).  The C ABI
! always takes three-element vectors; the macros below fill the axes a lower-dimensional code does not have.
#ifndef EPOCH_NDIMS
#define EPOCH_NDIMS 3
#endif
#if EPOCH_NDIMS == 1
#define I3(a, b, c) [a, 1, 1]
#define R3(a, b, c, fill) [a, fill, fill]
#define GHOST_DIMS DIMENSION(1-ng:)
#elif EPOCH_NDIMS == 2
#define I3(a, b, c) [a, b, 1]
#define R3(a, b, c, fill) [a, b, fill]
#define GHOST_DIMS DIMENSION(1-ng:,1-ng:)
#else
#define I3(a, b, c) [a, b, c]
#define R3(a, b, c, fill) [a, b, c]
#define GHOST_DIMS DIMENSION(1-ng:,1-ng:,1-ng:)
#endif
MODULE epoch_b200_hooks

  USE, INTRINSIC :: ISO_C_BINDING
  USE epoch_b200_iface
  USE shared_data          ! ex..bz, jx..jz, species_list, dt, time, step, nx.., x_grid_min_local.., bc_field, ...
  USE partlist             ! pack_particle, unpack_particle, create_empty_partlist, add_particle_to_partlist
  USE utilities            ! abort_code
  IMPLICIT NONE

  TYPE(C_PTR), SAVE :: gpu = C_NULL_PTR
  LOGICAL, SAVE :: host_stale = .FALSE.      ! device copy is newer than the host copy
  LOGICAL, SAVE :: device_stale = .TRUE.     ! host copy is newer than the device copy
#ifdef PARTICLE_SHAPE_TOPHAT
  INTEGER(C_INT32_T), PARAMETER :: gpu_shape = 0
#elif defined(PARTICLE_SHAPE_BSPLINE3)
  INTEGER(C_INT32_T), PARAMETER :: gpu_shape = 2
#else
  INTEGER(C_INT32_T), PARAMETER :: gpu_shape = 1
#endif

CONTAINS

  SUBROUTINE gpu_check(rc)
    ! reference convention: print and MPI_ABORT (housekeeping/utilities.f90:261-281)
    INTEGER(C_INT32_T), INTENT(IN) :: rc
    CHARACTER(KIND=C_CHAR), POINTER :: msg(:)
    INTEGER :: i
    IF (rc == 0) RETURN
    CALL C_F_POINTER(eb200_last_error(gpu), msg, [512])
    DO i = 1, 512
      IF (msg(i) == C_NULL_CHAR) EXIT
      WRITE(*, '(A)', ADVANCE='NO') msg(i)
    END DO
    WRITE(*,*)
    CALL abort_code(c_err_terminate)
  END SUBROUTINE gpu_check

  SUBROUTINE gpu_attach
    ! after mpi_initialise (mpi_routines.F90:393-593), setup_boundaries and set_dt
    CHARACTER(KIND=C_CHAR), TARGET :: id(128)
    INTEGER(C_INT32_T) :: nprocdir(3), crd(3), ispecies, bcp(6)
    INTEGER :: ierr, node_comm, local_rank
    REAL(num) :: species_weight
    IF (rank == 0) CALL gpu_check(eb200_nccl_unique_id(id))
    CALL MPI_BCAST(id, 128, MPI_CHARACTER, 0, comm, ierr)
    nprocdir = 1
    crd = 0
    nprocdir(1) = nprocx
    crd(1) = x_coords
#if EPOCH_NDIMS >= 2
    nprocdir(2) = nprocy
    crd(2) = y_coords
#endif
#if EPOCH_NDIMS >= 3
    nprocdir(3) = nprocz
    crd(3) = z_coords
#endif
    ! one rank per GPU: the device follows the rank's position among the ranks of ITS node, whatever the rank
    ! layout and however many GPUs the node has.  device = -(local_rank + 1) asks the library for
    ! MOD(local_rank, cudaGetDeviceCount) (include/epoch_b200.h, eb200_create)
    CALL MPI_COMM_SPLIT_TYPE(comm, MPI_COMM_TYPE_SHARED, rank, MPI_INFO_NULL, node_comm, ierr)
    CALL MPI_COMM_RANK(node_comm, local_rank, ierr)
    CALL MPI_COMM_FREE(node_comm, ierr)
    CALL gpu_check(eb200_create(gpu, c_ndims, gpu_shape, INT(-(local_rank + 1), C_INT32_T), rank, nproc, nprocdir, crd, &
        C_LOC(id)))
    CALL gpu_set_grid
    CALL gpu_check(eb200_set_field_bcs(gpu, INT(bc_field, C_INT32_T)))
    IF (cpml_boundaries) CALL gpu_set_cpml
    CALL gpu_check(eb200_set_solver(gpu, field_order, maxwell_solver, &
        [alphax, alphay, alphaz, betaxy, betaxz, betayx, betayz, betazx, betazy, &
         gammax, gammay, gammaz, deltax, deltay, deltaz]))
    ! current_finish options; initial_jx,jy,jz are valid after calc_initial_current (epoch3d.F90:140)
    IF (ALLOCATED(smooth_strides)) THEN
      CALL gpu_check(eb200_set_current_options(gpu, MERGE(1, 0, smooth_currents), smooth_its, smooth_comp_its, &
          SIZE(smooth_strides), INT(smooth_strides, C_INT32_T), MERGE(1, 0, use_current_correction), &
          [initial_jx, initial_jy, initial_jz]))
    ELSE
      CALL gpu_check(eb200_set_current_options(gpu, MERGE(1, 0, smooth_currents), smooth_its, smooth_comp_its, &
          0, [1_C_INT32_T], MERGE(1, 0, use_current_correction), [initial_jx, initial_jy, initial_jz]))
    END IF
    CALL gpu_resend_lasers
    DO ispecies = 1, n_species
      bcp = 0
      bcp(1:2*c_ndims) = INT(species_list(ispecies)%bc_particle(1:2*c_ndims), C_INT32_T)
#ifdef PER_SPECIES_WEIGHT
      species_weight = species_list(ispecies)%weight
#else
      species_weight = 0.0_num        ! the weight travels with every particle (shared_data.F90:91-137)
#endif
      CALL gpu_check(eb200_set_species(gpu, ispecies - 1, species_list(ispecies)%charge, &
          species_list(ispecies)%mass, species_weight, &
          MERGE(1, 0, species_list(ispecies)%immobile), MERGE(1, 0, species_list(ispecies)%zero_current), &
          MERGE(1, 0, species_list(ispecies)%species_type == c_species_id_photon), bcp))
      CALL gpu_send_thermal(ispecies)
    END DO
    device_stale = .TRUE.
  END SUBROUTINE gpu_attach

  SUBROUTINE gpu_send_thermal(ispecies)
    ! ext_temp_<face> of a species with thermal particle boundaries (mpi_routines.F90:561-584; filled by
    ! helper.F90:42-57); again after balance_workload re-sliced them (balance.F90:1170-1260).  The device draws
    ! from its own counter-based generator, keyed here by the rank so that ranks differ.
    INTEGER, INTENT(IN) :: ispecies
    INTEGER(C_INT64_T) :: seed
    seed = 7842432_C_INT64_T + INT(rank, C_INT64_T)
    IF (species_list(ispecies)%bc_particle(c_bd_x_min) == c_bc_thermal) CALL gpu_check(eb200_set_thermal_boundary( &
        gpu, ispecies - 1, c_bd_x_min, species_list(ispecies)%ext_temp_x_min, seed))
    IF (species_list(ispecies)%bc_particle(c_bd_x_max) == c_bc_thermal) CALL gpu_check(eb200_set_thermal_boundary( &
        gpu, ispecies - 1, c_bd_x_max, species_list(ispecies)%ext_temp_x_max, seed))
#if EPOCH_NDIMS >= 2
    IF (species_list(ispecies)%bc_particle(c_bd_y_min) == c_bc_thermal) CALL gpu_check(eb200_set_thermal_boundary( &
        gpu, ispecies - 1, c_bd_y_min, species_list(ispecies)%ext_temp_y_min, seed))
    IF (species_list(ispecies)%bc_particle(c_bd_y_max) == c_bc_thermal) CALL gpu_check(eb200_set_thermal_boundary( &
        gpu, ispecies - 1, c_bd_y_max, species_list(ispecies)%ext_temp_y_max, seed))
#endif
#if EPOCH_NDIMS >= 3
    IF (species_list(ispecies)%bc_particle(c_bd_z_min) == c_bc_thermal) CALL gpu_check(eb200_set_thermal_boundary( &
        gpu, ispecies - 1, c_bd_z_min, species_list(ispecies)%ext_temp_z_min, seed))
    IF (species_list(ispecies)%bc_particle(c_bd_z_max) == c_bc_thermal) CALL gpu_check(eb200_set_thermal_boundary( &
        gpu, ispecies - 1, c_bd_z_max, species_list(ispecies)%ext_temp_z_max, seed))
#endif
  END SUBROUTINE gpu_send_thermal

  SUBROUTINE gpu_set_grid
    ! call again after balance_workload / moving_window changed the local grid
    CALL gpu_check(eb200_set_grid(gpu, I3(nx, ny, nz), I3(nx_global, ny_global, nz_global), &
        I3(nx_global_min, ny_global_min, nz_global_min), R3(dx, dy, dz, 1.0_num), &
        R3(x_grid_min_local, y_grid_min_local, z_grid_min_local, 0.0_num), &
        R3(x_min_local, y_min_local, z_min_local, 0.0_num), R3(x_max_local, y_max_local, z_max_local, 0.0_num), &
        R3(x_min, y_min, z_min, 0.0_num), R3(x_max, y_max, z_max, 0.0_num), cpml_thickness))
  END SUBROUTINE gpu_set_grid

  SUBROUTINE gpu_set_cpml
    ! after set_cpml_helpers (boundary.F90:1989): profiles, slab ranges and laser planes of every axis;
    ! call again whenever set_cpml_helpers ran again (balance_workload)
    INTEGER :: lmin, lmax
    lmin = -1; lmax = -1
    IF (add_laser(c_bd_x_min)) lmin = cpml_x_min_laser_idx
    IF (add_laser(c_bd_x_max)) lmax = cpml_x_max_laser_idx
    CALL gpu_check(eb200_set_cpml(gpu, 0, cpml_kappa_ex, cpml_kappa_bx, cpml_sigma_ex, cpml_sigma_bx, &
        cpml_a_ex, cpml_a_bx, cpml_x_min_start, cpml_x_min_end, cpml_x_max_start, cpml_x_max_end, lmin, lmax))
#if EPOCH_NDIMS >= 2
    lmin = -1; lmax = -1
    IF (add_laser(c_bd_y_min)) lmin = cpml_y_min_laser_idx
    IF (add_laser(c_bd_y_max)) lmax = cpml_y_max_laser_idx
    CALL gpu_check(eb200_set_cpml(gpu, 1, cpml_kappa_ey, cpml_kappa_by, cpml_sigma_ey, cpml_sigma_by, &
        cpml_a_ey, cpml_a_by, cpml_y_min_start, cpml_y_min_end, cpml_y_max_start, cpml_y_max_end, lmin, lmax))
#endif
#if EPOCH_NDIMS >= 3
    lmin = -1; lmax = -1
    IF (add_laser(c_bd_z_min)) lmin = cpml_z_min_laser_idx
    IF (add_laser(c_bd_z_max)) lmax = cpml_z_max_laser_idx
    CALL gpu_check(eb200_set_cpml(gpu, 2, cpml_kappa_ez, cpml_kappa_bz, cpml_sigma_ez, cpml_sigma_bz, &
        cpml_a_ez, cpml_a_bz, cpml_z_min_start, cpml_z_min_end, cpml_z_max_start, cpml_z_max_end, lmin, lmax))
#endif
  END SUBROUTINE gpu_set_cpml

  SUBROUTINE gpu_sync_to_device
    ! host lists/arrays -> device (start of run, after restart_data, balance, window, injectors)
    INTEGER :: ispecies
    INTEGER(KIND=8) :: n, i
    REAL(num), ALLOCATABLE, TARGET :: buf(:)
    TYPE(particle), POINTER :: cur
    IF (.NOT. device_stale) RETURN
    CALL gpu_check(eb200_upload_fields(gpu, C_LOC(ex), C_LOC(ey), C_LOC(ez), C_LOC(bx), C_LOC(by), C_LOC(bz), &
        C_LOC(jx), C_LOC(jy), C_LOC(jz)))
    DO ispecies = 1, n_species
      n = species_list(ispecies)%attached_list%count
      ALLOCATE(buf(MAX(n, 1_8) * nvar))
      cur => species_list(ispecies)%attached_list%head
      DO i = 1, n
        CALL pack_particle(buf((i - 1) * nvar + 1 : i * nvar), cur)      ! partlist.F90:411
        cur => cur%next
      END DO
      CALL gpu_check(eb200_upload_particles(gpu, ispecies - 1, n, buf, nvar))
      DEALLOCATE(buf)
    END DO
    device_stale = .FALSE.
    host_stale = .FALSE.
  END SUBROUTINE gpu_sync_to_device

  SUBROUTINE gpu_sync_to_host
    ! device -> host lists/arrays (before a dump, balance_workload, moving_window, finalise)
    INTEGER :: ispecies
    INTEGER(KIND=8) :: n, got, i
    REAL(num), ALLOCATABLE, TARGET :: buf(:)
    TYPE(particle), POINTER :: cur
    IF (.NOT. host_stale) RETURN
    CALL gpu_check(eb200_download_fields(gpu, C_LOC(ex), C_LOC(ey), C_LOC(ez), C_LOC(bx), C_LOC(by), C_LOC(bz), &
        C_LOC(jx), C_LOC(jy), C_LOC(jz)))
    DO ispecies = 1, n_species
      CALL gpu_check(eb200_particle_count(gpu, ispecies - 1, n))
      ALLOCATE(buf(MAX(n, 1_8) * nvar))
      CALL gpu_check(eb200_download_particles(gpu, ispecies - 1, 0_8, n, buf, nvar, got))
      CALL destroy_partlist(species_list(ispecies)%attached_list)
      CALL create_empty_partlist(species_list(ispecies)%attached_list)
      DO i = 1, got
        CALL create_particle(cur)
        CALL unpack_particle(buf((i - 1) * nvar + 1 : i * nvar), cur)    ! partlist.F90:477
        CALL add_particle_to_partlist(species_list(ispecies)%attached_list, cur)
      END DO
      DEALLOCATE(buf)
    END DO
    host_stale = .FALSE.
  END SUBROUTINE gpu_sync_to_host

  SUBROUTINE gpu_resend_lasers
    ! TYPE(laser_block) amplitudes and the profile / phase planes on this rank's boundary
    ! (laser.f90:232-330); again after the planes were re-sliced by balance_workload
    TYPE(laser_block), POINTER :: current
    INTEGER :: ib
    DO ib = 1, 2 * c_ndims
      current => laser_boundary_head(ib)
      DO WHILE (ASSOCIATED(current))
        CALL gpu_check(eb200_set_laser(gpu, ib, current%id, current%amp, current%pol_angle, current%t_start, &
            current%t_end, C_LOC(current%profile), C_LOC(current%phase)))
        current => current%next
      END DO
    END DO
  END SUBROUTINE gpu_resend_lasers

  SUBROUTINE gpu_laser_step
    ! the host keeps evaluating the deck's time profiles (laser.f90:207-226,334-346)
    TYPE(laser_block), POINTER :: current
    INTEGER :: ib
    CALL update_laser_omegas
    DO ib = 1, 2 * c_ndims
      current => laser_boundary_head(ib)       ! laser_x_min, laser_x_max, ... in the reference
      DO WHILE (ASSOCIATED(current))
        CALL gpu_check(eb200_set_laser_step(gpu, ib, current%id, laser_time_profile(current), &
            current%current_integral_phase))
        current => current%next
      END DO
    END DO
  END SUBROUTINE gpu_laser_step

  ! ---- the four hot entries, reference names and (absent) argument lists --------

  SUBROUTINE update_eb_fields_half          ! fields.f90:819
    CALL gpu_sync_to_device
    CALL gpu_check(eb200_set_time(gpu, dt, time, step))
    CALL gpu_check(eb200_update_eb_fields_half(gpu))
    host_stale = .TRUE.
  END SUBROUTINE update_eb_fields_half

  SUBROUTINE push_particles                 ! particles.F90:28
    CALL gpu_check(eb200_push_particles(gpu))
    host_stale = .TRUE.
  END SUBROUTINE push_particles

  SUBROUTINE current_finish                 ! housekeeping/current_smooth.F90:29
    CALL gpu_check(eb200_current_finish(gpu))
  END SUBROUTINE current_finish

  ! Device half of shift_window (window.F90:62-94) for one cell.  In window.F90 the loop body becomes
  !   CALL insert_particles_to_gpu      ! the reference's insert_particles building a partlist that is
  !                                     ! packed (pack_particle) and passed to eb200_append_particles
  !   <shift x_grid_min, xb_min, x_min, x_max; CALL setup_grid_x>        (unchanged host code)
  !   CALL gpu_shift_window
  ! and remove_particles / shift_fields are no longer called on the host.
  SUBROUTINE gpu_shift_window
    CALL gpu_check(eb200_shift_window(gpu, x_grid_min_local, x_min_local, x_max_local, x_min, x_max))
  END SUBROUTINE gpu_shift_window

  ! ---- load balancing (housekeeping/balance.F90) ----
  ! part_load_func (:2468): the device counts this rank's particles per global cell; get_load's
  ! MPI_ALLREDUCE and push_per_field weighting (:2446-2460) stay as they are.
#if EPOCH_NDIMS == 3
  SUBROUTINE part_load_func(load_x, load_y, load_z)
    INTEGER(i8), DIMENSION(1-ng:), INTENT(OUT) :: load_x, load_y, load_z
    CALL gpu_check(eb200_get_load(gpu, [x_grid_min, y_grid_min, z_grid_min], load_x, load_y, load_z))
  END SUBROUTINE part_load_func
#elif EPOCH_NDIMS == 2
  SUBROUTINE part_load_func(load_x, load_y)       ! epoch2d/src/housekeeping/balance.F90: two histograms
    INTEGER(i8), DIMENSION(1-ng:), INTENT(OUT) :: load_x, load_y
    INTEGER(i8), TARGET :: unused(1)
    CALL gpu_check(eb200_get_load(gpu, [x_grid_min, y_grid_min, 0.0_num], load_x, load_y, unused))
  END SUBROUTINE part_load_func
#else
  SUBROUTINE part_load_func(load_x)               ! epoch1d: one histogram
    INTEGER(i8), DIMENSION(1-ng:), INTENT(OUT) :: load_x
    INTEGER(i8), TARGET :: unused(1)
    CALL gpu_check(eb200_get_load(gpu, [x_grid_min, 0.0_num, 0.0_num], load_x, unused, unused))
  END SUBROUTINE part_load_func
#endif

  ! redistribute_domain (:411-472): called by balance_workload after calculate_breaks filled
  ! new_cell_*_min/max and after the new local grid scalars were set (setup_grid_x etc. with the
  ! new nx_global_min); fields, psi arrays and particles move on the device, rank to rank.
  ! (3D argument list; the 2D / 1D callers pass 0 / 1 for the axes they do not have, as I3 / R3 do above)
#if EPOCH_NDIMS == 3
  SUBROUTINE gpu_redistribute_domain(x_grid_min_local_new, y_grid_min_local_new, z_grid_min_local_new, &
      mins_new, maxs_new)
    REAL(num), INTENT(IN) :: x_grid_min_local_new, y_grid_min_local_new, z_grid_min_local_new
    REAL(num), INTENT(IN) :: mins_new(3), maxs_new(3)
    CALL gpu_check(eb200_redistribute(gpu, &
        INT([cell_x_min, cell_y_min, cell_z_min], C_INT32_T), INT([cell_x_max, cell_y_max, cell_z_max], C_INT32_T), &
        INT([new_cell_x_min, new_cell_y_min, new_cell_z_min], C_INT32_T), &
        INT([new_cell_x_max, new_cell_y_max, new_cell_z_max], C_INT32_T), &
        [x_grid_min, y_grid_min, z_grid_min], &
        [x_grid_min_local_new, y_grid_min_local_new, z_grid_min_local_new], mins_new, maxs_new))
    ! the host keeps redistributing its own small arrays (laser profiles, window planes, CPML
    ! profiles via set_cpml_helpers) and hands them over again
    IF (cpml_boundaries) CALL gpu_set_cpml
    CALL gpu_resend_lasers
    host_stale = .TRUE.
  END SUBROUTINE gpu_redistribute_domain
#endif

  ! ---- derived grid variables for output_routines (io/calc_df.F90) ----
  ! e.g. calc_number_density(array, ispecies) becomes
  !   CALL gpu_calc_derived(EB200_DF_NUMBER_DENSITY, array, ispecies)
  SUBROUTINE gpu_calc_derived(kind, data_array, current_species, direction)
    INTEGER, INTENT(IN) :: kind, current_species
    REAL(num), GHOST_DIMS, INTENT(OUT) :: data_array
    INTEGER, INTENT(IN), OPTIONAL :: direction
    INTEGER :: dir
    dir = 0
    IF (PRESENT(direction)) dir = direction
    ! current_species <= 0 means the sum over species (-1 in the C ABI); species are 0-based there
    CALL gpu_check(eb200_calc_derived(gpu, kind, MAX(current_species, 0) - 1, dir, data_array))
  END SUBROUTINE gpu_calc_derived

  SUBROUTINE update_eb_fields_final         ! fields.f90:851
    CALL gpu_check(eb200_set_time(gpu, dt, time, step))
    CALL gpu_laser_step
    CALL gpu_check(eb200_update_eb_fields_final(gpu))
    host_stale = .TRUE.
  END SUBROUTINE update_eb_fields_final

END MODULE epoch_b200_hooks
