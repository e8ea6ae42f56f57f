! epoch_b200_iface.F90 -- ISO_C_BINDING interfaces of include/epoch_b200.h.
! One-to-one with the C declarations; the wrappers that carry the reference's
! subroutine names live in epoch_b200_hooks.F90.  Pattern precedent in the
! reference: housekeeping/terminal_controls.F90:22-29 (BIND(C, NAME='isatty')).
! Compiled only where a Fortran compiler exists (none in the authoring image).
MODULE epoch_b200_iface

  USE, INTRINSIC :: ISO_C_BINDING
  IMPLICIT NONE

  INTERFACE
    INTEGER(C_INT32_T) FUNCTION eb200_create(ctx, ndims, shape, device, rank, nproc, &
        nprocdir, coords, nccl_id) BIND(C, NAME='eb200_create')
      IMPORT
      TYPE(C_PTR), INTENT(OUT) :: ctx
      INTEGER(C_INT32_T), VALUE :: ndims, shape, device, rank, nproc
      INTEGER(C_INT32_T), INTENT(IN) :: nprocdir(3), coords(3)
      TYPE(C_PTR), VALUE :: nccl_id
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_host_alloc(out, bytes) BIND(C, NAME='eb200_host_alloc')
      IMPORT :: C_INT32_T, C_INT64_T, C_PTR
      TYPE(C_PTR), INTENT(OUT) :: out
      INTEGER(C_INT64_T), VALUE :: bytes
    END FUNCTION eb200_host_alloc
    INTEGER(C_INT32_T) FUNCTION eb200_host_free(p) BIND(C, NAME='eb200_host_free')
      IMPORT :: C_INT32_T, C_PTR
      TYPE(C_PTR), VALUE :: p
    END FUNCTION eb200_host_free
    INTEGER(C_INT32_T) FUNCTION eb200_set_thermal_boundary(ctx, ispecies, boundary, ext_temp, seed) &
        BIND(C, NAME='eb200_set_thermal_boundary')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: ispecies, boundary
      REAL(C_DOUBLE), INTENT(IN) :: ext_temp(*)
      INTEGER(C_INT64_T), VALUE :: seed
    END FUNCTION eb200_set_thermal_boundary
    INTEGER(C_INT32_T) FUNCTION eb200_set_probe(ctx, ispecies, iprobe, point, normal, ek_min, ek_max) &
        BIND(C, NAME='eb200_set_probe')
      IMPORT :: C_INT32_T, C_DOUBLE, C_PTR
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: ispecies, iprobe
      REAL(C_DOUBLE), INTENT(IN) :: point(*), normal(*)
      REAL(C_DOUBLE), VALUE :: ek_min, ek_max
    END FUNCTION eb200_set_probe
    INTEGER(C_INT32_T) FUNCTION eb200_download_probe(ctx, ispecies, iprobe, capacity, packed, nvar, n_out, clear) &
        BIND(C, NAME='eb200_download_probe')
      IMPORT :: C_INT32_T, C_INT64_T, C_PTR
      TYPE(C_PTR), VALUE :: ctx, packed
      INTEGER(C_INT32_T), VALUE :: ispecies, iprobe, nvar, clear
      INTEGER(C_INT64_T), VALUE :: capacity
      INTEGER(C_INT64_T), INTENT(OUT) :: n_out
    END FUNCTION eb200_download_probe
    INTEGER(C_INT32_T) FUNCTION eb200_dist_fn(ctx, ispecies, curdims, direction, ranges, resolution, dgrid, &
        restrictions, use_restrictions, array) BIND(C, NAME='eb200_dist_fn')
      IMPORT :: C_INT32_T, C_DOUBLE, C_PTR
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: ispecies, curdims
      INTEGER(C_INT32_T), INTENT(IN) :: direction(3), resolution(3)
      REAL(C_DOUBLE), INTENT(IN) :: ranges(2,3), dgrid(3)
      TYPE(C_PTR), VALUE :: restrictions, use_restrictions   ! (2,c_df_maxdirs) reals / c_df_maxdirs int32, or C_NULL_PTR
      REAL(C_DOUBLE), INTENT(OUT) :: array(*)
    END FUNCTION eb200_dist_fn
    INTEGER(C_INT32_T) FUNCTION eb200_nccl_unique_id(out128) BIND(C, NAME='eb200_nccl_unique_id')
      IMPORT
      CHARACTER(KIND=C_CHAR), INTENT(OUT) :: out128(128)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_destroy(ctx) BIND(C, NAME='eb200_destroy')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_grid(ctx, n_local, n_global, n_global_min, d, &
        grid_min_local, min_local, max_local, min_global, max_global, cpml_thickness) &
        BIND(C, NAME='eb200_set_grid')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), INTENT(IN) :: n_local(3), n_global(3), n_global_min(3)
      REAL(C_DOUBLE), INTENT(IN) :: d(3), grid_min_local(3), min_local(3), max_local(3)
      REAL(C_DOUBLE), INTENT(IN) :: min_global(3), max_global(3)
      INTEGER(C_INT32_T), VALUE :: cpml_thickness
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_time(ctx, dt, time, step) BIND(C, NAME='eb200_set_time')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      REAL(C_DOUBLE), VALUE :: dt, time
      INTEGER(C_INT32_T), VALUE :: step
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_field_bcs(ctx, bc_field) BIND(C, NAME='eb200_set_field_bcs')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), INTENT(IN) :: bc_field(6)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_solver(ctx, field_order, maxwell_solver, coeffs) &
        BIND(C, NAME='eb200_set_solver')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: field_order, maxwell_solver
      REAL(C_DOUBLE), INTENT(IN) :: coeffs(15)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_current_options(ctx, smooth, smooth_its, smooth_comp_its, &
        nstrides, strides, use_current_correction, initial_j) BIND(C, NAME='eb200_set_current_options')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: smooth, smooth_its, smooth_comp_its, nstrides, use_current_correction
      INTEGER(C_INT32_T), INTENT(IN) :: strides(*)
      REAL(C_DOUBLE), INTENT(IN) :: initial_j(3)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_cpml(ctx, axis, kappa_e, kappa_b, sigma_e, sigma_b, a_e, a_b, &
        min_start, min_end, max_start, max_end, laser_idx_min, laser_idx_max) BIND(C, NAME='eb200_set_cpml')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: axis, min_start, min_end, max_start, max_end, laser_idx_min, laser_idx_max
      REAL(C_DOUBLE), INTENT(IN) :: kappa_e(*), kappa_b(*), sigma_e(*), sigma_b(*), a_e(*), a_b(*)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_get_load(ctx, grid_min_global, load_x, load_y, load_z) &
        BIND(C, NAME='eb200_get_load')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      REAL(C_DOUBLE), INTENT(IN) :: grid_min_global(3)
      INTEGER(C_INT64_T), INTENT(OUT) :: load_x(*), load_y(*), load_z(*)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_redistribute(ctx, old_cell_min, old_cell_max, new_cell_min, new_cell_max, &
        grid_min_global, grid_min_local, min_local, max_local) BIND(C, NAME='eb200_redistribute')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), INTENT(IN) :: old_cell_min(*), old_cell_max(*), new_cell_min(*), new_cell_max(*)
      REAL(C_DOUBLE), INTENT(IN) :: grid_min_global(3), grid_min_local(3), min_local(3), max_local(3)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_calc_derived(ctx, kind, ispecies, direction, data_array) &
        BIND(C, NAME='eb200_calc_derived')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: kind, ispecies, direction
      REAL(C_DOUBLE), INTENT(OUT) :: data_array(*)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_species(ctx, ispecies, charge, mass, weight, immobile, &
        zero_current, species_type, bc_particle) BIND(C, NAME='eb200_set_species')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: ispecies, immobile, zero_current, species_type
      REAL(C_DOUBLE), VALUE :: charge, mass, weight
      INTEGER(C_INT32_T), INTENT(IN) :: bc_particle(6)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_upload_particles(ctx, ispecies, n, packed, nvar) &
        BIND(C, NAME='eb200_upload_particles')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: ispecies, nvar
      INTEGER(C_INT64_T), VALUE :: n
      REAL(C_DOUBLE), INTENT(IN) :: packed(*)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_append_particles(ctx, ispecies, n, packed, nvar) &
        BIND(C, NAME='eb200_append_particles')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: ispecies, nvar
      INTEGER(C_INT64_T), VALUE :: n
      REAL(C_DOUBLE), INTENT(IN) :: packed(*)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_particle_count(ctx, ispecies, n) BIND(C, NAME='eb200_particle_count')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: ispecies
      INTEGER(C_INT64_T), INTENT(OUT) :: n
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_download_particles(ctx, ispecies, first, capacity, packed, nvar, n_out) &
        BIND(C, NAME='eb200_download_particles')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: ispecies, nvar
      INTEGER(C_INT64_T), VALUE :: first, capacity
      REAL(C_DOUBLE), INTENT(OUT) :: packed(*)
      INTEGER(C_INT64_T), INTENT(OUT) :: n_out
    END FUNCTION
    ! fields: pass C_LOC(array) or C_NULL_PTR to skip a component
    INTEGER(C_INT32_T) FUNCTION eb200_upload_fields(ctx, ex, ey, ez, bx, by, bz, jx, jy, jz) &
        BIND(C, NAME='eb200_upload_fields')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx, ex, ey, ez, bx, by, bz, jx, jy, jz
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_download_fields(ctx, ex, ey, ez, bx, by, bz, jx, jy, jz) &
        BIND(C, NAME='eb200_download_fields')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx, ex, ey, ez, bx, by, bz, jx, jy, jz
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_laser(ctx, boundary, id, amp, pol_angle, t_start, t_end, &
        profile, phase) BIND(C, NAME='eb200_set_laser')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx, profile, phase
      INTEGER(C_INT32_T), VALUE :: boundary, id
      REAL(C_DOUBLE), VALUE :: amp, pol_angle, t_start, t_end
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_laser_step(ctx, boundary, id, t_env, integral_phase) &
        BIND(C, NAME='eb200_set_laser_step')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      INTEGER(C_INT32_T), VALUE :: boundary, id
      REAL(C_DOUBLE), VALUE :: t_env, integral_phase
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_update_eb_fields_half(ctx) BIND(C, NAME='eb200_update_eb_fields_half')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_push_particles(ctx) BIND(C, NAME='eb200_push_particles')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_current_finish(ctx) BIND(C, NAME='eb200_current_finish')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_update_eb_fields_final(ctx) BIND(C, NAME='eb200_update_eb_fields_final')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_shift_window(ctx, x_grid_min_local, x_min_local, x_max_local, x_min, x_max) &
        BIND(C, NAME='eb200_shift_window')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      REAL(C_DOUBLE), VALUE :: x_grid_min_local, x_min_local, x_max_local, x_min, x_max
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_window_fields(ctx, ex_x_max, ey_x_max, ez_x_max, bx_x_max, by_x_max, &
        bz_x_max) BIND(C, NAME='eb200_set_window_fields')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      REAL(C_DOUBLE), INTENT(IN) :: ex_x_max(*), ey_x_max(*), ez_x_max(*), bx_x_max(*), by_x_max(*), bz_x_max(*)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_particle_bcs(ctx) BIND(C, NAME='eb200_particle_bcs')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_efield_bcs(ctx) BIND(C, NAME='eb200_efield_bcs')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_bfield_final_bcs(ctx) BIND(C, NAME='eb200_bfield_final_bcs')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_energy(ctx, field_energy, particle_energy) BIND(C, NAME='eb200_energy')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      REAL(C_DOUBLE), INTENT(OUT) :: field_energy, particle_energy(*)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_charge_residual(ctx, max_res, max_rho) BIND(C, NAME='eb200_charge_residual')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      REAL(C_DOUBLE), INTENT(OUT) :: max_res, max_rho
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_set_option(ctx, key, val) BIND(C, NAME='eb200_set_option')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      CHARACTER(KIND=C_CHAR), INTENT(IN) :: key(*), val(*)
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_sync(ctx) BIND(C, NAME='eb200_sync')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_mark(ctx, phase, begin) BIND(C, NAME='eb200_mark')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      CHARACTER(KIND=C_CHAR), INTENT(IN) :: phase(*)
      INTEGER(C_INT32_T), VALUE :: begin
    END FUNCTION
    INTEGER(C_INT32_T) FUNCTION eb200_timer(ctx, phase, reset, ms, launches) BIND(C, NAME='eb200_timer')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
      CHARACTER(KIND=C_CHAR), INTENT(IN) :: phase(*)
      INTEGER(C_INT32_T), VALUE :: reset
      REAL(C_DOUBLE), INTENT(OUT) :: ms
      INTEGER(C_INT64_T), INTENT(OUT) :: launches
    END FUNCTION
    TYPE(C_PTR) FUNCTION eb200_last_error(ctx) BIND(C, NAME='eb200_last_error')
      IMPORT
      TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    TYPE(C_PTR) FUNCTION eb200_version() BIND(C, NAME='eb200_version')
      IMPORT
    END FUNCTION
  END INTERFACE

END MODULE epoch_b200_iface
