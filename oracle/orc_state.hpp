// oracle/orc_state.hpp -- CPU ORACLE (TEST INFRASTRUCTURE ONLY; never linked into
// or called by the product library).
//
// Plain restatement, in C++ with fp64 and no FMA contraction, of the EPOCH
// per-timestep hot path.  The reference itself (Fortran 90 + MPI + missing SDF
// submodule) cannot be built in this container (no Fortran compiler), so this
// file states the same arithmetic, citing the reference lines it follows.
//
// Parity pinning: the field solver + boundary + laser part is pinned against the
// reference's own golden sums (epoch{1,2,3}d/tests/test_laser.py).  The particle
// push / deposition has NO golden vectors in the reference ("parity unpinned");
// it is validated by analytic invariants (exact charge conservation etc.).
#pragma once
#include <cstdint>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

namespace orc {

// constants.F90:144-151 (CODATA-2010 era literals -- keep these, not modern ones)
constexpr double PI = 3.141592653589793238462643383279503;
constexpr double Q0 = 1.602176565e-19;
constexpr double M0 = 9.10938291e-31;
constexpr double C = 2.99792458e8;
constexpr double KB = 1.3806488e-23;
constexpr double C_TINY = 2.2250738585072014e-308;   // TINY(1.0_num), constants.F90:29
constexpr double MU0 = 4.e-7 * PI;
constexpr double EPS0 = 8.854187817620389850536563031710750e-12;

// constants.F90:56-71
enum { BC_PERIODIC = 1, BC_OTHER = 2, BC_SIMPLE_LASER = 3, BC_SIMPLE_OUTFLOW = 4, BC_OPEN = 5,
       BC_DUMP = 6, BC_ZERO_GRADIENT = 7, BC_CLAMP = 8, BC_REFLECT = 9, BC_CONDUCT = 10,
       BC_THERMAL = 11, BC_CPML_LASER = 12, BC_CPML_OUTFLOW = 13, BC_MIXED = 14, BC_HEAT_BATH = 15 };

enum { EX = 0, EY, EZ, BX, BY, BZ, JX, JY, JZ, NFIELD };

struct Laser {
  int id = 0;
  double amp = 0, pol = 0, t_start = 0, t_end = 0;
  double t_env = 0, integral_phase = 0;     // set each step by the host
  std::vector<double> profile, phase;       // planes over 0..n of tangential dims
};

struct Species {
  bool defined = false;
  double charge = 0, mass = 0, weight = 0;
  int immobile = 0, zero_current = 0, species_type = 0;
  int bc[6] = {BC_PERIODIC, BC_PERIODIC, BC_PERIODIC, BC_PERIODIC, BC_PERIODIC, BC_PERIODIC};
  // SoA: pos[d], p[3], w
  std::vector<double> pos[3], p[3], w;
  // particle probes (shared_data.F90:392-404) and their sampled particles (packed pos, p(3), weight)
  struct Probe { bool on = false; double point[3] = {0, 0, 0}, normal[3] = {0, 0, 0}, ek_min = 0, ek_max = 0; std::vector<double> sampled; };
  Probe probes[8];
  // thermal particle boundaries: ext_temp_{x,y,z}_{min,max}(transverse extents incl. ghosts, 3), mpi_routines.F90:561-584
  std::vector<double> ext_temp[6];
  size_t count() const { return w.size(); }
};

// global_random of the reference (orc_loader.cpp: KISS + polar Box-Muller, random_generator.f90:45-173)
double global_random();
double global_random_box_muller(double stdev, double mu);

struct State {
  int ndims = 3, shape = 1;
  int rank = 0, nproc = 1, nprocdir[3] = {1, 1, 1}, coords[3] = {0, 0, 0};
  // shape constants (constants.F90:472-492)
  int sf_min = -1, sf_max = 1, png = 3, ng = 5;
  // grid
  bool grid_set = false;
  int n[3] = {1, 1, 1};          // local cells
  int nglob[3] = {1, 1, 1};
  int nglob_min[3] = {1, 1, 1};
  int ngd[3] = {0, 0, 0};        // ghost width per dim (ng or 0 for missing dims)
  int64_t st[3] = {1, 1, 1};     // strides
  int64_t total = 1;
  double d[3] = {1, 1, 1};
  double grid_min_local[3] = {0, 0, 0};
  double min_local[3] = {0, 0, 0}, max_local[3] = {0, 0, 0};
  double gmin[3] = {0, 0, 0}, gmax[3] = {0, 0, 0};
  int cpml_thickness = 0;
  bool min_bnd[3] = {true, true, true}, max_bnd[3] = {true, true, true};
  // time
  double dt = 0, time = 0;
  int step = 0;
  // field BCs (after setup_boundaries normalisation) and solver
  int bc_field[6] = {BC_PERIODIC, BC_PERIODIC, BC_PERIODIC, BC_PERIODIC, BC_PERIODIC, BC_PERIODIC};
  bool add_laser[6] = {false, false, false, false, false, false};
  int field_order = 2, maxwell_solver = 0;
  double coeffs[15] = {0};
  // current_finish options (housekeeping/current_smooth.F90)
  bool smooth = false, use_current_correction = false;
  int smooth_its = 1, smooth_comp_its = 0;
  std::vector<int> smooth_strides;
  double initial_j[3] = {0, 0, 0};
  // CPML (set_cpml_helpers): per axis kappa_e, kappa_b, sigma_e, sigma_b, a_e, a_b over (1-ng : n+ng),
  // slab ranges per boundary, laser plane index per boundary, psi[axis][0..3] = E(next), E(next-next),
  // B(next), B(next-next)
  bool cpml_on = false;
  std::vector<double> cpml_prof[3][6];
  int cpml_start[6] = {0, 0, 0, 0, 0, 0}, cpml_end[6] = {-1, -1, -1, -1, -1, -1};
  int cpml_laser_idx[6] = {1, 1, 1, 1, 1, 1};
  std::vector<double> cpml_psi[3][4];
  std::vector<double> win_planes[6];   // moving window: e*_x_max, b*_x_max (setup.F90:425-470)
  // setup_field_boundaries (setup.F90:426-510): per boundary the planes ex..bz over the other axes incl. ghosts
  // (x fastest in (lower axis, higher axis) order); empty = zero
  std::vector<double> bnd_planes[6][6];
  // half-step scalars (fields.f90:821-830)
  double hdt = 0, hdtd[3] = {0, 0, 0}, cn[3] = {0, 0, 0}, fac = 0;
  std::vector<double> f[NFIELD];
  std::vector<Species> species;
  std::vector<Laser> lasers[6];
  std::string err;
  std::vector<double> res_ref;   // captured div E - rho/eps0 (option capture_charge_reference)

  // linear index of Fortran (i,j,k) (1-based interior); missing dims use index 1
  inline int64_t idx(int i, int j, int k) const {
    return (int64_t)(i + ngd[0] - 1) * st[0] + (int64_t)(j + ngd[1] - 1) * st[1] +
           (int64_t)(k + ngd[2] - 1) * st[2];
  }
  // loop bounds "0..n" of the reference for an existing dim, "1..1" otherwise
  inline int lo0(int dim) const { return dim < ndims ? 0 : 1; }
  inline int hi(int dim) const { return dim < ndims ? n[dim] : 1; }
};

// orc_fields.cpp
void update_e_field(State& s);
void update_b_field(State& s);
void set_half_step_scalars(State& s);
void field_bc(State& s, int comp);                 // halo (single rank: periodic wrap)
void field_bc_array(State& s, double* a);          // the same for any array of the field shape
void field_zero_gradient_array(State& s, double* a, bool staggered, int boundary);   // boundary.F90:623-698
// bc6 != null: the species form (current_bcs(species) with mixed per-species BCs): that species' boundary codes,
// ghosts cleared afterwards (particle_clear_bcs, boundary.F90:1088-1119)
void processor_summation_array(State& s, double* a, int flip_dir, const int* bc6 = nullptr);   // boundary.F90:1123-1144
bool bc_mixed(const State& s);                                                      // ANY(bc_allspecies == c_bc_mixed)
void weights_fac(int shape, double f, double* out);                                  // include/*/gxfac.inc
void calc_derived(State& s, int kind, int ispecies, int direction, double* out);     // io/calc_df.F90
void dist_fn(const State& s, int ispecies, int curdims, const int32_t* direction, const double* ranges,
             const int32_t* resolution, const double* dgrid, const double* restrictions,
             const int32_t* use_restrictions, double* out);                            // io/dist_fn.F90
void efield_bcs(State& s);
void bfield_bcs(State& s, bool mpi_only);
void bfield_final_bcs(State& s);
void setup_field_boundaries(State& s);
void update_eb_fields_half(State& s);
void update_eb_fields_final(State& s);
void current_finish(State& s);
void cpml_advance_currents(State& s, bool efield, double tstep);
void shift_window(State& s, double grid_min_local, double min_local, double max_local, double gmin, double gmax);
// orc_particles.cpp
void push_particles(State& s);
void particle_bcs(State& s);
void energy(const State& s, double* field_energy, double* per_species);
void charge_residual(State& s, double* max_res, double* max_rho, bool capture);

}  // namespace orc
