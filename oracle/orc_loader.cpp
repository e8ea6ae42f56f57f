// oracle/orc_loader.cpp -- CPU ORACLE (TEST INFRASTRUCTURE ONLY; never linked into or called by
// the product library): the reference's initial particle load, so that "identical decks and
// initial particle loads" (BASELINE.json north_star) means the loads EPOCH itself would make.
//
//   random(), random_init(), random_box_muller():
//       epoch{1,2,3}d/src/housekeeping/random_generator.f90:45-78, :82-108, :112-173
//       (KISS: 32-bit wrap-around integers; ISHFT with a negative count is a LOGICAL shift)
//   seed: housekeeping/setup.F90:505-510 (1D numbering): seed = 7842432 + rank
//   placement: user_interaction/helper.F90 load_particles -- 1D :500-524, 3D :575-607: cells in
//       Fortran order (x fastest), npart_per_cell particles each at x(ix) + (random() - 1/2) dx,
//       then y, then z drawn in that order
//   weights: setup_particle_density -- 1D :601-700, 3D :735-832: the normalised shape weights
//       (include/*/gxfac.inc) fold the cell densities, then x dV / npart_in_cell(nearest cell)
//   momenta: user_interaction/particle_temperature.F90 setup_particle_temperature (1D :30-77,
//       3D :30-85) once per direction over the whole list, momentum_from_temperature :380-390
//       (stdev = sqrt(T kB m), mu = drift) through the polar Box-Muller with its cached second value
// Parity: unpinned -- the reference holds no vector of its generator or loader; the restatement
// is cross-checked against an independent numpy restatement (tests/test_loader.py).
//
// Restricted to what the tests and the bench need: uniform density / temperature / drift over
// the whole local grid (every cell valid), an integer number of particles per cell.
#include "orc_state.hpp"
#include "../include/epoch_b200.h"
#include <algorithm>

namespace orc {

struct Rng {
  uint32_t x = 123456789u, y = 362436069u, z = 521288629u, w = 916191069u;
  bool cached = false;
  double cached_value = 0.0;
};
static Rng g_rng;

// random_generator.f90:45-78
static double kiss(Rng& s) {
  s.x = 69069u * s.x + 1327217885u;
  const uint32_t a1 = s.y;
  const uint32_t a2 = a1 ^ (a1 << 13);
  const uint32_t a3 = a2 ^ (a2 >> 17);          // ISHFT(a2, -17): logical
  s.y = a3 ^ (a3 << 5);
  s.z = 18000u * (s.z & 65535u) + (s.z >> 16);
  s.w = 30903u * (s.w & 65535u) + (s.w >> 16);
  const uint32_t k = s.x + s.y + (s.z << 16) + s.w;
  return ((double)(int32_t)k + 2147483648.0) / 4294967296.0;
}

// random_generator.f90:82-108
static void kiss_init(Rng& s, int32_t seed) {
  s.x = 123456789u + (uint32_t)seed;
  s.y = 362436069u + (uint32_t)seed;
  s.z = 521288629u + (uint32_t)seed;
  s.w = 916191069u + (uint32_t)seed;
  s.cached = false;
  s.cached_value = 0.0;
  for (int i = 0; i < 1000; ++i) kiss(s);
}

// random_generator.f90:112-173
static double box_muller(Rng& s, double stdev, double mu) {
  if (s.cached) {
    s.cached = false;
    return s.cached_value * stdev + mu;
  }
  s.cached = true;
  double rand1, rand2, w;
  for (;;) {
    rand1 = kiss(s);
    rand2 = kiss(s);
    rand1 = 2.0 * rand1 - 1.0;
    rand2 = 2.0 * rand2 - 1.0;
    w = rand1 * rand1 + rand2 * rand2;
    if (w > C_TINY && w < 1.0) break;
  }
  w = std::sqrt((-2.0 * std::log(w)) / w);
  s.cached_value = rand2 * w;
  return rand1 * w * stdev + mu;
}

double global_random() { return kiss(g_rng); }
double global_random_box_muller(double stdev, double mu) { return box_muller(g_rng, stdev, mu); }

// particle_to_grid.inc: nearest cell (1-based) and normalised weights of one axis
static int to_grid(const State& s, int d, double pos, double* g) {
  double cr = (pos - s.grid_min_local[d]) / s.d[d];
  if (s.shape == 0) cr = cr - 0.5;
  const int cell = (int)std::floor(cr + 0.5);
  const double frac = (double)cell - cr;
  weights_fac(s.shape, frac, g);
  return cell + 1;
}

// auto_load for one species with uniform density, temperature(3) and drift(3): appends to the species
static int64_t load_uniform(State& s, int isp, int64_t ppc, double density, const double* temperature,
                            const double* drift) {
  Species& sp = s.species[isp];
  const int nd = s.ndims, nw = s.sf_max - s.sf_min + 1;
  const int64_t ncell = (int64_t)s.n[0] * s.n[1] * s.n[2];
  const int64_t np = ncell * ppc;
  const size_t first = sp.count();
  for (int d = 0; d < nd; ++d) sp.pos[d].resize(first + np);
  for (int d = 0; d < 3; ++d) sp.p[d].resize(first + np, 0.0);
  sp.w.resize(first + np, 0.0);
  // load_particles: cells in Fortran order, ppc particles each, x then y then z
  size_t ip = first;
  for (int iz = 1; iz <= s.n[2]; ++iz)
    for (int iy = 1; iy <= s.n[1]; ++iy)
      for (int ix = 1; ix <= s.n[0]; ++ix)
        for (int64_t k = 0; k < ppc; ++k, ++ip) {
          const int c[3] = {ix, iy, iz};
          for (int d = 0; d < nd; ++d) {
            // x(ix) = x_global(nx_global_min + ix - 1) = x_grid_min + (global index - 1) dx  (setup.F90:164-181)
            const double x_grid_min = (s.gmin[d] - s.d[d] * (double)s.cpml_thickness) + s.d[d] / 2.0;
            const double xc = x_grid_min + (double)(s.nglob_min[d] + c[d] - 2) * s.d[d];
            sp.pos[d][ip] = xc + (kiss(g_rng) - 0.5) * s.d[d];
          }
        }
  // setup_particle_density: weight = sum of g * density over the stencil (z outermost, x innermost) ...
  std::vector<int32_t> npart_in_cell((size_t)s.total, 0);
  double gw[3][5];
  for (ip = first; ip < first + np; ++ip) {
    int cell[3] = {1, 1, 1};
    for (int d = 0; d < nd; ++d) cell[d] = to_grid(s, d, sp.pos[d][ip], gw[d]);
    double wdata = 0.0;
    for (int k = 0; k < (nd >= 3 ? nw : 1); ++k)
      for (int j = 0; j < (nd >= 2 ? nw : 1); ++j)
        for (int i = 0; i < nw; ++i) {
          double g = gw[0][i];
          if (nd >= 2) g = g * gw[1][j];
          if (nd >= 3) g = g * gw[2][k];
          wdata = wdata + g * density;
        }
    sp.w[ip] = wdata;
    if (s.shape == 0)        // tophat: the cell that holds the particle
      for (int d = 0; d < nd; ++d) if (gw[d][1] > gw[d][0]) cell[d] += 1;
    npart_in_cell[(size_t)s.idx(cell[0], cell[1], cell[2])] += 1;
  }
  // ... renormalised by the cell volume over the particles of the nearest cell
  double dv = s.d[0];
  if (nd >= 2) dv = dv * s.d[1];
  if (nd >= 3) dv = dv * s.d[2];
  for (ip = first; ip < first + np; ++ip) {
    int cell[3] = {1, 1, 1};
    for (int d = 0; d < nd; ++d) cell[d] = (int)std::floor((sp.pos[d][ip] - s.grid_min_local[d]) / s.d[d] + 1.5);
    sp.w[ip] = sp.w[ip] * dv / (double)npart_in_cell[(size_t)s.idx(cell[0], cell[1], cell[2])];
  }
  // setup_particle_temperature: direction by direction over the whole list
  for (int dir = 0; dir < 3; ++dir)
    for (ip = first; ip < first + np; ++ip) {
      for (int d = 0; d < nd; ++d) to_grid(s, d, sp.pos[d][ip], gw[d]);
      double temp_local = 0.0, drift_local = 0.0;
      for (int k = 0; k < (nd >= 3 ? nw : 1); ++k)
        for (int j = 0; j < (nd >= 2 ? nw : 1); ++j)
          for (int i = 0; i < nw; ++i) {
            double g = gw[0][i];
            if (nd >= 2) g = g * gw[1][j];
            if (nd >= 3) g = g * gw[2][k];
            temp_local = temp_local + g * temperature[dir];
            drift_local = drift_local + g * drift[dir];
          }
      const double stdev = std::sqrt(temp_local * KB * sp.mass);
      sp.p[dir][ip] = box_muller(g_rng, stdev, drift_local);
    }
  return np;
}

}  // namespace orc

struct orc_ctx { orc::State s; };     // as in orc_api.cpp

extern "C" {
// the generator on its own (global_random of the reference)
int32_t orc_random_init(int32_t seed) { orc::kiss_init(orc::g_rng, seed); return 0; }
double orc_random(void) { return orc::kiss(orc::g_rng); }
double orc_random_box_muller(double stdev, double mu) { return orc::box_muller(orc::g_rng, stdev, mu); }
// auto_load (helper.F90:89-163) of one species of the context's grid; the species must exist (set_species)
int32_t orc_load_uniform(orc_ctx* c, int32_t ispecies, int64_t npart_per_cell, double density,
                         const double temperature[3], const double drift[3], int64_t* n_out) {
  if (!c || !c->s.grid_set) return EB200_E_STATE;
  if (ispecies < 0 || ispecies >= (int32_t)c->s.species.size() || !c->s.species[ispecies].defined) return EB200_E_ARG;
  if (npart_per_cell < 1 || !temperature || !drift) return EB200_E_ARG;
  const int64_t n = orc::load_uniform(c->s, ispecies, npart_per_cell, density, temperature, drift);
  if (n_out) *n_out = n;
  return 0;
}
}
