// TEST INFRASTRUCTURE ONLY -- CPU oracle, never linked into the product.
// Restates the derived grid variables of epochNd/src/io/calc_df.F90 (3D line numbers):
// calc_mass_density :35, calc_ekbar :119, calc_ekflux :232, calc_poynt_flux :382,
// calc_charge_density :439, calc_number_density :523, calc_ppc :598,
// calc_average_weight :652, calc_temperature :723, calc_per_species_current :994,
// calc_average_momentum :1104 -- with the weights of include/particle_to_grid.inc.
#include "orc_state.hpp"
#include <vector>
#include <algorithm>
#include <cmath>

namespace orc {

namespace {

struct P2G {           // particle_to_grid.inc
  int cell[3] = {1, 1, 1};
  double w[3][5] = {{1, 0, 0, 0, 0}, {1, 0, 0, 0, 0}, {1, 0, 0, 0, 0}};
};

P2G particle_to_grid(const State& s, const Species& sp, size_t ip) {
  P2G r;
  for (int d = 0; d < s.ndims; ++d) {
    double cr = (sp.pos[d][ip] - s.grid_min_local[d]) / s.d[d];
    if (s.shape == 0) cr -= 0.5;
    r.cell[d] = (int)std::floor(cr + 0.5);
    const double frac = (double)r.cell[d] - cr;
    r.cell[d] += 1;
    weights_fac(s.shape, frac, r.w[d]);
  }
  return r;
}

// fn(index, gx*gy*gz) over the stencil sf_min..sf_max of every existing axis
template <class F>
void stencil(const State& s, const P2G& g, F fn) {
  const int nd = s.ndims, nw = s.sf_max - s.sf_min + 1;
  const int n1 = nd >= 2 ? nw : 1, n2 = nd >= 3 ? nw : 1;
  for (int k = 0; k < n2; ++k)
    for (int j = 0; j < n1; ++j)
      for (int i = 0; i < nw; ++i) {
        const int cx = g.cell[0] + s.sf_min + i;
        const int cy = nd >= 2 ? g.cell[1] + s.sf_min + j : 1;
        const int cz = nd >= 3 ? g.cell[2] + s.sf_min + k : 1;
        double gf = g.w[0][i];
        if (nd >= 2) gf = gf * g.w[1][j];
        if (nd >= 3) gf = gf * g.w[2][k];
        fn(s.idx(cx, cy, cz), gf);
      }
}

// the cell a particle is counted in by calc_ppc / calc_average_weight (:627-639)
int64_t count_cell(const State& s, const Species& sp, size_t ip) {
  int cell[3] = {1, 1, 1};
  for (int d = 0; d < s.ndims; ++d) {
    double cr = (sp.pos[d][ip] - s.grid_min_local[d]) / s.d[d];
    if (s.shape != 0) cr += 0.5;
    cell[d] = (int)std::floor(cr) + 1;
  }
  return s.idx(cell[0], cell[1], cell[2]);
}

bool skip(const Species& sp, bool spec_sum, bool photons_always) {
  if (!sp.defined) return true;
  if ((photons_always || spec_sum) && sp.species_type != 0) return true;
  if (spec_sum && sp.zero_current) return true;
  return false;
}

void zero_gradient_all(State& s, double* a) {
  for (int b = 0; b < 2 * s.ndims; ++b) field_zero_gradient_array(s, a, false, b);
}

void poynt_flux(const State& s, int dir, double* out) {
  const int nd = s.ndims;
  const double* F[6];
  for (int k = 0; k < 6; ++k) F[k] = s.f[k].data();
  // centre a component: E_a is staggered along a, B_a along the other two axes
  auto cc = [&](int comp, int64_t i) {
    const bool isb = comp >= 3;
    const int a = comp % 3;
    int ax[2], n = 0;
    for (int d = 0; d < nd; ++d)
      if (isb ? d != a : d == a) ax[n++] = d;
    const double* f = F[comp];
    if (n == 0) return f[i];
    if (n == 1) return 0.5 * (f[i - s.st[ax[0]]] + f[i]);
    // (lo,lo) + (hi_first,lo) + (lo,hi) + (hi,hi) with the first listed axis fastest
    return 0.25 * (f[i - s.st[ax[0]] - s.st[ax[1]]] + f[i - s.st[ax[1]]] + f[i - s.st[ax[0]]] + f[i]);
  };
  const int a = dir - 1, b = (a + 1) % 3, c = (a + 2) % 3;
  for (int iz = 1; iz <= (nd >= 3 ? s.n[2] : 1); ++iz)
    for (int iy = 1; iy <= (nd >= 2 ? s.n[1] : 1); ++iy)
      for (int ix = 1; ix <= s.n[0]; ++ix) {
        const int64_t i = s.idx(ix, iy, iz);
        out[i] = (cc(b, i) * cc(3 + c, i) - cc(c, i) * cc(3 + b, i)) / MU0;
      }
}

}  // namespace

// kind: 0 mass_density, 1 ekbar, 2 ekflux, 3 poynt_flux, 4 charge_density, 5 number_density,
// 6 ppc, 7 average_weight, 8 temperature, 9 per_species_current, 10 average_momentum.
// ispecies: 0-based, or -1 for the sum over species (current_species <= 0).
// direction: 1..3 (x,y,z), negative for the backward ekflux, 0 = absent (temperature: all three).
void calc_derived(State& s, int kind, int ispecies, int direction, double* out) {
  const int nd = s.ndims;
  const bool spec_sum = ispecies < 0;
  const size_t s0 = spec_sum ? 0 : (size_t)ispecies, s1 = spec_sum ? s.species.size() : (size_t)ispecies + 1;
  std::fill(out, out + s.total, 0.0);
  double vol = 1.0, idx = 1.0;
  for (int d = 0; d < nd; ++d) { vol *= s.d[d]; idx = idx / s.d[d]; }
  if (kind == 3) { poynt_flux(s, direction, out); return; }
  std::vector<double> wt;
  if (kind == 1 || kind == 2 || kind == 7 || kind == 10) wt.assign((size_t)s.total, 0.0);

  if (kind == 8) {                    // calc_temperature: two passes
    const int dir = direction >= 1 ? direction : -1;
    const double dof = direction >= 1 ? 1.0 : 3.0;
    std::vector<double> mean[3], cnt((size_t)s.total, 0.0);
    for (int k = 0; k < 3; ++k) mean[k].assign((size_t)s.total, 0.0);
    for (size_t is = s0; is < s1; ++is) {
      const Species& sp = s.species[is];
      if (skip(sp, spec_sum, true)) continue;
      const double sqm = std::sqrt(sp.mass);
      for (size_t ip = 0; ip < sp.count(); ++ip) {
        const double pm[3] = {sp.p[0][ip] / sqm, sp.p[1][ip] / sqm, sp.p[2][ip] / sqm};
        const double pw = sp.w[ip];
        stencil(s, particle_to_grid(s, sp, ip), [&](int64_t i, double g) {
          const double gf = g * pw;
          for (int k = 0; k < 3; ++k)
            if (dir < 0 || dir == k + 1) mean[k][i] += gf * pm[k];
          cnt[i] += gf;
        });
      }
    }
    for (int k = 0; k < 3; ++k)
      if (dir < 0 || dir == k + 1) processor_summation_array(s, mean[k].data(), -1);
    processor_summation_array(s, cnt.data(), -1);
    for (int64_t i = 0; i < s.total; ++i) {
      cnt[i] = std::max(cnt[i], 1.e-6);
      for (int k = 0; k < 3; ++k) mean[k][i] = mean[k][i] / cnt[i];
    }
    for (int k = 0; k < 3; ++k)
      if (dir < 0 || dir == k + 1) field_bc_array(s, mean[k].data());
    std::fill(cnt.begin(), cnt.end(), 0.0);
    for (size_t is = s0; is < s1; ++is) {
      const Species& sp = s.species[is];
      if (skip(sp, spec_sum, true)) continue;
      const double sqm = std::sqrt(sp.mass);
      for (size_t ip = 0; ip < sp.count(); ++ip) {
        const double pm[3] = {sp.p[0][ip] / sqm, sp.p[1][ip] / sqm, sp.p[2][ip] / sqm};
        stencil(s, particle_to_grid(s, sp, ip), [&](int64_t i, double gf) {
          double wdata = 0.0;
          for (int k = 0; k < 3; ++k)
            if (dir < 0 || dir == k + 1) { const double q = pm[k] - mean[k][i]; wdata = wdata + q * q; }
          out[i] += gf * wdata;
          cnt[i] += gf;
        });
      }
    }
    processor_summation_array(s, out, -1);
    processor_summation_array(s, cnt.data(), -1);
    for (int64_t i = 0; i < s.total; ++i) out[i] = out[i] / std::max(cnt[i], 1.e-6) / KB / dof;
    return;
  }

  for (size_t is = s0; is < s1; ++is) {
    const Species& sp = s.species[is];
    // which species are skipped differs per routine (photons: always for mass / charge /
    // current / momentum, only in a species sum for ekbar / ekflux / number density; ppc
    // and average weight count photons)
    const bool photons_always = kind == 0 || kind == 4 || kind == 9 || kind == 10;
    if (kind == 6 || kind == 7) { if (!sp.defined || (spec_sum && sp.zero_current)) continue; }
    else if (skip(sp, spec_sum, photons_always)) continue;
    const double part_mc = C * sp.mass;
    for (size_t ip = 0; ip < sp.count(); ++ip) {
      const double pw = sp.w[ip];
      if (kind == 6) { out[count_cell(s, sp, ip)] += 1.0; continue; }
      if (kind == 7) { const int64_t i = count_cell(s, sp, ip); out[i] += pw; wt[i] += 1.0; continue; }
      double wdata = 0.0;
      if (kind == 0) wdata = sp.mass * pw;
      else if (kind == 4) wdata = sp.charge * pw;
      else if (kind == 5) wdata = pw;
      else if (kind == 1 || kind == 2) {
        const double fac = part_mc * pw * C;
        const double u[3] = {sp.p[0][ip] / part_mc, sp.p[1][ip] / part_mc, sp.p[2][ip] / part_mc};
        const double u2 = u[0] * u[0] + u[1] * u[1] + u[2] * u[2];
        const double gamma_rel = std::sqrt(u2 + 1.0);
        wdata = u2 / (gamma_rel + 1.0) * fac;
        if (kind == 2) {
          const int a = std::abs(direction) - 1;
          double afac = C;                       // xfac = c * dy * dz and cyclic (:258-260)
          for (int d = 0; d < nd; ++d) if (d != a) afac = afac * s.d[d];
          const double flux = afac * u[a] / gamma_rel;
          wdata = direction < 0 ? -wdata * std::min(flux, 0.0) : wdata * std::max(flux, 0.0);
        }
      } else if (kind == 9) {
        const double px = sp.p[0][ip], py = sp.p[1][ip], pz = sp.p[2][ip];
        const double root = 1.0 / std::sqrt(part_mc * part_mc + px * px + py * py + pz * pz);
        wdata = sp.charge * pw * sp.p[direction - 1][ip] * root;
      } else if (kind == 10) wdata = pw * sp.p[direction - 1][ip];
      stencil(s, particle_to_grid(s, sp, ip), [&](int64_t i, double gf) {
        out[i] += gf * wdata;
        if (!wt.empty()) wt[i] += gf * pw;
      });
    }
  }
  if (kind == 6) return;
  if (kind == 7) {
    for (int64_t i = 0; i < s.total; ++i) out[i] = out[i] / std::max(wt[i], C_TINY);
    return;
  }
  processor_summation_array(s, out, -1);
  if (!wt.empty()) {
    processor_summation_array(s, wt.data(), -1);
    for (int64_t i = 0; i < s.total; ++i) out[i] = out[i] / std::max(wt[i], C_TINY);
  } else {
    const double scale = kind == 9 ? C * idx : idx;
    for (int64_t i = 0; i < s.total; ++i) out[i] = out[i] * scale;
  }
  zero_gradient_all(s, out);
}

// general_dist_fn, the particle loop (io/dist_fn.F90:532-628), statement by statement; `out` is
// array(resolution(1), resolution(2), resolution(3)) in Fortran order.  xy_max etc.: :505-510.
void dist_fn(const State& s, int ispecies, int curdims, const int32_t* direction, const double* ranges,
             const int32_t* resolution, const double* dgrid, const double* restrictions,
             const int32_t* use_restrictions, double* out) {
  const Species& sp = s.species[ispecies];
  const int nd = s.ndims, maxdirs = nd + 9;
  const int c_dir_px = nd + 1, c_dir_py = nd + 2, c_dir_pz = nd + 3, c_dir_en = nd + 4, c_dir_gamma_m1 = nd + 5,
            c_dir_xy_angle = nd + 6, c_dir_yz_angle = nd + 7, c_dir_zx_angle = nd + 8, c_dir_mod_p = nd + 9;
  long long tot = 1;
  bool use_xy = false, use_yz = false, use_zx = false;
  double xy_max = 0, yz_max = 0, zx_max = 0;
  for (int i = 0; i < curdims; ++i) {
    tot *= resolution[i];
    if (direction[i] == c_dir_xy_angle) { use_xy = true; xy_max = ranges[2 * i + 1]; }
    if (direction[i] == c_dir_yz_angle) { use_yz = true; yz_max = ranges[2 * i + 1]; }
    if (direction[i] == c_dir_zx_angle) { use_zx = true; zx_max = ranges[2 * i + 1]; }
  }
  for (long long i = 0; i < tot; ++i) out[i] = 0.0;
  const double part_mc = sp.mass * C, part_mc2 = part_mc * C, pi2 = 2.0 * PI;
  std::vector<double> particle_data(maxdirs + 1, 0.0);
  for (size_t ip = 0; ip < sp.count(); ++ip) {
    const double part_weight = sp.w[ip];
    const double px = sp.p[0][ip], py = sp.p[1][ip], pz = sp.p[2][ip];
    const double a = px / part_mc, b = py / part_mc, cc = pz / part_mc;
    const double part_u2 = a * a + b * b + cc * cc;
    const double gamma_rel = std::sqrt(part_u2 + 1.0);
    const double gamma_rel_m1 = part_u2 / (gamma_rel + 1.0);
    std::fill(particle_data.begin(), particle_data.end(), 0.0);
    for (int d = 0; d < nd; ++d) particle_data[1 + d] = sp.pos[d][ip];
    if (sp.species_type == 1) {
      particle_data[c_dir_gamma_m1] = 1.0;
    } else {
      particle_data[c_dir_px] = px; particle_data[c_dir_py] = py; particle_data[c_dir_pz] = pz;
      particle_data[c_dir_en] = gamma_rel_m1 * part_mc2;
      particle_data[c_dir_gamma_m1] = gamma_rel_m1;
      particle_data[c_dir_mod_p] = std::sqrt(px * px + py * py + pz * pz);
    }
    bool skip = false;
    auto angle = [&](double q0, double q1, double amax, int code) {
      const double p = std::sqrt(q0 * q0 + q1 * q1);
      if (p < C_TINY) { skip = true; return; }
      double theta = std::asin(q1 / p);
      if (q0 < 0.0) theta = std::copysign(1.0, q1) * PI - theta;
      if ((theta + pi2) < amax) theta = theta + pi2;
      particle_data[code] = theta;
    };
    if (use_xy) angle(px, py, xy_max, c_dir_xy_angle);
    if (!skip && use_yz) angle(py, pz, yz_max, c_dir_yz_angle);
    if (!skip && use_zx) angle(pz, px, zx_max, c_dir_zx_angle);
    if (skip) continue;
    if (restrictions) {
      for (int idir = 1; idir <= maxdirs && !skip; ++idir)
        if (use_restrictions[idir - 1] &&
            (particle_data[idir] < restrictions[2 * (idir - 1)] || particle_data[idir] > restrictions[2 * (idir - 1) + 1]))
          skip = true;
      if (skip) continue;
    }
    long long idx = 0, mul = 1;
    for (int idim = 0; idim < curdims && !skip; ++idim) {
      const double current_data = particle_data[direction[idim]];
      const double cf = std::floor((current_data - ranges[2 * idim]) / dgrid[idim]) + 1.0;
      if (!(cf >= 1.0) || cf > (double)resolution[idim]) { skip = true; break; }
      idx += mul * ((long long)cf - 1);
      mul *= resolution[idim];
    }
    if (skip) continue;
    out[idx] += part_weight;
  }
}

}  // namespace orc
