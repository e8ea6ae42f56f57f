// oracle/orc_particles.cpp -- CPU ORACLE (test infrastructure only).
// Restates epochNd/src/particles.F90 push_particles (3D :28-614, 2D, 1D) with the
// shape include files epochNd/src/include/{tophat,triangle,bspline3}/*.inc, in the
// reference's operation order, plus particle_bcs
// (boundary.F90:1316-1972) and the diagnostics formulas of io/calc_df.F90.
//
// PARITY UNPINNED BY THE REFERENCE: EPOCH's own tests hold no particle-dependent number and EPOCH cannot be built
// here (Fortran + MPI + the missing SDF submodule), so nothing in this file has been compared with reference output
// (tools/ref_build/ is the recipe that would do it).  What pins it instead, stage by stage (tests/test_oracle.py,
// DESIGN.md section 5): the gather's Yee staggering by exact reproduction of linear fields, the shape weights by the
// B-spline central moments, the rotation by the textbook Boris form, the position update, the deposit by Esirkepov's
// scheme written from the paper, particle_bcs by placed particles, the whole by exact charge conservation and the
// reference's two-stream deck (growth rates, Langmuir frequency, energy bound).
#include "orc_state.hpp"
#include <algorithm>

namespace orc {

// un-normalised 1-D weights: include/*/gx.inc, hx_dcell.inc.  out[0..nw-1] are the
// slots sf_min..sf_max.
static inline void weights(int shape, double f, double* out) {
  if (shape == 0) {            // tophat/gx.inc:3-4
    out[0] = 0.5 + f;
    out[1] = 0.5 - f;
  } else if (shape == 1) {     // triangle/gx.inc:1-4
    const double cf2 = f * f;
    out[0] = 0.25 + cf2 + f;
    out[1] = 1.5 - 2.0 * cf2;
    out[2] = 0.25 + cf2 - f;
  } else {                     // bspline3/gx.inc:3-10
    const double cf2 = f * f;
    double a = 0.5 + f; a = a * a; out[0] = a * a;
    out[1] = 4.75 + 11.0 * f + 4.0 * cf2 * (1.5 - f - cf2);
    out[2] = 14.375 + 6.0 * cf2 * (cf2 - 2.5);
    out[3] = 4.75 - 11.0 * f + 4.0 * cf2 * (1.5 + f - cf2);
    double b = 0.5 - f; b = b * b; out[4] = b * b;
  }
}

// tensor-product gather in the reference's nesting order
// sum_k wz(k)*( sum_j wy(j)*( sum_i wx(i)*F ) ), include/*/e_part.inc, b_part.inc
static inline double gather(const State& s, const double* F, const double* const w[3],
                            const int cell[3], int nw) {
  const int nd = s.ndims;
  const int sfm = s.sf_min;
  if (nd == 1) {
    double acc = 0.0;
    for (int i = 0; i < nw; ++i) {
      const double t = w[0][i] * F[s.idx(cell[0] + sfm + i, 1, 1)];
      acc = i == 0 ? t : acc + t;
    }
    return acc;
  }
  if (nd == 2) {
    double accy = 0.0;
    for (int j = 0; j < nw; ++j) {
      double accx = 0.0;
      for (int i = 0; i < nw; ++i) {
        const double t = w[0][i] * F[s.idx(cell[0] + sfm + i, cell[1] + sfm + j, 1)];
        accx = i == 0 ? t : accx + t;
      }
      const double t = w[1][j] * accx;
      accy = j == 0 ? t : accy + t;
    }
    return accy;
  }
  double accz = 0.0;
  for (int k = 0; k < nw; ++k) {
    double accy = 0.0;
    for (int j = 0; j < nw; ++j) {
      double accx = 0.0;
      for (int i = 0; i < nw; ++i) {
        const double t = w[0][i] * F[s.idx(cell[0] + sfm + i, cell[1] + sfm + j, cell[2] + sfm + k)];
        accx = i == 0 ? t : accx + t;
      }
      const double t = w[1][j] * accx;
      accy = j == 0 ? t : accy + t;
    }
    const double t = w[2][k] * accy;
    accz = k == 0 ? t : accz + t;
  }
  return accz;
}

void push_particles(State& s) {
  const int nd = s.ndims, shape = s.shape;
  const int sf_min = s.sf_min, sf_max = s.sf_max;
  const int nw = sf_max - sf_min + 1;
  const int NS = nw + 2;                 // slots sf_min-1 .. sf_max+1
  auto slot = [&](int i) { return i - (sf_min - 1); };
  // particles.F90:128-136
  double fac;
  if (shape == 2) { fac = 1.0; for (int d = 0; d < nd; ++d) fac *= (1.0 / 24.0); }
  else if (shape == 0) fac = 1.0;
  else { fac = 1.0; for (int d = 0; d < nd; ++d) fac *= 0.5; }

  // particles.F90:147-149
  for (int c = JX; c <= JZ; ++c) std::fill(s.f[c].begin(), s.f[c].end(), 0.0);

  // particles.F90:153-168
  double id[3];
  for (int d = 0; d < 3; ++d) id[d] = 1.0 / s.d[d];
  const double idt = 1.0 / s.dt, dto2 = s.dt / 2.0, dtco2 = C * dto2;
  const double dtfac = 0.5 * s.dt * fac, third = 1.0 / 3.0;
  // 3D: idtyz, idtxz, idtxy; 2D: idty, idtx, idxy; 1D: idtf, idxf
  double fc0 = 0, fc1 = 0, fc2 = 0;
  if (nd == 3) { fc0 = idt * id[1] * id[2] * fac; fc1 = idt * id[0] * id[2] * fac; fc2 = idt * id[0] * id[1] * fac; }
  else if (nd == 2) { fc0 = idt * id[1] * fac; fc1 = idt * id[0] * fac; fc2 = id[0] * id[1] * fac; }
  else { fc0 = idt * fac; fc1 = id[0] * fac; fc2 = 0; }

  const double* Fex = s.f[EX].data(); const double* Fey = s.f[EY].data(); const double* Fez = s.f[EZ].data();
  const double* Fbx = s.f[BX].data(); const double* Fby = s.f[BY].data(); const double* Fbz = s.f[BZ].data();
  double* Jx = s.f[JX].data(); double* Jy = s.f[JY].data(); double* Jz = s.f[JZ].data();

  std::vector<double> g[3], h[3];
  for (int d = 0; d < 3; ++d) { g[d].assign(NS, 0.0); h[d].assign(NS, 0.0); }
  std::vector<double> jyh(NS), jzh((size_t)NS * NS);

  for (Species& sp : s.species) {
    if (!sp.defined || sp.immobile) continue;
    if (sp.species_type == 1) continue;  // photon species: not on this path
    const double part_q = sp.charge;
    const double part_mc = C * sp.mass;
    const double ipart_mc = 1.0 / part_mc;
    const double cmratio = part_q * dtfac * ipart_mc;
    const double ccmratio = C * cmratio;
    const size_t np = sp.count();
    for (size_t ip = 0; ip < np; ++ip) {
      const double part_weight = sp.w[ip];
      const double fcx = fc0 * part_weight, fcy = fc1 * part_weight, fcz = fc2 * part_weight;
      double px[3] = {0, 0, 0};
      double init_pos[3] = {0, 0, 0};                       // init_part_x, .. (particles.F90:226-228)
      for (int d = 0; d < nd; ++d) init_pos[d] = sp.pos[d][ip];
      for (int d = 0; d < nd; ++d) px[d] = sp.pos[d][ip] - s.grid_min_local[d];
      double ux = sp.p[0][ip] * ipart_mc, uy = sp.p[1][ip] * ipart_mc, uz = sp.p[2][ip] * ipart_mc;
      double gamma_rel = std::sqrt(ux * ux + uy * uy + uz * uz + 1.0);
      double root = dtco2 / gamma_rel;
      const double uvec0[3] = {ux, uy, uz};
      for (int d = 0; d < nd; ++d) px[d] = px[d] + uvec0[d] * root;

      int cell1[3] = {1, 1, 1}, cell2[3] = {1, 1, 1};
      double cell_r[3];
      double wg[3][5], wh[3][5];
      for (int d = 0; d < nd; ++d) {
        cell_r[d] = shape == 0 ? px[d] * id[d] - 0.5 : px[d] * id[d];
        cell1[d] = (int)std::floor(cell_r[d] + 0.5);
        double frac = (double)cell1[d] - cell_r[d];
        cell1[d] += 1;
        weights(shape, frac, wg[d]);
        cell2[d] = (int)std::floor(cell_r[d]);
        frac = (double)cell2[d] - cell_r[d] + 0.5;
        cell2[d] += 1;
        weights(shape, frac, wh[d]);
        // g arrays keep zero end slots (particles.F90:151-153)
        std::fill(g[d].begin(), g[d].end(), 0.0);
        for (int i = 0; i < nw; ++i) g[d][slot(sf_min + i)] = wg[d][i];
      }
      // gather: stagger table in SURVEY a6 (include/*/e_part.inc, b_part.inc)
      auto G = [&](const double* F, bool sx, bool sy, bool sz) {
        const double* w[3] = {sx ? wh[0] : wg[0], sy ? wh[1] : wg[1], sz ? wh[2] : wg[2]};
        const int cell[3] = {sx ? cell2[0] : cell1[0], sy ? cell2[1] : cell1[1], sz ? cell2[2] : cell1[2]};
        return gather(s, F, w, cell, nw);
      };
      const double ex_part = G(Fex, true, false, false);
      const double ey_part = G(Fey, false, true, false);
      const double ez_part = G(Fez, false, false, true);
      const double bx_part = G(Fbx, false, true, true);
      const double by_part = G(Fby, true, false, true);
      const double bz_part = G(Fbz, true, true, false);

      // Boris: particles.F90:347-394
      const double uxm = ux + cmratio * ex_part;
      const double uym = uy + cmratio * ey_part;
      const double uzm = uz + cmratio * ez_part;
      gamma_rel = std::sqrt(uxm * uxm + uym * uym + uzm * uzm + 1.0);
      root = ccmratio / gamma_rel;
      const double taux = bx_part * root, tauy = by_part * root, tauz = bz_part * root;
      const double taux2 = taux * taux, tauy2 = tauy * tauy, tauz2 = tauz * tauz;
      const double tau = 1.0 / (1.0 + taux2 + tauy2 + tauz2);
      const double uxp = ((1.0 + taux2 - tauy2 - tauz2) * uxm
          + 2.0 * ((taux * tauy + tauz) * uym + (taux * tauz - tauy) * uzm)) * tau;
      const double uyp = ((1.0 - taux2 + tauy2 - tauz2) * uym
          + 2.0 * ((tauy * tauz + taux) * uzm + (tauy * taux - tauz) * uxm)) * tau;
      const double uzp = ((1.0 - taux2 - tauy2 + tauz2) * uzm
          + 2.0 * ((tauz * taux + tauy) * uxm + (tauz * tauy - taux) * uym)) * tau;
      ux = uxp + cmratio * ex_part;
      uy = uyp + cmratio * ey_part;
      uz = uzp + cmratio * ez_part;

      // drift: 3D particles.F90:396-408, 2D :378-390, 1D :356-366
      const double part_u2 = ux * ux + uy * uy + uz * uz;
      gamma_rel = std::sqrt(part_u2 + 1.0);
      double delta[3] = {0, 0, 0};
      double part_vy = 0, part_vz = 0;
      if (nd == 3) {
        root = dtco2 / gamma_rel;
        delta[0] = ux * root; delta[1] = uy * root; delta[2] = uz * root;
      } else if (nd == 2) {
        const double igamma = 1.0 / gamma_rel;
        root = dtco2 * igamma;
        delta[0] = ux * root; delta[1] = uy * root;
        part_vz = uz * C * igamma;
      } else {
        root = C / gamma_rel;
        delta[0] = ux * root * dto2;
        part_vy = uy * root;
        part_vz = uz * root;
      }
      for (int d = 0; d < nd; ++d) px[d] = px[d] + delta[d];
      for (int d = 0; d < nd; ++d) sp.pos[d][ip] = px[d] + s.grid_min_local[d];
      sp.p[0][ip] = part_mc * ux; sp.p[1][ip] = part_mc * uy; sp.p[2][ip] = part_mc * uz;
      {
        // particle probes (particles.F90:567-606; in the reference after the deposit, which does not touch
        // the quantities used here)
        const double gamma_rel_m1 = part_u2 / (gamma_rel + 1.0);
        const double probe_energy = gamma_rel_m1 * (C * part_mc);
        for (Species::Probe& pr : sp.probes) {
          if (!pr.on) continue;
          if (probe_energy > pr.ek_min && probe_energy < pr.ek_max) {
            double d_init = 0.0, d_final = 0.0;
            for (int d = 0; d < nd; ++d) {
              d_init += pr.normal[d] * (pr.point[d] - init_pos[d]);
              d_final += pr.normal[d] * (pr.point[d] - sp.pos[d][ip]);
            }
            if (d_final < 0.0 && d_init >= 0.0) {
              for (int d = 0; d < nd; ++d) pr.sampled.push_back(sp.pos[d][ip]);
              for (int k = 0; k < 3; ++k) pr.sampled.push_back(sp.p[k][ip]);
              pr.sampled.push_back(part_weight);
            }
          }
        }
      }

      if (sp.zero_current) continue;

      // deposit: particles.F90:443-563
      int dcell[3] = {0, 0, 0}, lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
      for (int d = 0; d < nd; ++d) {
        px[d] = px[d] + delta[d];
        const double cr = shape == 0 ? px[d] * id[d] - 0.5 : px[d] * id[d];
        int cell3 = (int)std::floor(cr + 0.5);
        const double frac = (double)cell3 - cr;
        cell3 += 1;
        std::fill(h[d].begin(), h[d].end(), 0.0);
        dcell[d] = cell3 - cell1[d];
        double w3[5];
        weights(shape, frac, w3);
        for (int i = 0; i < nw; ++i) {
          const int sl = slot(dcell[d] + sf_min + i);
          if (sl >= 0 && sl < NS) h[d][sl] = w3[i];
        }
        for (int i = 0; i < NS; ++i) h[d][i] = h[d][i] - g[d][i];
        lo[d] = sf_min + (dcell[d] - 1) / 2;   // C++ '/' truncates toward zero like Fortran
        hi[d] = sf_max + (dcell[d] + 1) / 2;
      }
      const double* gx = g[0].data(); const double* hx = h[0].data();
      const double* gy = g[1].data(); const double* hy = h[1].data();
      const double* gz = g[2].data(); const double* hz = h[2].data();
      if (nd == 3) {
        const double fjx = fcx * part_q, fjy = fcy * part_q, fjz = fcz * part_q;
        std::fill(jzh.begin(), jzh.end(), 0.0);
        for (int iz = lo[2]; iz <= hi[2]; ++iz) {
          const int cz = cell1[2] + iz, kz = slot(iz);
          const double zfac1 = gz[kz] + 0.5 * hz[kz];
          const double zfac2 = third * hz[kz] + 0.5 * gz[kz];
          const double gz_iz = gz[kz], hz_iz = hz[kz];
          std::fill(jyh.begin(), jyh.end(), 0.0);
          for (int iy = lo[1]; iy <= hi[1]; ++iy) {
            const int cy = cell1[1] + iy, ky = slot(iy);
            const double yfac1 = gy[ky] + 0.5 * hy[ky];
            const double yfac2 = third * hy[ky] + 0.5 * gy[ky];
            const double hygz = hy[ky] * gz_iz, hyhz = hy[ky] * hz_iz;
            const double yzfac = gy[ky] * zfac1 + hy[ky] * zfac2;
            const double hzyfac1 = hz_iz * yfac1, hzyfac2 = hz_iz * yfac2;
            double jxh = 0.0;
            for (int ix = lo[0]; ix <= hi[0]; ++ix) {
              const int cx = cell1[0] + ix, kx = slot(ix);
              const double xfac1 = gx[kx] + 0.5 * hx[kx];
              const double xfac2 = third * hx[kx] + 0.5 * gx[kx];
              const double wx = hx[kx] * yzfac;
              const double wy = xfac1 * hygz + xfac2 * hyhz;
              const double wz = gx[kx] * hzyfac1 + hx[kx] * hzyfac2;
              jxh = jxh - fjx * wx;
              jyh[kx] = jyh[kx] - fjy * wy;
              jzh[(size_t)ky * NS + kx] = jzh[(size_t)ky * NS + kx] - fjz * wz;
              const int64_t i = s.idx(cx, cy, cz);
              Jx[i] = Jx[i] + jxh;
              Jy[i] = Jy[i] + jyh[kx];
              Jz[i] = Jz[i] + jzh[(size_t)ky * NS + kx];
            }
          }
        }
      } else if (nd == 2) {
        // epoch2d particles.F90:470-509
        const double fjx = fcx * part_q, fjy = fcy * part_q, fjz = fcz * part_q * part_vz;
        std::fill(jyh.begin(), jyh.end(), 0.0);
        for (int iy = lo[1]; iy <= hi[1]; ++iy) {
          const int cy = cell1[1] + iy, ky = slot(iy);
          const double yfac1 = gy[ky] + 0.5 * hy[ky];
          const double yfac2 = third * hy[ky] + 0.5 * gy[ky];
          const double hy_iy = hy[ky];
          double jxh = 0.0;
          for (int ix = lo[0]; ix <= hi[0]; ++ix) {
            const int cx = cell1[0] + ix, kx = slot(ix);
            const double xfac1 = gx[kx] + 0.5 * hx[kx];
            const double wx = hx[kx] * yfac1;
            const double wy = hy_iy * xfac1;
            const double wz = gx[kx] * yfac1 + hx[kx] * yfac2;
            jxh = jxh - fjx * wx;
            jyh[kx] = jyh[kx] - fjy * wy;
            const double jzh1 = fjz * wz;
            const int64_t i = s.idx(cx, cy, 1);
            Jx[i] = Jx[i] + jxh;
            Jy[i] = Jy[i] + jyh[kx];
            Jz[i] = Jz[i] + jzh1;
          }
        }
      } else {
        // epoch1d particles.F90:432-457
        const double fjx = fcx * part_q;
        const double fjy = fcy * part_q * part_vy;
        const double fjz = fcy * part_q * part_vz;
        double jxh = 0.0;
        for (int ix = lo[0]; ix <= hi[0]; ++ix) {
          const int cx = cell1[0] + ix, kx = slot(ix);
          const double wx = hx[kx];
          const double wy = gx[kx] + 0.5 * hx[kx];
          jxh = jxh - fjx * wx;
          const double jyh1 = fjy * wy;
          const double jzh1 = fjz * wy;
          const int64_t i = s.idx(cx, 1, 1);
          Jx[i] = Jx[i] + jxh;
          Jy[i] = Jy[i] + jyh1;
          Jz[i] = Jz[i] + jzh1;
        }
      }
    }
    // current_bcs(species) (particles.F90:609, boundary.F90:1976-1985): a no-op unless per-species BCs are mixed
    if (bc_mixed(s)) {
      processor_summation_array(s, s.f[JX].data(), 0, sp.bc);
      processor_summation_array(s, s.f[JY].data(), 1, sp.bc);
      processor_summation_array(s, s.f[JZ].data(), 2, sp.bc);
    }
  }
  particle_bcs(s);
}

// momentum_from_temperature / flux_momentum_from_temperature with drift = 0 (particle_temperature.F90:396-469)
static double momentum_from_temperature(double mass, double temperature) {
  const double stdev = std::sqrt(temperature * KB * mass);
  return global_random_box_muller(stdev, 0.0);
}
static double flux_momentum_from_temperature(double mass, double temperature, double direction) {
  // drift == 0: Rayleigh distributed, Z = SQRT(X**2 + Y**2) of two normal deviates
  const double mom1 = momentum_from_temperature(mass, temperature);
  const double mom2 = momentum_from_temperature(mass, temperature);
  return direction * std::sqrt(mom1 * mom1 + mom2 * mom2);
}

// Thermal boundary (boundary.F90:1405-1461 for x_min, the five others alike): the temperature at the particle's
// transverse position (always the triangle weighting), new momenta drawn in the order x, y, z -- the normal
// component from the flux distribution, heading back in -- and the position mirrored at the outer edge.
static void thermal_reemit(State& s, Species& sp, size_t ip, int d, int side, double edge, double part_pos) {
  const int nd = s.ndims;
  int t[2], nt = 0;
  for (int a = 0; a < nd; ++a) if (a != d) t[nt++] = a;
  int cell[2] = {1, 1};
  double g[2][3] = {{0.0, 1.0, 0.0}, {0.0, 1.0, 0.0}};
  for (int k = 0; k < nt; ++k) {
    const int a = t[k];
    const double cell_r = (sp.pos[a][ip] - s.grid_min_local[a]) / s.d[a];
    const int c0 = (int)std::floor(cell_r + 0.5);
    const double cell_frac = (double)c0 - cell_r;
    cell[k] = c0 + 1;
    const double cf2 = cell_frac * cell_frac;
    g[k][0] = 0.5 * (0.25 + cf2 + cell_frac);
    g[k][1] = 0.75 - cf2;
    g[k][2] = 0.5 * (0.25 + cf2 - cell_frac);
  }
  const std::vector<double>& et = sp.ext_temp[2 * d + side];
  const int64_t e1 = nt >= 1 ? s.n[t[0]] + 2 * s.ng : 1, e2 = nt >= 2 ? s.n[t[1]] + 2 * s.ng : 1;
  double temp[3];
  for (int i = 0; i < 3; ++i) {
    temp[i] = 0.0;
    if (nt == 0) { temp[i] = et[i]; continue; }
    for (int i2 = (nt >= 2 ? -1 : 0); i2 <= (nt >= 2 ? 1 : 0); ++i2)
      for (int i1 = -1; i1 <= 1; ++i1) {
        const int64_t j1 = cell[0] + i1 - (1 - s.ng), j2 = nt >= 2 ? cell[1] + i2 - (1 - s.ng) : 0;
        const double w2 = nt >= 2 ? g[1][i2 + 1] : 1.0;
        temp[i] = temp[i] + g[0][i1 + 1] * w2 * et[((int64_t)i * e2 + j2) * e1 + j1];
      }
  }
  const double direction = side == 0 ? 1.0 : -1.0;      // -REAL(sgn, num)
  for (int i = 0; i < 3; ++i)
    sp.p[i][ip] = i == d ? flux_momentum_from_temperature(sp.mass, temp[i], direction)
                         : momentum_from_temperature(sp.mass, temp[i]);
  sp.pos[d][ip] = 2.0 * edge - part_pos;
}

// boundary.F90:1316-1972 (single rank: a particle that leaves through a periodic
// face wraps by the domain length; reflect mirrors; anything else is "open" and
// the particle is removed once beyond x_min_outer / x_max_outer).
void particle_bcs(State& s) {
  const int nd = s.ndims;
  for (Species& sp : s.species) {
    if (!sp.defined) continue;
    const size_t np = sp.count();
    std::vector<char> dead(np, 0);
    bool any_dead = false;
    for (size_t ip = 0; ip < np; ++ip) {
      bool out_of_bounds = false;
      for (int d = 0; d < nd; ++d) {
        const double boundary_shift = s.d[d] * (double)((1 + s.png + s.cpml_thickness) / 2);
        const double min_outer = s.gmin[d] - boundary_shift, max_outer = s.gmax[d] + boundary_shift;
        const double shift = (s.gmax[d] - s.gmin[d]) + 2.0 * s.d[d] * (double)s.cpml_thickness;
        const double part_pos = sp.pos[d][ip];
        const int fmin = s.bc_field[2 * d], fmax = s.bc_field[2 * d + 1];
        const bool cpml_min = fmin == BC_CPML_LASER || fmin == BC_CPML_OUTFLOW;
        const bool cpml_max = fmax == BC_CPML_LASER || fmax == BC_CPML_OUTFLOW;
        if (cpml_min) {        // boundary.F90:1386-1402: through a CPML the particle only ever leaves
          if (s.min_bnd[d] && part_pos < min_outer) out_of_bounds = true;
        } else if (part_pos < s.min_local[d]) {
          if (s.min_bnd[d]) {
            const int bc = sp.bc[2 * d];
            if (bc == BC_REFLECT) {
              sp.pos[d][ip] = 2.0 * s.gmin[d] - part_pos;
              sp.p[d][ip] = -sp.p[d][ip];
            } else if (bc == BC_PERIODIC) {
              sp.pos[d][ip] = part_pos - (-1.0) * shift;
            }
            if (part_pos < min_outer && bc != BC_PERIODIC) {
              if (bc == BC_THERMAL) thermal_reemit(s, sp, ip, d, 0, min_outer, part_pos);
              else out_of_bounds = true;
            }
          }
        }
        if (cpml_max) {
          if (s.max_bnd[d] && part_pos >= max_outer) out_of_bounds = true;
        } else if (part_pos >= s.max_local[d]) {
          if (s.max_bnd[d]) {
            const int bc = sp.bc[2 * d + 1];
            if (bc == BC_REFLECT) {
              sp.pos[d][ip] = 2.0 * s.gmax[d] - part_pos;
              sp.p[d][ip] = -sp.p[d][ip];
            } else if (bc == BC_PERIODIC) {
              sp.pos[d][ip] = part_pos - shift;
            }
            if (part_pos >= max_outer && bc != BC_PERIODIC) {
              if (bc == BC_THERMAL) thermal_reemit(s, sp, ip, d, 1, max_outer, part_pos);
              else out_of_bounds = true;
            }
          }
        }
      }
      if (out_of_bounds) { dead[ip] = 1; any_dead = true; }
    }
    if (any_dead) {
      size_t o = 0;
      for (size_t ip = 0; ip < np; ++ip) {
        if (dead[ip]) continue;
        for (int d = 0; d < nd; ++d) sp.pos[d][o] = sp.pos[d][ip];
        for (int d = 0; d < 3; ++d) sp.p[d][o] = sp.p[d][ip];
        sp.w[o] = sp.w[ip];
        ++o;
      }
      for (int d = 0; d < nd; ++d) sp.pos[d].resize(o);
      for (int d = 0; d < 3; ++d) sp.p[d].resize(o);
      sp.w.resize(o);
    }
  }
}

// io/calc_df.F90:1190-1288 calc_total_energy_sum (local part)
void energy(const State& s, double* field_energy, double* per_species) {
  for (size_t is = 0; is < s.species.size(); ++is) {
    const Species& sp = s.species[is];
    double part_energy = 0.0;
    if (sp.defined && !sp.zero_current) {
      const double part_mc = C * sp.mass;
      for (size_t ip = 0; ip < sp.count(); ++ip) {
        const double fac = part_mc * sp.w[ip] * C;
        const double ux = sp.p[0][ip] / part_mc, uy = sp.p[1][ip] / part_mc, uz = sp.p[2][ip] / part_mc;
        const double u2 = ux * ux + uy * uy + uz * uz;
        const double gamma_rel = std::sqrt(u2 + 1.0);
        part_energy = part_energy + (u2 / (gamma_rel + 1.0)) * fac;
      }
    }
    if (per_species) per_species[is] = part_energy;
  }
  double fe = 0.0;
  const double c2 = C * C;
  for (int k = 1; k <= s.n[2]; ++k)
    for (int j = 1; j <= s.n[1]; ++j)
      for (int i = 1; i <= s.n[0]; ++i) {
        const int64_t q = s.idx(i, j, k);
        const double ex = s.f[EX][q], ey = s.f[EY][q], ez = s.f[EZ][q];
        const double bx = s.f[BX][q], by = s.f[BY][q], bz = s.f[BZ][q];
        fe = fe + ex * ex + ey * ey + ez * ez + c2 * (bx * bx + by * by + bz * bz);
      }
  double vol = 1.0;
  for (int d = 0; d < s.ndims; ++d) vol *= s.d[d];
  if (field_energy) *field_energy = 0.5 * EPS0 * fe * vol;
}

// normalised weights of include/*/gxfac.inc
void weights_fac(int shape, double f, double* out) {
  if (shape == 0) { out[0] = 0.5 + f; out[1] = 0.5 - f; }
  else if (shape == 1) {
    const double c2 = f * f;
    out[0] = 0.5 * (0.25 + c2 + f); out[1] = 0.75 - c2; out[2] = 0.5 * (0.25 + c2 - f);
  } else {
    const double fac1 = 1.0 / 24.0, fac2 = 1.0 / 6.0, fac3 = 115.0 / 48.0;
    const double c2 = f * f;
    double a = 0.5 + f; a = a * a; out[0] = fac1 * (a * a);
    out[1] = fac2 * (1.1875 + 2.75 * f + c2 * (1.5 - f - c2));
    out[2] = 0.25 * (fac3 + c2 * (c2 - 2.5));
    out[3] = fac2 * (1.1875 - 2.75 * f + c2 * (1.5 + f - c2));
    double b = 0.5 - f; b = b * b; out[4] = fac1 * (b * b);
  }
}

// rho as calc_charge_density (io/calc_df.F90:439-519) on a single periodic rank,
// then max |div E - rho/eps0| over interior cells.
// Positions are first advanced by the half drift x + u*c*dt/(2*gamma)
// (particles.F90:242-257): the Esirkepov current links rho at the t+1/2 and t+3/2
// positions, so div E(t_n) - rho(x_n + half drift)/eps0 is the exact invariant.
void charge_residual(State& s, double* max_res, double* max_rho, bool capture) {
  const int nd = s.ndims, nw = s.sf_max - s.sf_min + 1;
  std::vector<double> rho(s.total, 0.0);
  const double dtco2 = C * (s.dt / 2.0);
  double idvol = 1.0;
  for (int d = 0; d < nd; ++d) idvol /= s.d[d];
  for (const Species& sp : s.species) {
    if (!sp.defined) continue;
    for (size_t ip = 0; ip < sp.count(); ++ip) {
      int cell[3] = {1, 1, 1};
      double w[3][5] = {{1, 0, 0, 0, 0}, {1, 0, 0, 0, 0}, {1, 0, 0, 0, 0}};
      const double ipmc = 1.0 / (C * sp.mass);
      const double uu[3] = {sp.p[0][ip] * ipmc, sp.p[1][ip] * ipmc, sp.p[2][ip] * ipmc};
      const double root = dtco2 / std::sqrt(uu[0] * uu[0] + uu[1] * uu[1] + uu[2] * uu[2] + 1.0);
      for (int d = 0; d < nd; ++d) {
        double cr = ((sp.pos[d][ip] - s.grid_min_local[d]) + uu[d] * root) / s.d[d];
        if (s.shape == 0) cr -= 0.5;
        cell[d] = (int)std::floor(cr + 0.5);
        const double frac = (double)cell[d] - cr;
        cell[d] += 1;
        weights_fac(s.shape, frac, w[d]);
      }
      const double wdata = sp.charge * sp.w[ip] * idvol;
      const int n1 = nd >= 2 ? nw : 1, n2 = nd >= 3 ? nw : 1;
      for (int k = 0; k < n2; ++k)
        for (int j = 0; j < n1; ++j)
          for (int i = 0; i < nw; ++i) {
            const int cx = cell[0] + s.sf_min + i;
            const int cy = nd >= 2 ? cell[1] + s.sf_min + j : 1;
            const int cz = nd >= 3 ? cell[2] + s.sf_min + k : 1;
            rho[s.idx(cx, cy, cz)] += w[0][i] * (nd >= 2 ? w[1][j] : 1.0) * (nd >= 3 ? w[2][k] : 1.0) * wdata;
          }
    }
  }
  // fold ghosts (periodic dims only)
  for (int ax = 0; ax < nd; ++ax) {
    if (s.bc_field[2 * ax] != BC_PERIODIC) continue;
    const int ng = s.ng, nn = s.n[ax];
    const int64_t sa = s.st[ax];
    int a1 = (ax + 1) % 3, a2 = (ax + 2) % 3;
    int e1 = s.n[a1] + 2 * s.ngd[a1], e2 = s.n[a2] + 2 * s.ngd[a2];
    for (int q = 0; q < e2; ++q)
      for (int p = 0; p < e1; ++p) {
        const int64_t base = (int64_t)p * s.st[a1] + (int64_t)q * s.st[a2];
        for (int i = 1; i <= ng; ++i) {
          rho[base + (i + ng - 1) * sa] += rho[base + (nn + i + ng - 1) * sa];
          rho[base + (nn - ng + i + ng - 1) * sa] += rho[base + (i - ng + ng - 1) * sa];
        }
      }
  }
  double mr = 0.0, mrho = 0.0;
  if (capture) s.res_ref.assign((size_t)s.total, 0.0);
  const bool have_ref = !capture && s.res_ref.size() == (size_t)s.total;
  for (int k = 1; k <= s.n[2]; ++k)
    for (int j = 1; j <= s.n[1]; ++j)
      for (int i = 1; i <= s.n[0]; ++i) {
        const int64_t q = s.idx(i, j, k);
        double div = (s.f[EX][q] - s.f[EX][q - s.st[0]]) / s.d[0];
        if (nd >= 2) div += (s.f[EY][q] - s.f[EY][q - s.st[1]]) / s.d[1];
        if (nd >= 3) div += (s.f[EZ][q] - s.f[EZ][q - s.st[2]]) / s.d[2];
        const double r = rho[q] / EPS0;
        double res = div - r;
        if (capture) s.res_ref[q] = res;
        if (have_ref) res -= s.res_ref[q];
        mr = std::max(mr, std::fabs(res));
        mrho = std::max(mrho, std::fabs(r));
      }
  if (max_res) *max_res = mr;
  if (max_rho) *max_rho = mrho;
}

}  // namespace orc
