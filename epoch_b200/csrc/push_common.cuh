// push_common.cuh -- device helpers shared by the push+deposit kernels.
// Arithmetic follows epochNd/src/particles.F90 and include/*/{gx,hx_dcell}.inc
// (SURVEY Appendix A.3-A.5).
#pragma once
#include "ctx.cuh"

namespace eb {

template <int SHAPE> struct Sh;
template <> struct Sh<SHAPE_TOPHAT>   { static constexpr int NW = 2, SFMIN = 0; };
template <> struct Sh<SHAPE_TRIANGLE> { static constexpr int NW = 3, SFMIN = -1; };
template <> struct Sh<SHAPE_BSPLINE3> { static constexpr int NW = 5, SFMIN = -2; };

// un-normalised 1-D weights W_s(f), slots SFMIN..SFMIN+NW-1
template <int SHAPE>
__device__ __forceinline__ void shape_weights(double f, double* w) {
  if (SHAPE == SHAPE_TOPHAT) {            // tophat/gx.inc:3-4
    w[0] = 0.5 + f;
    w[1] = 0.5 - f;
  } else if (SHAPE == SHAPE_TRIANGLE) {   // triangle/gx.inc:1-4
    const double cf2 = f * f;
    w[0] = 0.25 + cf2 + f;
    w[1] = 1.5 - 2.0 * cf2;
    w[2] = 0.25 + cf2 - f;
  } else {                                // bspline3/gx.inc:3-10
    const double cf2 = f * f;
    double a = 0.5 + f; a = a * a;
    w[0] = a * a;
    w[1] = 4.75 + 11.0 * f + 4.0 * cf2 * (1.5 - f - cf2);
    w[2] = 14.375 + 6.0 * cf2 * (cf2 - 2.5);
    w[3] = 4.75 - 11.0 * f + 4.0 * cf2 * (1.5 + f - cf2);
    double b = 0.5 - f; b = b * b;
    w[4] = b * b;
  }
}

// per-launch constants of push_particles (particles.F90:153-212)
struct PushP {
  long long n;
  double idx[3];
  double gml[3];
  double dto2, dtco2;
  double cmratio, ccmratio, part_mc, ipart_mc, q;
  double fc[3];          // idtyz,idtxz,idtxy (3D) | idty,idtx,idxy (2D) | idtf,idxf,- (1D), without weight
  int zero_current;
};

__host__ inline double shape_fac(int shape, int nd) {   // particles.F90:128-136
  double fac = 1.0;
  if (shape == SHAPE_BSPLINE3) for (int d = 0; d < nd; ++d) fac *= (1.0 / 24.0);
  else if (shape == SHAPE_TRIANGLE) for (int d = 0; d < nd; ++d) fac *= 0.5;
  return fac;
}

// cell coordinate x*idx [-0.5 tophat] (particles.F90:270-291), rounded as the reference
// rounds it (two roundings, no FMA contraction) so that FLOOR lands in the same cell
template <int SHAPE>
__device__ __forceinline__ double cell_coord(double px, double idx) {
  const double r = __dmul_rn(px, idx);
  return SHAPE == SHAPE_TOPHAT ? __dadd_rn(r, -0.5) : r;
}
// nearest-node cell and fraction: cell = FLOOR(r+0.5); frac = cell - r; cell+1
__device__ __forceinline__ int cell_round(double r, double& frac) {
  const double fl = floor(r + 0.5);
  frac = fl - r;
  return (int)fl + 1;
}
// staggered cell: cell = FLOOR(r); frac = cell - r + 0.5; cell+1
__device__ __forceinline__ int cell_floor(double r, double& frac) {
  const double fl = floor(r);
  frac = fl - r + 0.5;
  return (int)fl + 1;
}

// half drift to t+dt/2 (particles.F90:242-257) with every rounding pinned by
// intrinsics, so that the sort-key kernel and the push kernels place a particle in
// the same cell bit for bit.  pos[] are absolute positions, out: px[] relative to
// x_grid_min_local, u = p / (m c).
template <int ND>
__device__ __forceinline__ void half_drift(const PushP& p, const double* pos, double mx, double my, double mz,
                                           double* px, double& ux, double& uy, double& uz) {
  ux = __dmul_rn(mx, p.ipart_mc); uy = __dmul_rn(my, p.ipart_mc); uz = __dmul_rn(mz, p.ipart_mc);
  const double g2 = __fma_rn(uz, uz, __fma_rn(uy, uy, __fma_rn(ux, ux, 1.0)));
#ifdef EB_EXACT_DIV
  const double root = __ddiv_rn(p.dtco2, __dsqrt_rn(g2));
#else
  const double root = __dmul_rn(p.dtco2, rsqrt(g2));     // (c dt / 2) / gamma, ~1 ulp; the same function keys and pushes
#endif
  const double u0[3] = {ux, uy, uz};
#pragma unroll
  for (int d = 0; d < ND; ++d) px[d] = __fma_rn(u0[d], root, __dadd_rn(pos[d], -p.gml[d]));
}

// tile-major sort key of the (1-based) cell triple, clamped into the extended grid
template <int ND>
__device__ __forceinline__ uint32_t cell_key(const GridP& g, const int* cell) {
  uint32_t tile = 0, tmul = 1, loc = 0, lmul = 1;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    int e = cell[d] - (1 - KM);
    e = e < 0 ? 0 : (e >= g.kn[d] ? g.kn[d] - 1 : e);
    const int td = tile_dim(ND, d);
    tile += tmul * (uint32_t)(e / td); tmul *= (uint32_t)g.ntile[d];
    loc += lmul * (uint32_t)(e % td); lmul *= (uint32_t)td;
  }
  return tile * (uint32_t)g.cpt + loc;
}

// particle boundary condition of one axis (particle_bcs, boundary.F90:1316-1972)
struct BcP {
  int axis, min_bnd, max_bnd, bc_min, bc_max, multi;   // multi: nprocdir[axis] > 1
  int cpml_min, cpml_max;   // the FIELD boundary of that side is a CPML: particles are only removed beyond *_outer
  double min_local, max_local, gmin, gmax, min_outer, max_outer, shift;
};
struct BcP3 { BcP b[3]; int on[3]; };

// particle_bcs of the axes that are not decomposed (periodic wrap, reflection, open-boundary
// removal; boundary.F90:1390-1461: the min side tests <, the max side >=), then the sort key of
// the cell the next push gathers from.  pos/mom are updated in place; returns -1 when the
// particle left through an open boundary, else a bit mask of the axes that were changed.
template <int ND, int SHAPE>
__device__ __forceinline__ int bc_and_key(const GridP& g, const PushP& p, const BcP3& bc, double* pos, double* mom,
                                          uint32_t& key) {
  int changed = 0;
  bool out = false;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    if (!bc.on[d]) continue;
    const BcP& b = bc.b[d];
    const double part_pos = pos[d];
    if (b.cpml_min) {                       // boundary.F90:1390-1402: left the system through the CPML
      if (part_pos < b.min_outer) out = true;
    } else if (part_pos < b.min_local) {
      if (b.bc_min == EB200_BC_REFLECT) { pos[d] = 2.0 * b.gmin - part_pos; mom[d] = -mom[d]; changed |= 1 << d; }
      else if (b.bc_min == EB200_BC_PERIODIC) { pos[d] = part_pos + b.shift; changed |= 1 << d; }
      if (part_pos < b.min_outer && b.bc_min != EB200_BC_PERIODIC) out = true;
    }
    if (b.cpml_max) {
      if (part_pos >= b.max_outer) out = true;
    } else if (part_pos >= b.max_local) {
      if (b.bc_max == EB200_BC_REFLECT) { pos[d] = 2.0 * b.gmax - part_pos; mom[d] = -mom[d]; changed |= 1 << d; }
      else if (b.bc_max == EB200_BC_PERIODIC) { pos[d] = part_pos - b.shift; changed |= 1 << d; }
      if (part_pos >= b.max_outer && b.bc_max != EB200_BC_PERIODIC) out = true;
    }
  }
  if (out) return -1;
  double px[3] = {0, 0, 0}, ux, uy, uz;
  half_drift<ND>(p, pos, mom[0], mom[1], mom[2], px, ux, uy, uz);
  int cell[3] = {1, 1, 1};
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    double frac;
    cell[d] = cell_round(cell_coord<SHAPE>(px[d], p.idx[d]), frac);   // 1-based
  }
  key = cell_key<ND>(g, cell);
  return changed;
}

// Boris rotation (particles.F90:347-394); u is updated in place
__device__ __forceinline__ void boris(double& ux, double& uy, double& uz, double ex, double ey, double ez,
                                      double bx, double by, double bz, double cmratio, double ccmratio) {
  const double uxm = ux + cmratio * ex, uym = uy + cmratio * ey, uzm = uz + cmratio * ez;
#ifdef EB_EXACT_DIV
  const double gamma_rel = sqrt(uxm * uxm + uym * uym + uzm * uzm + 1.0);
  const double root = ccmratio / gamma_rel;
#else
  const double root = ccmratio * rsqrt(uxm * uxm + uym * uym + uzm * uzm + 1.0);
#endif
  const double taux = bx * root, tauy = by * root, tauz = bz * root;
  const double taux2 = taux * taux, tauy2 = tauy * tauy, tauz2 = tauz * tauz;
  const double tau = 1.0 / (1.0 + taux2 + tauy2 + tauz2);
  const double uxp = ((1.0 + taux2 - tauy2 - tauz2) * uxm
      + 2.0 * ((taux * tauy + tauz) * uym + (taux * tauz - tauy) * uzm)) * tau;
  const double uyp = ((1.0 - taux2 + tauy2 - tauz2) * uym
      + 2.0 * ((tauy * tauz + taux) * uzm + (tauy * taux - tauz) * uxm)) * tau;
  const double uzp = ((1.0 - taux2 - tauy2 + tauz2) * uzm
      + 2.0 * ((tauz * taux + tauy) * uxm + (tauz * tauy - taux) * uym)) * tau;
  ux = uxp + cmratio * ex;
  uy = uyp + cmratio * ey;
  uz = uzp + cmratio * ez;
}

}  // namespace eb
