// push_baseline.cu -- reference-order push + deposit, one thread per particle,
// fields gathered through the read-only path and currents scattered with
// RED.ADD.F64 straight to L2.  This is the general kernel: every ndims x shape
// combination, any particle order, any boundary.  It restates particles.F90
// (:214-608) statement by statement (running sums included) and is the kernel
// the tiled fast path (push_tiled.cu) is checked against on the device.
#include "push_common.cuh"

namespace eb {

template <int ND, int SHAPE>
__device__ __forceinline__ double gather1(const GridP& g, const double* __restrict__ F,
                                          const double* wx, const double* wy, const double* wz,
                                          int cx, int cy, int cz) {
  constexpr int NW = Sh<SHAPE>::NW, SFM = Sh<SHAPE>::SFMIN;
  const long long base = g.idx(cx + SFM, ND >= 2 ? cy + SFM : 1, ND >= 3 ? cz + SFM : 1);
  double accz = 0.0;
#pragma unroll
  for (int k = 0; k < (ND >= 3 ? NW : 1); ++k) {
    double accy = 0.0;
#pragma unroll
    for (int j = 0; j < (ND >= 2 ? NW : 1); ++j) {
      double accx = 0.0;
#pragma unroll
      for (int i = 0; i < NW; ++i) {
        const double t = wx[i] * __ldg(F + base + i + j * g.st[1] + k * g.st[2]);
        accx = i == 0 ? t : accx + t;
      }
      if (ND >= 2) { const double t = wy[j] * accx; accy = j == 0 ? t : accy + t; }
      else accy = accx;
    }
    if (ND >= 3) { const double t = wz[k] * accy; accz = k == 0 ? t : accz + t; }
    else accz = accy;
  }
  return accz;
}

template <int ND, int SHAPE>
__global__ void __launch_bounds__(128) k_push_baseline(GridP g, FieldP F, PushP p,
    double* __restrict__ X, double* __restrict__ Y, double* __restrict__ Z,
    double* __restrict__ PX, double* __restrict__ PY, double* __restrict__ PZ,
    const double* __restrict__ W, const ProbeSet pr) {
  constexpr int NW = Sh<SHAPE>::NW, SFM = Sh<SHAPE>::SFMIN, NS = NW + 2;
  const long long ip = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (ip >= p.n) return;
  double* POS[3] = {X, Y, Z};
  const double part_weight = W[ip];
  double px[3] = {0, 0, 0}, pos0[3] = {0, 0, 0};
#pragma unroll
  for (int d = 0; d < ND; ++d) pos0[d] = POS[d][ip];
  double ux, uy, uz, gamma_rel, root;
  half_drift<ND>(p, pos0, PX[ip], PY[ip], PZ[ip], px, ux, uy, uz);
  int cell1[3] = {1, 1, 1}, cell2[3] = {1, 1, 1};
  double wg[3][NW], wh[3][NW];
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    const double cr = cell_coord<SHAPE>(px[d], p.idx[d]);
    double frac;
    cell1[d] = cell_round(cr, frac);
    shape_weights<SHAPE>(frac, wg[d]);
    cell2[d] = cell_floor(cr, frac);
    shape_weights<SHAPE>(frac, wh[d]);
  }
  // stagger table: ex (h,g,g) ey (g,h,g) ez (g,g,h) bx (g,h,h) by (h,g,h) bz (h,h,g)
  const double ex_part = gather1<ND, SHAPE>(g, F.f[EX], wh[0], wg[1], wg[2], cell2[0], cell1[1], cell1[2]);
  const double ey_part = gather1<ND, SHAPE>(g, F.f[EY], wg[0], wh[1], wg[2], cell1[0], cell2[1], cell1[2]);
  const double ez_part = gather1<ND, SHAPE>(g, F.f[EZ], wg[0], wg[1], wh[2], cell1[0], cell1[1], cell2[2]);
  const double bx_part = gather1<ND, SHAPE>(g, F.f[BX], wg[0], wh[1], wh[2], cell1[0], cell2[1], cell2[2]);
  const double by_part = gather1<ND, SHAPE>(g, F.f[BY], wh[0], wg[1], wh[2], cell2[0], cell1[1], cell2[2]);
  const double bz_part = gather1<ND, SHAPE>(g, F.f[BZ], wh[0], wh[1], wg[2], cell2[0], cell2[1], cell1[2]);

  boris(ux, uy, uz, ex_part, ey_part, ez_part, bx_part, by_part, bz_part, p.cmratio, p.ccmratio);

  const double part_u2 = ux * ux + uy * uy + uz * uz;
  gamma_rel = sqrt(part_u2 + 1.0);
  double delta[3] = {0, 0, 0};
  double part_vy = 0, part_vz = 0;
  if (ND == 3) {
    root = p.dtco2 / gamma_rel;
    delta[0] = ux * root; delta[1] = uy * root; delta[2] = uz * root;
  } else if (ND == 2) {
    const double igamma = 1.0 / gamma_rel;
    root = p.dtco2 * igamma;
    delta[0] = ux * root; delta[1] = uy * root;
    part_vz = uz * C0 * igamma;
  } else {
    root = C0 / gamma_rel;
    delta[0] = ux * root * p.dto2;
    part_vy = uy * root;
    part_vz = uz * root;
  }
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    px[d] = px[d] + delta[d];
    POS[d][ip] = px[d] + p.gml[d];
  }
  PX[ip] = p.part_mc * ux; PY[ip] = p.part_mc * uy; PZ[ip] = p.part_mc * uz;
  if (pr.n > 0) {
    // particle probes (particles.F90:567-606): energy window, then the plane crossed along its normal
    const double gamma_rel_m1 = part_u2 / (gamma_rel + 1.0);
    const double probe_energy = gamma_rel_m1 * (C0 * p.part_mc);
    for (int k = 0; k < pr.n; ++k) {
      if (!pr.on[k] || !(probe_energy > pr.ek_min[k]) || !(probe_energy < pr.ek_max[k])) continue;
      double d_init = 0.0, d_final = 0.0;
#pragma unroll
      for (int d = 0; d < ND; ++d) {
        d_init += pr.normal[k][d] * (pr.point[k][d] - pos0[d]);
        d_final += pr.normal[k][d] * (pr.point[k][d] - (px[d] + p.gml[d]));
      }
      if (d_final < 0.0 && d_init >= 0.0) {
        const unsigned long long slot = atomicAdd(pr.count + k, 1ULL);
        if ((long long)slot < pr.cap) {
          double* q = pr.buf + ((size_t)k * pr.cap + slot) * (ND + 4);
#pragma unroll
          for (int d = 0; d < ND; ++d) q[d] = px[d] + p.gml[d];
          q[ND] = p.part_mc * ux; q[ND + 1] = p.part_mc * uy; q[ND + 2] = p.part_mc * uz; q[ND + 3] = part_weight;
        }
      }
    }
  }
  if (p.zero_current) return;

  // ---- deposit (particles.F90:443-563) ----
  double gs[3][NS], hs[3][NS];
  int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    px[d] = px[d] + delta[d];
    const double cr = cell_coord<SHAPE>(px[d], p.idx[d]);
    double frac;
    const int cell3 = cell_round(cr, frac);
    const int dcell = cell3 - cell1[d];
    double w3[NW];
    shape_weights<SHAPE>(frac, w3);
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const double gv = (s >= 1 && s <= NW) ? wg[d][s >= 1 && s <= NW ? s - 1 : 0] : 0.0;
      double hv = 0.0;
#pragma unroll
      for (int i = 0; i < NW; ++i) if (s == dcell + 1 + i) hv = w3[i];
      gs[d][s] = gv;
      hs[d][s] = hv - gv;
    }
    // xmin = sf_min + (dcell-1)/2 ; xmax = sf_max + (dcell+1)/2 (truncating division), as slots
    lo[d] = 1 + (dcell - 1) / 2;
    hi[d] = NW + (dcell + 1) / 2;
  }
  const double third = 1.0 / 3.0;
  double* __restrict__ Jx = F.f[JX]; double* __restrict__ Jy = F.f[JY]; double* __restrict__ Jz = F.f[JZ];
  const double fcx = p.fc[0] * part_weight, fcy = p.fc[1] * part_weight, fcz = p.fc[2] * part_weight;
  if (ND == 3) {
    const double fjx = fcx * p.q, fjy = fcy * p.q, fjz = fcz * p.q;
    double jzh[NS][NS];
#pragma unroll
    for (int a = 0; a < NS; ++a)
#pragma unroll
      for (int b = 0; b < NS; ++b) jzh[a][b] = 0.0;
#pragma unroll
    for (int kz = 0; kz < NS; ++kz) {
      const bool inz = kz >= lo[2] && kz <= hi[2];
      const double gz_iz = gs[2][kz], hz_iz = hs[2][kz];
      const double zfac1 = gz_iz + 0.5 * hz_iz;
      const double zfac2 = third * hz_iz + 0.5 * gz_iz;
      double jyh[NS];
#pragma unroll
      for (int a = 0; a < NS; ++a) jyh[a] = 0.0;
#pragma unroll
      for (int ky = 0; ky < NS; ++ky) {
        const bool iny = inz && ky >= lo[1] && ky <= hi[1];
        const double gyv = gs[1][ky], hyv = hs[1][ky];
        const double yfac1 = gyv + 0.5 * hyv;
        const double yfac2 = third * hyv + 0.5 * gyv;
        const double hygz = hyv * gz_iz, hyhz = hyv * hz_iz;
        const double yzfac = gyv * zfac1 + hyv * zfac2;
        const double hzyfac1 = hz_iz * yfac1, hzyfac2 = hz_iz * yfac2;
        double jxh = 0.0;
#pragma unroll
        for (int kx = 0; kx < NS; ++kx) {
          const bool in = iny && kx >= lo[0] && kx <= hi[0];
          if (in) {
            const double gxv = gs[0][kx], hxv = hs[0][kx];
            const double xfac1 = gxv + 0.5 * hxv;
            const double xfac2 = third * hxv + 0.5 * gxv;
            const double wx = hxv * yzfac;
            const double wy = xfac1 * hygz + xfac2 * hyhz;
            const double wz = gxv * hzyfac1 + hxv * hzyfac2;
            jxh = jxh - fjx * wx;
            jyh[kx] = jyh[kx] - fjy * wy;
            jzh[ky][kx] = jzh[ky][kx] - fjz * wz;
            const long long i = g.idx(cell1[0] + SFM - 1 + kx, cell1[1] + SFM - 1 + ky, cell1[2] + SFM - 1 + kz);
            atomicAdd(Jx + i, jxh);
            atomicAdd(Jy + i, jyh[kx]);
            atomicAdd(Jz + i, jzh[ky][kx]);
          }
        }
      }
    }
  } else if (ND == 2) {
    const double fjx = fcx * p.q, fjy = fcy * p.q, fjz = fcz * p.q * part_vz;
    double jyh[NS];
#pragma unroll
    for (int a = 0; a < NS; ++a) jyh[a] = 0.0;
#pragma unroll
    for (int ky = 0; ky < NS; ++ky) {
      const bool iny = ky >= lo[1] && ky <= hi[1];
      const double gyv = gs[1][ky], hyv = hs[1][ky];
      const double yfac1 = gyv + 0.5 * hyv;
      const double yfac2 = third * hyv + 0.5 * gyv;
      double jxh = 0.0;
#pragma unroll
      for (int kx = 0; kx < NS; ++kx) {
        if (iny && kx >= lo[0] && kx <= hi[0]) {
          const double gxv = gs[0][kx], hxv = hs[0][kx];
          const double xfac1 = gxv + 0.5 * hxv;
          const double wx = hxv * yfac1;
          const double wy = hyv * xfac1;
          const double wz = gxv * yfac1 + hxv * yfac2;
          jxh = jxh - fjx * wx;
          jyh[kx] = jyh[kx] - fjy * wy;
          const double jzh = fjz * wz;
          const long long i = g.idx(cell1[0] + SFM - 1 + kx, cell1[1] + SFM - 1 + ky, 1);
          atomicAdd(Jx + i, jxh);
          atomicAdd(Jy + i, jyh[kx]);
          atomicAdd(Jz + i, jzh);
        }
      }
    }
  } else {
    const double fjx = fcx * p.q, fjy = fcy * p.q * part_vy, fjz = fcy * p.q * part_vz;
    double jxh = 0.0;
#pragma unroll
    for (int kx = 0; kx < NS; ++kx) {
      if (kx >= lo[0] && kx <= hi[0]) {
        const double gxv = gs[0][kx], hxv = hs[0][kx];
        const double wx = hxv;
        const double wy = gxv + 0.5 * hxv;
        jxh = jxh - fjx * wx;
        const long long i = g.idx(cell1[0] + SFM - 1 + kx, 1, 1);
        atomicAdd(Jx + i, jxh);
        atomicAdd(Jy + i, fjy * wy);
        atomicAdd(Jz + i, fjz * wy);
      }
    }
  }
}

int32_t launch_push_baseline(eb200_ctx* c, const SpeciesD& sp, const PushP& p) {
  if (p.n <= 0) return 0;
  const int block = 128;
  const unsigned grid = (unsigned)((p.n + block - 1) / block);
#define L(ND, SH)                                                                              \
  if (c->ndims == ND && c->shape == SH)                                                        \
    k_push_baseline<ND, SH><<<grid, block, 0, c->stream>>>(c->g, c->f, p, sp.a[0], sp.a[1],    \
        sp.a[2], sp.a[3], sp.a[4], sp.a[5], sp.a[6], sp.probes);
  L(1, 0) L(1, 1) L(1, 2) L(2, 0) L(2, 1) L(2, 2) L(3, 0) L(3, 1) L(3, 2)
#undef L
  c->launches++;
  EB_CUDA(c, cudaGetLastError());
  return 0;
}

}  // namespace eb
