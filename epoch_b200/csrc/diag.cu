// diag.cu -- derived grid variables of io/calc_df.F90 on the device: the particle
// moments output_routines asks for are reduced onto the grid where the particles live, so the
// host pulls one field-shaped array instead of the particle lists.
//   calc_mass_density :35, calc_ekbar :119, calc_ekflux :232, calc_poynt_flux :382,
//   calc_charge_density :439, calc_number_density :523, calc_ppc :598, calc_average_weight :652,
//   calc_temperature :723, calc_per_species_current :994, calc_average_momentum :1104
// One scatter kernel serves every moment: a thread per particle evaluates the weights of
// include/particle_to_grid.inc and adds gx*gy*gz*wdata (and, where the routine normalises, the
// weight sum) with fp64 RED atomics.  The particles are cell sorted, so the atomics of a warp
// fall on a few neighbouring lines and are merged in L2; the pass reads 24-56 B per particle and
// is a diagnostic (once per dump), not part of the step.
#include "ctx.cuh"
#include "push_common.cuh"

namespace eb {

namespace {

constexpr double C_TINY = 2.2250738585072014e-308;   // TINY(1.0_num), constants.F90:29
constexpr double PI_ = 3.141592653589793238462643383279503;
constexpr double MU0 = 4.e-7 * PI_;

// normalised weights (include/*/gxfac.inc)
template <int SHAPE>
__device__ __forceinline__ void gxfac(double f, double* w) {
  if (SHAPE == SHAPE_TOPHAT) { w[0] = 0.5 + f; w[1] = 0.5 - f; }
  else if (SHAPE == SHAPE_TRIANGLE) {
    const double c2 = f * f;
    w[0] = 0.5 * (0.25 + c2 + f); w[1] = 0.75 - c2; w[2] = 0.5 * (0.25 + c2 - f);
  } else {
    const double fac1 = 1.0 / 24.0, fac2 = 1.0 / 6.0, fac3 = 115.0 / 48.0;
    const double c2 = f * f;
    double a = 0.5 + f; a = a * a; w[0] = fac1 * (a * a);
    w[1] = fac2 * (1.1875 + 2.75 * f + c2 * (1.5 - f - c2));
    w[2] = 0.25 * (fac3 + c2 * (c2 - 2.5));
    w[3] = fac2 * (1.1875 - 2.75 * f + c2 * (1.5 + f - c2));
    double b = 0.5 - f; b = b * b; w[4] = fac1 * (b * b);
  }
}

struct DiagP {
  long long n;
  int kind, dir;          // dir: signed direction (0 = absent)
  int pass;               // temperature: 1 = means, 2 = variance
  double mass, charge, afac;
  const double* a[7];     // x y z px py pz w
  double* out;            // data_array / sigma
  double* wt;             // weight sum / part_count (or nullptr)
  double* mean[3];        // temperature means
};

template <int ND, int SHAPE>
__global__ void __launch_bounds__(128) k_derived(GridP g, DiagP p) {
  constexpr int NW = Sh<SHAPE>::NW, SFM = Sh<SHAPE>::SFMIN;
  const long long ip = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (ip >= p.n) return;
  const double pw = p.a[6][ip];
  if (p.kind == EB200_DF_PPC || p.kind == EB200_DF_AVERAGE_WEIGHT) {
    // the cell the particle is counted in (:627-639)
    int cell[3] = {1, 1, 1};
#pragma unroll
    for (int d = 0; d < ND; ++d) {
      double cr = (p.a[d][ip] - g.gml[d]) / g.d[d];
      if (SHAPE != SHAPE_TOPHAT) cr += 0.5;
      cell[d] = (int)floor(cr) + 1;
    }
    const long long q = g.idx(cell[0], cell[1], cell[2]);
    if (p.kind == EB200_DF_PPC) atomicAdd(p.out + q, 1.0);
    else { atomicAdd(p.out + q, pw); atomicAdd(p.wt + q, 1.0); }
    return;
  }
  int cell[3] = {1, 1, 1};
  double w[3][NW];
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    double cr = (p.a[d][ip] - g.gml[d]) / g.d[d];
    if (SHAPE == SHAPE_TOPHAT) cr -= 0.5;
    double frac;
    cell[d] = cell_round(cr, frac);
    gxfac<SHAPE>(frac, w[d]);
  }
  const double mom[3] = {p.a[3][ip], p.a[4][ip], p.a[5][ip]};
  double wdata = 0.0, wsum = pw;
  double pm[3] = {0.0, 0.0, 0.0};
  const int ad = p.dir < 0 ? -p.dir : p.dir;
  switch (p.kind) {
    case EB200_DF_MASS_DENSITY: wdata = p.mass * pw; break;
    case EB200_DF_CHARGE_DENSITY: wdata = p.charge * pw; break;
    case EB200_DF_NUMBER_DENSITY: wdata = pw; break;
    case EB200_DF_EKBAR:
    case EB200_DF_EKFLUX: {
      const double part_mc = C0 * p.mass;
      const double fac = part_mc * pw * C0;
      const double ux = mom[0] / part_mc, uy = mom[1] / part_mc, uz = mom[2] / part_mc;
      const double u2 = ux * ux + uy * uy + uz * uz;
      const double gamma_rel = sqrt(u2 + 1.0);
      wdata = u2 / (gamma_rel + 1.0) * fac;
      if (p.kind == EB200_DF_EKFLUX) {
        const double ua = ad == 1 ? ux : ad == 2 ? uy : uz;
        const double flux = p.afac * ua / gamma_rel;
        wdata = p.dir < 0 ? -wdata * fmin(flux, 0.0) : wdata * fmax(flux, 0.0);
      }
      break;
    }
    case EB200_DF_SPECIES_CURRENT: {
      const double part_mc = C0 * p.mass;
      const double root = 1.0 / sqrt(part_mc * part_mc + mom[0] * mom[0] + mom[1] * mom[1] + mom[2] * mom[2]);
      wdata = p.charge * pw * mom[ad - 1] * root;
      break;
    }
    case EB200_DF_AVERAGE_MOMENTUM: wdata = pw * mom[ad - 1]; break;
    case EB200_DF_TEMPERATURE: {
      const double sqm = sqrt(p.mass);
      pm[0] = mom[0] / sqm; pm[1] = mom[1] / sqm; pm[2] = mom[2] / sqm;
      break;
    }
  }
#pragma unroll
  for (int k = 0; k < (ND >= 3 ? NW : 1); ++k)
#pragma unroll
    for (int j = 0; j < (ND >= 2 ? NW : 1); ++j)
#pragma unroll
      for (int i = 0; i < NW; ++i) {
        const long long q = g.idx(cell[0] + SFM + i, ND >= 2 ? cell[1] + SFM + j : 1, ND >= 3 ? cell[2] + SFM + k : 1);
        double gf = w[0][i];
        if (ND >= 2) gf = gf * w[1][j];
        if (ND >= 3) gf = gf * w[2][k];
        if (p.kind == EB200_DF_TEMPERATURE) {
          if (p.pass == 1) {
            const double gw = gf * pw;
#pragma unroll
            for (int m = 0; m < 3; ++m)
              if (ad == 0 || ad == m + 1) atomicAdd(p.mean[m] + q, gw * pm[m]);
            atomicAdd(p.wt + q, gw);
          } else {
            double s2 = 0.0;
#pragma unroll
            for (int m = 0; m < 3; ++m)
              if (ad == 0 || ad == m + 1) { const double dq = pm[m] - p.mean[m][q]; s2 = s2 + dq * dq; }
            atomicAdd(p.out + q, gf * s2);
            atomicAdd(p.wt + q, gf);
          }
        } else {
          atomicAdd(p.out + q, gf * wdata);
          if (p.wt) atomicAdd(p.wt + q, gf * wsum);
        }
      }
}

__global__ void k_scale(double* a, long long n, double s) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = a[i] * s;
}
// a = a / max(b, floor) [* s1 * s2 applied as two divisions when div2 is set, as the reference writes it]
__global__ void k_divide(double* a, const double* b, long long n, double floor_, double d1, double d2, int div2) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double v = a[i] / fmax(b[i], floor_);
  if (div2) v = v / d1 / d2;
  a[i] = v;
}
// part_count = max(part_count, 1e-6); mean* = mean* / part_count (:877-880)
__global__ void k_temp_means(double* m0, double* m1, double* m2, double* cnt, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double c = fmax(cnt[i], 1.e-6);
  m0[i] = m0[i] / c; m1[i] = m1[i] / c; m2[i] = m2[i] / c;
}

// calc_poynt_flux: E_a is staggered along a, B_a along the other two axes; a component is
// centred by averaging over the staggered axes that exist
struct PoyntP { const double* f[6]; double* out; int dir; };
template <int ND>
__device__ __forceinline__ double centre(const GridP& g, const double* f, int comp, long long i) {
  const bool isb = comp >= 3;
  const int a = comp % 3;
  long long st[2];
  int n = 0;
#pragma unroll
  for (int d = 0; d < ND; ++d)
    if (isb ? d != a : d == a) st[n++] = g.st[d];
  if (n == 0) return f[i];
  if (n == 1) return 0.5 * (f[i - st[0]] + f[i]);
  return 0.25 * (f[i - st[0] - st[1]] + f[i - st[1]] + f[i - st[0]] + f[i]);
}
template <int ND>
__global__ void k_poynt(GridP g, PoyntP p, long long count) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count) return;
  const int ix = (int)(t % g.n[0]) + 1;
  const long long r = t / g.n[0];
  const int iy = ND >= 2 ? (int)(r % g.n[1]) + 1 : 1;
  const int iz = ND >= 3 ? (int)(r / g.n[1]) + 1 : 1;
  const long long i = g.idx(ix, iy, iz);
  const int a = p.dir - 1, b = (a + 1) % 3, c = (a + 2) % 3;
  p.out[i] = (centre<ND>(g, p.f[b], b, i) * centre<ND>(g, p.f[3 + c], 3 + c, i) -
              centre<ND>(g, p.f[c], c, i) * centre<ND>(g, p.f[3 + b], 3 + b, i)) / MU0;
}

}  // namespace

int32_t calc_derived(eb200_ctx* c, int kind, int ispecies, int direction, double* out_host) {
  const GridP& g = c->g;
  const size_t tot = (size_t)g.total;
  if (!c->diag_wk) EB_CUDA(c, cudaMalloc(&c->diag_wk, 5 * tot * sizeof(double)));
  double* out = c->diag_wk;
  double* wt = out + tot;
  double* mean[3] = {out + 2 * tot, out + 3 * tot, out + 4 * tot};
  const bool temp = kind == EB200_DF_TEMPERATURE;
  const bool normalised = kind == EB200_DF_EKBAR || kind == EB200_DF_EKFLUX || kind == EB200_DF_AVERAGE_WEIGHT ||
                          kind == EB200_DF_AVERAGE_MOMENTUM;
  EB_CUDA(c, cudaMemsetAsync(out, 0, (temp ? 5 : 2) * tot * sizeof(double), c->stream));
  const unsigned gblocks = (unsigned)((tot + 255) / 256);
  const int ad = direction < 0 ? -direction : direction;

  if (kind == EB200_DF_POYNT_FLUX) {
    PoyntP p;
    for (int k = 0; k < 6; ++k) p.f[k] = c->f.f[k];
    p.out = out; p.dir = direction;
    long long count = 1;
    for (int d = 0; d < c->ndims; ++d) count *= g.n[d];
    const unsigned nb = (unsigned)((count + 255) / 256);
    if (c->ndims == 1) k_poynt<1><<<nb, 256, 0, c->stream>>>(g, p, count);
    if (c->ndims == 2) k_poynt<2><<<nb, 256, 0, c->stream>>>(g, p, count);
    if (c->ndims == 3) k_poynt<3><<<nb, 256, 0, c->stream>>>(g, p, count);
    c->launches++;
  } else {
    const bool spec_sum = ispecies < 0;
    const int s0 = spec_sum ? 0 : ispecies, s1 = spec_sum ? (int)c->species.size() : ispecies + 1;
    for (int pass = 1; pass <= (temp ? 2 : 1); ++pass) {
      for (int is = s0; is < s1; ++is) {
        SpeciesD& sp = c->species[is];
        if (!sp.defined || sp.n == 0) continue;
        if (spec_sum && sp.zero_current) continue;
        DiagP p;
        p.n = sp.n; p.kind = kind; p.dir = direction; p.pass = pass;
        p.mass = sp.mass; p.charge = sp.charge;
        p.afac = C0;                            // xfac = c * dy * dz and cyclic (:258-260)
        for (int d = 0; d < c->ndims; ++d) if (d != ad - 1) p.afac = p.afac * g.d[d];
        for (int k = 0; k < 7; ++k) p.a[k] = sp.a[k];
        p.out = out;
        p.wt = (normalised || temp) ? wt : nullptr;
        for (int k = 0; k < 3; ++k) p.mean[k] = mean[k];
        const unsigned nb = (unsigned)((sp.n + 127) / 128);
#define L(ND, SH) if (c->ndims == ND && c->shape == SH) k_derived<ND, SH><<<nb, 128, 0, c->stream>>>(g, p);
        L(1, 0) L(1, 1) L(1, 2) L(2, 0) L(2, 1) L(2, 2) L(3, 0) L(3, 1) L(3, 2)
#undef L
        c->launches++;
      }
      if (temp && pass == 1) {
        // calc_boundary on the means and the count, normalise, restore the ghosts of the means
        for (int k = 0; k < 3; ++k)
          if (ad == 0 || ad == k + 1) EB_TRY(summation_bcs(c, mean[k], -1));
        EB_TRY(summation_bcs(c, wt, -1));
        k_temp_means<<<gblocks, 256, 0, c->stream>>>(mean[0], mean[1], mean[2], wt, (long long)tot);
        c->launches++;
        for (int k = 0; k < 3; ++k)
          if (ad == 0 || ad == k + 1) EB_TRY(field_bc_array(c, mean[k]));
        EB_CUDA(c, cudaMemsetAsync(wt, 0, tot * sizeof(double), c->stream));
      }
    }
    if (kind == EB200_DF_PPC) {
      // counts only, no boundary sums (:598-648)
    } else if (kind == EB200_DF_AVERAGE_WEIGHT) {
      k_divide<<<gblocks, 256, 0, c->stream>>>(out, wt, (long long)tot, C_TINY, 1.0, 1.0, 0);
      c->launches++;
    } else {
      EB_TRY(summation_bcs(c, out, -1));
      if (temp) {
        EB_TRY(summation_bcs(c, wt, -1));
        k_divide<<<gblocks, 256, 0, c->stream>>>(out, wt, (long long)tot, 1.e-6, KB, ad ? 1.0 : 3.0, 1);
        c->launches++;
      } else {
        if (normalised) {
          EB_TRY(summation_bcs(c, wt, -1));
          k_divide<<<gblocks, 256, 0, c->stream>>>(out, wt, (long long)tot, C_TINY, 1.0, 1.0, 0);
        } else {
          double idx = 1.0;
          for (int d = 0; d < c->ndims; ++d) idx = idx / g.d[d];
          k_scale<<<gblocks, 256, 0, c->stream>>>(out, (long long)tot, kind == EB200_DF_SPECIES_CURRENT ? C0 * idx : idx);
        }
        c->launches++;
        EB_TRY(zero_gradient_centre(c, out));
      }
    }
  }
  EB_CUDA(c, cudaGetLastError());
  const char* dev = (const char*)out;
  EB_TRY(staged_d2h(c, out_host, tot * sizeof(double), sizeof(double),
                    [dev](size_t o, size_t) { return (const void*)(dev + o); }));
  EB_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

// ---------------------------------------------------------------------------------
// general_dist_fn (io/dist_fn.F90:118-762): the particle loop (:532-628).  Thread per particle,
// one fp64 RED per particle into the histogram.
// ---------------------------------------------------------------------------------
struct DistP {
  long long n;
  const double* a[7];
  int nd, curdims, photon;
  int dir[3];                 // 1-based c_dir_* codes
  double lo[3], dgrid[3];
  int res[3];
  double ang_max[3];          // xy_max, yz_max, zx_max (:505-510)
  int use_ang[3];
  int use_r[12];
  double r_lo[12], r_hi[12];
  double part_mc, part_mc2;
  double* out;
};

__global__ void k_dist_fn(DistP p) {
  const long long ip = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (ip >= p.n) return;
  const double PI_ = 3.141592653589793238462643383279503, pi2 = 2.0 * PI_;
  const int nd = p.nd;
  double data[13];            // particle_data(1 : c_df_maxdirs), index = code
#pragma unroll
  for (int k = 0; k < 13; ++k) data[k] = 0.0;
  for (int d = 0; d < nd; ++d) data[1 + d] = p.a[d][ip];
  const double px = p.a[3][ip], py = p.a[4][ip], pz = p.a[5][ip];
  const double ux = px / p.part_mc, uy = py / p.part_mc, uz = pz / p.part_mc;
  const double part_u2 = ux * ux + uy * uy + uz * uz;
  const double gamma_rel = sqrt(part_u2 + 1.0);
  const double gamma_rel_m1 = part_u2 / (gamma_rel + 1.0);
  if (p.photon) {             // (:554-566 without PHOTONS / BREMSSTRAHLUNG)
    data[nd + EB200_DIR_GAMMA_M1] = 1.0;
  } else {
    data[nd + EB200_DIR_PX] = px; data[nd + EB200_DIR_PY] = py; data[nd + EB200_DIR_PZ] = pz;
    data[nd + EB200_DIR_EN] = gamma_rel_m1 * p.part_mc2;
    data[nd + EB200_DIR_GAMMA_M1] = gamma_rel_m1;
    data[nd + EB200_DIR_MOD_P] = sqrt(px * px + py * py + pz * pz);
  }
  const double pa[3] = {px, py, pz};
#pragma unroll
  for (int k = 0; k < 3; ++k) {          // xy: (px, py), yz: (py, pz), zx: (pz, px)   (:576-604)
    if (!p.use_ang[k]) continue;
    const double a0 = pa[k], a1 = pa[(k + 1) % 3];
    const double pp = sqrt(a0 * a0 + a1 * a1);
    if (pp < C_TINY) return;
    double theta = asin(a1 / pp);
    if (a0 < 0.0) theta = copysign(1.0, a1) * PI_ - theta;
    if ((theta + pi2) < p.ang_max[k]) theta = theta + pi2;
    data[nd + EB200_DIR_XY_ANGLE + k] = theta;
  }
  for (int idir = 1; idir <= nd + 9; ++idir)
    if (p.use_r[idir - 1] && (data[idir] < p.r_lo[idir - 1] || data[idir] > p.r_hi[idir - 1])) return;
  long long idx = 0, mul = 1;
  for (int idim = 0; idim < p.curdims; ++idim) {
    const double cur = data[p.dir[idim]];
    const double cf = floor((cur - p.lo[idim]) / p.dgrid[idim]) + 1.0;
    if (!(cf >= 1.0) || cf > (double)p.res[idim]) return;
    idx += mul * ((long long)cf - 1);
    mul *= p.res[idim];
  }
  atomicAdd(p.out + idx, p.a[6][ip]);
}

int32_t dist_fn(eb200_ctx* c, int ispecies, int curdims, const int32_t* direction, const double* ranges,
                const int32_t* resolution, const double* dgrid, const double* restrictions,
                const int32_t* use_restrictions, double* out_host) {
  SpeciesD& sp = c->species[ispecies];
  const int nd = c->ndims;
  DistP p;
  p.n = sp.n; p.nd = nd; p.curdims = curdims; p.photon = sp.species_type == 1;
  for (int k = 0; k < 7; ++k) p.a[k] = sp.a[k];
  long long tot = 1;
  for (int k = 0; k < 3; ++k) { p.use_ang[k] = 0; p.ang_max[k] = 0.0; p.dir[k] = 1; p.lo[k] = 0; p.dgrid[k] = 1; p.res[k] = 1; }
  for (int i = 0; i < curdims; ++i) {
    p.dir[i] = direction[i]; p.lo[i] = ranges[2 * i]; p.dgrid[i] = dgrid[i]; p.res[i] = resolution[i];
    tot *= resolution[i];
    for (int k = 0; k < 3; ++k)
      if (direction[i] == nd + EB200_DIR_XY_ANGLE + k) { p.use_ang[k] = 1; p.ang_max[k] = ranges[2 * i + 1]; }
  }
  for (int k = 0; k < 12; ++k) {
    p.use_r[k] = (use_restrictions && k < nd + 9) ? (use_restrictions[k] != 0) : 0;
    p.r_lo[k] = (restrictions && k < nd + 9) ? restrictions[2 * k] : 0.0;
    p.r_hi[k] = (restrictions && k < nd + 9) ? restrictions[2 * k + 1] : 0.0;
  }
  p.part_mc = sp.mass * C0; p.part_mc2 = p.part_mc * C0;
  double* dev = nullptr;
  EB_CUDA(c, cudaMalloc(&dev, (size_t)tot * sizeof(double)));
  EB_CUDA(c, cudaMemsetAsync(dev, 0, (size_t)tot * sizeof(double), c->stream));
  p.out = dev;
  if (sp.n > 0) {
    k_dist_fn<<<(unsigned)((sp.n + 255) / 256), 256, 0, c->stream>>>(p);
    c->launches++;
  }
  int32_t rc = cudaGetLastError() == cudaSuccess ? 0 : EB200_E_CUDA;
  const char* d8 = (const char*)dev;
  if (rc == 0) rc = staged_d2h(c, out_host, (size_t)tot * sizeof(double), sizeof(double),
                               [d8](size_t o, size_t) { return (const void*)(d8 + o); });
  cudaStreamSynchronize(c->stream);
  cudaFree(dev);
  return rc;
}

}  // namespace eb
