// push_tiled.cu -- tiled push + deposit for cell-sorted particles (2D/3D, tophat
// and triangle shapes).  Replaces the per-particle loop of push_particles
// (epoch3d/src/particles.F90:214-608, epoch2d :199-540).
//
// The particles are sorted by a tile-major cell key (3D tiles 8x8x4 cells, 2D 16x16;
// ctx.cuh tile_dim), so a tile's particles are contiguous.  Two kernels run over the
// tiles of a slab, passing a 64 B deposit record per particle through a global scratch:
//   k_push_split, CTA per tile, thread per particle: E,B of the tile + gather halo are
//     staged in shared memory once; half drift, gather (same-cell lanes -> broadcast
//     shared loads), Boris, drift, store, particle BCs + next sort key (single rank);
//     the record (fractions at t+dt/2 and t+3dt/2, cell shift, weight[, v_z]) goes to
//     the scratch through a warp transpose (full-line stores).
//   k_deposit_split, CTA per half tile, thread per CELL: walks the records of its own
//     cell (cp.async ring) and accumulates all three current components on the cell's
//     NW^ND stencil box in REGISTERS -- no atomics, no shuffles: the reduction over the
//     particles of a cell is a serial sum in one thread.  A particle that changed cell
//     (|dcell| = 1 on some axis) reaches one layer beyond the box; those few values
//     are added afterwards by work items (thread per (particle, moving axis)) with
//     shared-memory atomics.  The register boxes are added to the shared J tile with
//     plain read-modify-writes, one stencil offset per barrier step (all threads apply
//     the same offset to distinct cells), and the tile is flushed to HBM once
//     (RED.ADD.F64 -- neighbouring tiles overlap in their halos).
// The Esirkepov sums are evaluated in factored form,
//   Jx(i,j,k) = -fjx * cumsum_i(hx) * (gy(j) zf1(k) + hy(j) zf2(k)),   (cyclic)
// which is the reference's running sum (particles.F90:522-563) with the common
// factor taken out; the result differs from the reference order by rounding only.
#include "push_common.cuh"
#include <cuda.h>          // CUtensorMap and its enums (the encoder is fetched with cudaGetDriverEntryPoint)
#include <algorithm>

namespace eb {

namespace {

#ifndef EB_MINB_A
#define EB_MINB_A 4
#endif
#ifndef EB_C0_SMEM
#define EB_C0_SMEM 1
#endif
#ifndef EB_SUB
#define EB_SUB 2                 // a deposit CTA owns 1/EB_SUB of a tile, split along the slowest tile axis
#endif
template <int ND, int SHAPE>
struct TL {
  static constexpr int NW = Sh<SHAPE>::NW, SFM = Sh<SHAPE>::SFMIN;
  static constexpr int TX = tile_dim(ND, 0), TY = tile_dim(ND, 1), TZ = tile_dim(ND, 2);
  static constexpr int CPT = TX * TY * TZ;
  // gather region: cells c0+SFM-1 .. c0+T-1+SFM+NW-1  (T + NW wide)
  // (x extent rounded up to an even number of doubles: a TMA box row is a multiple of 16 bytes)
#ifndef EB_FXPAD
#define EB_FXPAD 1
#endif
  static constexpr int FX = EB_FXPAD ? ((TX + NW + 1) & ~1) : TX + NW, FY = ND >= 2 ? TY + NW : 1, FZ = ND >= 3 ? TZ + NW : 1;
  static constexpr int FN = FX * FY * FZ;
  // shared layout of the staged E,B: [z][component][y][x].  With FX = 12 (3D triangle) the strides of a step in
  // x, y, z are 1, 12, 6 * 132 = 8 (mod 16 doubles), so the 2 x 2 x 2 corner offsets of neighbouring
  // particles (g- and h-type cells) and neighbouring cells in x fall into distinct banks.
#ifndef EB_ZCOMP
#define EB_ZCOMP 1
#endif
  static constexpr int FCS = EB_ZCOMP ? FX * FY : FN;         // component stride  (EB_ZCOMP = 0: [component][z][y][x])
  static constexpr int FZS = EB_ZCOMP ? 6 * FX * FY : FX * FY;   // z stride
  // deposit region: cells c0+SFM-1 .. c0+T-1+SFM+NW  (T + NW + 1 wide)
  static constexpr int JX = TX + NW + 1, JY = ND >= 2 ? TY + NW + 1 : 1, JZ = ND >= 3 ? TZ + NW + 1 : 1;
  static constexpr int JN = JX * JY * JZ;
  // ... of a deposit CTA's sub-tile (half a tile along the slowest axis)
  static constexpr int JNS = ND == 3 ? JX * JY * (TZ / EB_SUB + NW + 1) : (ND == 2 ? JX * (TY / EB_SUB + NW + 1) : TX / EB_SUB + NW + 1);
  // register accumulators per component.  3D: the last box slot along the running-sum axis is
  // left out -- the sum has returned to zero there (rounding residue) unless the particle moved
  // up a cell, and then the slot is deposited by a work item (mover_item3, m == A, dcell = +1)
  static constexpr int NACC = ND == 3 ? (NW - 1) * NW * NW : (ND == 2 ? NW * NW : NW);
};

// deposit record: doubles f1[ND], f3[ND], weight[, v_z] and an int = packed cell shift
// (2 bits per axis, value dcell + 1) | tile-local cell << 8.  Two views:
struct RecR {            // 8 doubles already in registers (static indices only)
  const double* r;
  __device__ __forceinline__ double get(int k) const { return r[k]; }
  __device__ __forceinline__ int pack() const { return (int)__double_as_longlong(r[7]); }
};
struct RecG {            // global scratch, 8 doubles per particle (split kernels)
  const double* r;
  __device__ __forceinline__ double get(int k) const { return __ldg(r + k); }
  __device__ __forceinline__ int pack() const { return (int)__double_as_longlong(__ldg(r + 7)); }
};

// weights at t+3dt/2 expressed on the t+dt/2 stencil: box part hb[i] (slot SFM+i),
// and the single slots one beyond the box on either side
template <int SHAPE>
__device__ __forceinline__ void axis_weights(double f1, double f3, int dc, double* g, double* hb, double& h_lo, double& h_hi) {
  constexpr int NW = Sh<SHAPE>::NW;
  double w3[NW];
  shape_weights<SHAPE>(f1, g);
  shape_weights<SHAPE>(f3, w3);
  h_lo = dc < 0 ? w3[0] : 0.0;
  h_hi = dc > 0 ? w3[NW - 1] : 0.0;
#pragma unroll
  for (int i = 0; i < NW; ++i) {
    const double same = w3[i];
    const double up = i >= 1 ? w3[i >= 1 ? i - 1 : 0] : 0.0;        // dc = +1: new weight index i-1
    const double dn = i < NW - 1 ? w3[i < NW - 1 ? i + 1 : 0] : 0.0;   // dc = -1: new weight index i+1
    const double nw = dc == 0 ? same : (dc > 0 ? up : dn);
    hb[i] = nw - g[i];
  }
}

}  // namespace

namespace {

// interpolation of one staged field component (include/*/e_part.inc, b_part.inc):
// nested sums over x (innermost), y, z with the weights of the component's stagger
template <int ND, int SHAPE>
__device__ __forceinline__ double gather_tile(const double* __restrict__ f, const double* wx, const double* wy,
                                              const double* wz, int bx, int by, int bz) {
  using T = TL<ND, SHAPE>;
  constexpr int NW = T::NW;
  const double* b = f + (ND >= 3 ? bz : 0) * T::FZS + (ND >= 2 ? by : 0) * T::FX + bx;
  double accz = 0.0;
#pragma unroll
  for (int k = 0; k < (ND >= 3 ? NW : 1); ++k) {
    double accy = 0.0;
#pragma unroll
    for (int j = 0; j < (ND >= 2 ? NW : 1); ++j) {
      double accx = 0.0;
#pragma unroll
      for (int i = 0; i < NW; ++i) {
        const double t = wx[i] * b[k * T::FZS + j * T::FX + i];
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

// 3D box accumulation of component A (cumsum axis A, product axes P = A+1, Q = A+2)
template <int A, int NW>
__device__ __forceinline__ void box3(double* acc, const double (*g)[NW], const double (*hb)[NW], const double* hlo,
                                     double fj) {
  constexpr int P = (A + 1) % 3, Q = (A + 2) % 3;
  const double third = 1.0 / 3.0;
  double ca[NW - 1], qf1[NW], qf2[NW];
  double c = hlo[A];
#pragma unroll
  for (int i = 0; i < NW - 1; ++i) { c += hb[A][i]; ca[i] = fj * c; }
#pragma unroll
  for (int v = 0; v < NW; ++v) {
    qf1[v] = g[Q][v] + 0.5 * hb[Q][v];
    qf2[v] = third * hb[Q][v] + 0.5 * g[Q][v];
  }
#pragma unroll
  for (int v = 0; v < NW; ++v)
#pragma unroll
    for (int u = 0; u < NW; ++u) {
      const double w = g[P][u] * qf1[v] + hb[P][u] * qf2[v];
#pragma unroll
      for (int i = 0; i < NW - 1; ++i) acc[(v * NW + u) * (NW - 1) + i] += ca[i] * w;
    }
}

// The same interpolation with the x row fetched as aligned 16-byte pairs (LDS.128): the row starts at the even
// column e = bx & ~1 and the NW weights are shifted by the parity of bx inside a vector of 2 * NP weights
// (zero in the unused slot), so a row costs NP loads instead of NW.  Same value as gather_tile: the extra term is
// an exact zero at either end of the sum.
// MEASURED SLOWER (512^3: push 140 -> 166 ms, profiles/r02_variants_512.txt): the shared-memory data stage delivers
// 128 B per wavefront whatever the access width -- an LDS.128 of a warp costs 4 wavefronts, an LDS.64 two, broadcast
// or not (ncu: 4.19 wavefronts per shared load) -- so pairing saves no wavefronts and adds 54 DFMA.  Kept as an option.
#ifndef EB_PAIR
#define EB_PAIR 0
#endif
template <int SHAPE>
struct PairW {
  static constexpr int NW = Sh<SHAPE>::NW, NP = (NW + 2) / 2;
  double w[2 * NP];
  int e;
  __device__ __forceinline__ void set(const double* wx, int bx) {
    const bool odd = bx & 1;
    e = bx & ~1;
#pragma unroll
    for (int i = 0; i < 2 * NP; ++i) {
      const double lo = i < NW ? wx[i < NW ? i : 0] : 0.0;            // even start: slot i holds weight i
      const double hi = i >= 1 && i - 1 < NW ? wx[i >= 1 && i - 1 < NW ? i - 1 : 0] : 0.0;   // odd start: weight i - 1
      w[i] = odd ? hi : lo;
    }
  }
};
template <int ND, int SHAPE>
__device__ __forceinline__ double gather_tile_pair(const double* __restrict__ f, const PairW<SHAPE>& wx, const double* wy,
                                                   const double* wz, int by, int bz) {
  using T = TL<ND, SHAPE>;
  constexpr int NW = T::NW, NP = PairW<SHAPE>::NP;
  const double2* b = reinterpret_cast<const double2*>(f + (ND >= 3 ? bz : 0) * T::FZS + (ND >= 2 ? by : 0) * T::FX + wx.e);
  double accz = 0.0;
#pragma unroll
  for (int k = 0; k < (ND >= 3 ? NW : 1); ++k) {
    double accy = 0.0;
#pragma unroll
    for (int j = 0; j < (ND >= 2 ? NW : 1); ++j) {
      double accx = 0.0;
#pragma unroll
      for (int i = 0; i < NP; ++i) {
        const double2 v = b[(k * T::FZS + j * T::FX) / 2 + i];
        const double t = wx.w[2 * i] * v.x;
        accx = i == 0 ? t : accx + t;
        accx = accx + wx.w[2 * i + 1] * v.y;
      }
      if (ND >= 2) { const double t = wy[j] * accx; accy = j == 0 ? t : accy + t; }
      else accy = accx;
    }
    if (ND >= 3) { const double t = wz[k] * accy; accz = k == 0 ? t : accz + t; }
    else accz = accy;
  }
  return accz;
}

template <int N>
__device__ __forceinline__ double pick(const double* a, int t) {
  double r = a[0];
#pragma unroll
  for (int i = 1; i < N; ++i) r = t == i ? a[i] : r;
  return r;
}

}  // namespace

namespace {

// Overflow contributions go to the shared J tile with atomics.  FX (deterministic mode): as 64-bit
// fixed-point integers into a second tile `fx.delta` doubles further on -- integer sums do not depend
// on the order of the adds; the tile is folded into J once all items are done.
struct FxP { long long delta; double inv_unit; };
template <bool FX>
__device__ __forceinline__ void ovf_add(double* addr, double v, const FxP& fx) {
  if (FX) atomicAdd(reinterpret_cast<unsigned long long*>(addr + fx.delta), (unsigned long long)__double2ll_rn(v * fx.inv_unit));
  else atomicAdd(addr, v);
}

// One work item of a 3D particle that changed cell along axis m: everything the particle
// deposits outside the register boxes because of that axis.  The window cells not covered
// by the boxes are split, per current component A (P = A+1, Q = A+2 cyclic), by the first
// such axis in the order (A, P, Q):
//   m == A: dcell = -1: cells one below the box along A, any (u, v) of the P x Q window;
//           dcell = +1: the last box slot along A (left out of the register boxes), (u, v)
//           inside the box; one above the box the running sum has returned to zero;
//   m == P: u beyond the box, a inside it, any v of the Q window;
//   m == Q: v beyond the box, a and u inside it.
// With the axes rotated to (a0, a1, a2) = (m, m+1, m+2) the item is role "A" for component
// a0, role "P" for component a2 and role "Q" for component a1, so the weights of the three
// axes are evaluated once.  The values go to the shared J tile with atomics.
template <int SHAPE, class R, bool FX = false>
__device__ __forceinline__ void mover_item3(const PushP& p, const R& rec, double* __restrict__ sJ, int m,
                                            int jn = TL<3, SHAPE>::JN, int sub_off = 0, const FxP fx = FxP{0, 0.0}) {
  using T = TL<3, SHAPE>;
  constexpr int NW = T::NW, NS = NW + 2;
  const int a0 = m, a1 = m == 2 ? 0 : m + 1, a2 = a1 == 2 ? 0 : a1 + 1;
  const int pk = rec.pack();
  const int lc = pk >> 8;
  const int l0 = lc % T::TX, l1 = (lc / T::TX) % T::TY, l2 = lc / (T::TX * T::TY);
  const int dc0 = ((pk >> (2 * a0)) & 3) - 1, dc1 = ((pk >> (2 * a1)) & 3) - 1, dc2 = ((pk >> (2 * a2)) & 3) - 1;
  double g0[NW], h0[NW], g1[NW], h1[NW], g2[NW], h2[NW], lo0, hi0, lo1, hi1, lo2, hi2;
  axis_weights<SHAPE>(rec.get(a0), rec.get(3 + a0), dc0, g0, h0, lo0, hi0);
  axis_weights<SHAPE>(rec.get(a1), rec.get(3 + a1), dc1, g1, h1, lo1, hi1);
  axis_weights<SHAPE>(rec.get(a2), rec.get(3 + a2), dc2, g2, h2, lo2, hi2);
  const double wq = rec.get(6) * p.q;
  const double third = 1.0 / 3.0;
  const int s0 = a0 == 0 ? 1 : (a0 == 1 ? T::JX : T::JX * T::JY);
  const int s1 = a1 == 0 ? 1 : (a1 == 1 ? T::JX : T::JX * T::JY);
  const int s2 = a2 == 0 ? 1 : (a2 == 1 ? T::JX : T::JX * T::JY);
  double* cell = sJ + l0 + l1 * T::JX + (l2 - sub_off) * T::JX * T::JY;   // window index t = 0 <-> region index l
  const double ov0 = dc0 < 0 ? lo0 : hi0;          // the weight of axis m in its slot beyond the box
  const int t0 = dc0 < 0 ? 0 : NW + 1;             // ... and that slot's window index
  {
    // component a0, role A: dcell = -1: slot t = 0, running sum = h_lo, any (u, v) of the window;
    // dcell = +1: the last box slot t = NW, running sum = -h_hi, (u, v) inside the box
    const double fj = -(p.fc[a0] * wq);
    const double cov = dc0 < 0 ? fj * lo0 : -(fj * hi0);
    const int ta = dc0 < 0 ? 0 : NW;
    const int lo = dc0 < 0 ? 0 : 1, hi = dc0 < 0 ? NS : NS - 1;
    double g5[NS], h5[NS], q1[NS], q2[NS];
    g5[0] = 0.0; g5[NS - 1] = 0.0; h5[0] = lo1; h5[NS - 1] = hi1;
    q1[0] = 0.5 * lo2; q2[0] = third * lo2; q1[NS - 1] = 0.5 * hi2; q2[NS - 1] = third * hi2;
#pragma unroll
    for (int u = 0; u < NW; ++u) { g5[u + 1] = g1[u]; h5[u + 1] = h1[u]; }
#pragma unroll
    for (int v = 0; v < NW; ++v) { q1[v + 1] = g2[v] + 0.5 * h2[v]; q2[v + 1] = third * h2[v] + 0.5 * g2[v]; }
    double* base = cell + a0 * jn + ta * s0;
#pragma unroll
    for (int v = 0; v < NS; ++v)
#pragma unroll
      for (int u = 0; u < NS; ++u) {
        // window membership first: the slots beyond the box exist only where P / Q moved too
        const bool in = u >= lo && u < hi && v >= lo && v < hi && (u != 0 || dc1 < 0) && (u != NS - 1 || dc1 > 0) &&
                        (v != 0 || dc2 < 0) && (v != NS - 1 || dc2 > 0);
        if (in) ovf_add<FX>(base + u * s1 + v * s2, cov * (g5[u] * q1[v] + h5[u] * q2[v]), fx);
      }
  }
  {
    // component a2 (P = a0 = m, Q = a1), role P: u in the slot beyond the box (g = 0 there), a inside
    // the box, any v of the Q window
    const double fj = -(p.fc[a2] * wq);
    double ca[NW];
    double c = lo2;
#pragma unroll
    for (int i = 0; i < NW; ++i) { c += h2[i]; ca[i] = fj * c * ov0; }
    double* base = cell + a2 * jn + t0 * s0 + s2;
#pragma unroll
    double q2[NS];
    q2[0] = third * lo1; q2[NS - 1] = third * hi1;
#pragma unroll
    for (int v = 0; v < NW; ++v) q2[v + 1] = third * h1[v] + 0.5 * g1[v];
#pragma unroll
    for (int v = 0; v < NS; ++v) {
      if ((v != 0 || dc1 < 0) && (v != NS - 1 || dc1 > 0)) {
#pragma unroll
        for (int i = 0; i < NW; ++i) ovf_add<FX>(base + i * s2 + v * s1, ca[i] * q2[v], fx);
      }
    }
  }
  {
    // component a1 (P = a2, Q = a0 = m), role Q: v in the slot beyond the box, a and u inside it
    const double fj = -(p.fc[a1] * wq);
    double ca[NW];
    double c = lo1;
#pragma unroll
    for (int i = 0; i < NW; ++i) { c += h1[i]; ca[i] = fj * c; }
    double* base = cell + a1 * jn + t0 * s0 + s1 + s2;
#pragma unroll
    for (int u = 0; u < NW; ++u) {
      const double w = g2[u] * (0.5 * ov0) + h2[u] * (third * ov0);
#pragma unroll
      for (int i = 0; i < NW; ++i) ovf_add<FX>(base + i * s1 + u * s2, ca[i] * w, fx);
    }
  }
}

}  // namespace

namespace {

struct PArr { double* x[3]; double* m[3]; double* w; };
// when on: the push also applies the particle BCs of the (undecomposed) axes to the values it
// stores and produces the next sort key + key histogram, saving the sort's own pass
// mig_axis >= 0 (one decomposed, periodic axis): the push also lists the slots that leave the sub-domain along
// that axis (low side / high side), so that particle_bcs needs no pass over the particles to find them
struct KeySink {
  int on; BcP3 bc; uint32_t* key; uint32_t* count; uint32_t nbucket;
  int mig_axis; double mig_min, mig_max; uint32_t* mig_list[2]; unsigned long long* mig_cnt; uint32_t mig_cap;
};

// deferred permute: slot s of the pre-sort order, or entry s - n_main of the side buffer
template <int ND>
__device__ __forceinline__ void load_particle_g(const PArr& a, const PArr& side, uint32_t s, uint32_t n_main, double* in) {
  const bool far = s >= n_main;
  const long long ip = far ? (long long)(s - n_main) : (long long)s;
#pragma unroll
  for (int d = 0; d < ND; ++d) in[d] = (far ? side.x[d] : a.x[d])[ip];
#pragma unroll
  for (int d = 0; d < 3; ++d) in[3 + d] = (far ? side.m[d] : a.m[d])[ip];
  in[6] = (far ? side.w : a.w)[ip];
}
template <int ND>
__device__ __forceinline__ void load_particle(const PArr& a, long long ip, double* in) {
#pragma unroll
  for (int d = 0; d < ND; ++d) in[d] = a.x[d][ip];
#pragma unroll
  for (int d = 0; d < 3; ++d) in[3 + d] = a.m[d][ip];
  in[6] = a.w[ip];
}

// phase 1 for one particle: particles.F90:219-452 (gather stage, Boris, drift, store,
// extra half drift); leaves the deposit inputs in the record arrays
template <int ND, int SHAPE, bool GATHER, bool PROBE = false, class C0P = const int*>
__device__ __forceinline__ bool push_one(const PushP& p, const double* in /* pos[3], mom[3], weight */, const PArr& a /* output */,
                                         long long ip, C0P c0,
                                         const double* __restrict__ sF, unsigned long long* counters,
                                         double* rec /* NREC doubles */, int& rec_pack, const GridP* g = nullptr,
                                         const KeySink* ks = nullptr, const ProbeSet* pr = nullptr) {
  using T = TL<ND, SHAPE>;
  constexpr int NW = T::NW;
  double pos[3] = {0, 0, 0}, px[3] = {0, 0, 0}, ux, uy, uz;
#pragma unroll
  for (int d = 0; d < ND; ++d) pos[d] = in[d];
  const double part_weight = in[6];
  half_drift<ND>(p, pos, in[3], in[4], in[5], px, ux, uy, uz);
  int cell1[3] = {1, 1, 1}, l1[3] = {0, 0, 0}, l2[3] = {0, 0, 0};
  double f1[3] = {0, 0, 0};
  double wg[3][NW], wh[3][NW];
  bool bad = false;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    const double cr = cell_coord<SHAPE>(px[d], p.idx[d]);
    cell1[d] = cell_round(cr, f1[d]);
    shape_weights<SHAPE>(f1[d], wg[d]);
    double f2;
    const int cell2 = cell_floor(cr, f2);
    shape_weights<SHAPE>(f2, wh[d]);
    l1[d] = cell1[d] - c0[d];
    l2[d] = cell2 - c0[d];
    const int td = tile_dim(ND, d);
    if ((unsigned)l1[d] >= (unsigned)td) { bad = true; l1[d] = l1[d] < 0 ? 0 : td - 1; l2[d] = l1[d]; }
  }
  if (bad) atomicAdd(counters + 8, 1ULL);   // key and push disagree on the cell: must never happen
  // tile-local cell of the record; from here on the cell at t + dt/2 is carried by `lc` alone (the cell shift at the
  // end is taken relative to the tile origin too), which frees the three cell1 registers for the gather
  const int lc = (l1[2] * T::TY + l1[1]) * T::TX + l1[0];
  // region index of weight slot i: g-type l1 + 1 + i, h-type l2 + 1 + i
  const int gx = l1[0] + 1, gy = l1[1] + 1, gz = l1[2] + 1, hx = l2[0] + 1, hy = l2[1] + 1, hz = l2[2] + 1;
  double ex_part, ey_part, ez_part, bx_part, by_part, bz_part;
  if (EB_PAIR && SHAPE != SHAPE_TOPHAT && (T::FX % 2 == 0) && (T::FZS % 2 == 0)) {
    PairW<SHAPE> pg, ph;
    pg.set(wg[0], gx);
    ph.set(wh[0], hx);
    ex_part = gather_tile_pair<ND, SHAPE>(sF + 0 * T::FCS, ph, wg[1], wg[2], gy, gz);
    ey_part = gather_tile_pair<ND, SHAPE>(sF + 1 * T::FCS, pg, wh[1], wg[2], hy, gz);
    ez_part = gather_tile_pair<ND, SHAPE>(sF + 2 * T::FCS, pg, wg[1], wh[2], gy, hz);
    bx_part = gather_tile_pair<ND, SHAPE>(sF + 3 * T::FCS, pg, wh[1], wh[2], hy, hz);
    by_part = gather_tile_pair<ND, SHAPE>(sF + 4 * T::FCS, ph, wg[1], wh[2], gy, hz);
    bz_part = gather_tile_pair<ND, SHAPE>(sF + 5 * T::FCS, ph, wh[1], wg[2], hy, gz);
  } else {
    ex_part = gather_tile<ND, SHAPE>(sF + 0 * T::FCS, wh[0], wg[1], wg[2], hx, gy, gz);
    ey_part = gather_tile<ND, SHAPE>(sF + 1 * T::FCS, wg[0], wh[1], wg[2], gx, hy, gz);
    ez_part = gather_tile<ND, SHAPE>(sF + 2 * T::FCS, wg[0], wg[1], wh[2], gx, gy, hz);
    bx_part = gather_tile<ND, SHAPE>(sF + 3 * T::FCS, wg[0], wh[1], wh[2], gx, hy, hz);
    by_part = gather_tile<ND, SHAPE>(sF + 4 * T::FCS, wh[0], wg[1], wh[2], hx, gy, hz);
    bz_part = gather_tile<ND, SHAPE>(sF + 5 * T::FCS, wh[0], wh[1], wg[2], hx, hy, gz);
  }

  boris(ux, uy, uz, ex_part, ey_part, ez_part, bx_part, by_part, bz_part, p.cmratio, p.ccmratio);

  double delta[3] = {0, 0, 0}, part_vy = 0.0, part_vz = 0.0;
#ifdef EB_EXACT_DIV
  const double gamma_rel = sqrt(ux * ux + uy * uy + uz * uz + 1.0);
  const double igamma3 = 1.0 / gamma_rel;
#else
  const double igamma3 = rsqrt(ux * ux + uy * uy + uz * uz + 1.0);
#endif
  if (ND == 3) {
    const double root = p.dtco2 * igamma3;
    delta[0] = ux * root; delta[1] = uy * root; delta[2] = uz * root;
  } else if (ND == 2) {
    const double igamma = igamma3;
    const double root = p.dtco2 * igamma;
    delta[0] = ux * root; delta[1] = uy * root;
    part_vz = uz * C0 * igamma;
  } else {          // epoch1d/src/particles.F90:330-340
    const double root = C0 * igamma3;
    delta[0] = ux * root * p.dto2;
    part_vy = uy * root;
    part_vz = uz * root;
  }
  double posn[3] = {0, 0, 0}, momn[3] = {p.part_mc * ux, p.part_mc * uy, p.part_mc * uz};
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    px[d] = px[d] + delta[d];
    posn[d] = px[d] + p.gml[d];
  }
  if (PROBE) {
    // particle probes (particles.F90:567-606): energy window, then the plane crossed along its normal between the
    // position the push started from and the one it stores (before the particle boundaries are applied)
    const double part_u2 = ux * ux + uy * uy + uz * uz;
    const double gamma_rel = sqrt(part_u2 + 1.0);
    const double probe_energy = part_u2 / (gamma_rel + 1.0) * (C0 * p.part_mc);
    for (int k = 0; k < pr->n; ++k) {
      if (!pr->on[k] || !(probe_energy > pr->ek_min[k]) || !(probe_energy < pr->ek_max[k])) continue;
      double d_init = 0.0, d_final = 0.0;
#pragma unroll
      for (int d = 0; d < ND; ++d) {
        d_init += pr->normal[k][d] * (pr->point[k][d] - in[d]);
        d_final += pr->normal[k][d] * (pr->point[k][d] - posn[d]);
      }
      if (d_final < 0.0 && d_init >= 0.0) {
        const unsigned long long slot = atomicAdd(pr->count + k, 1ULL);
        if ((long long)slot < pr->cap) {
          double* q = pr->buf + ((size_t)k * pr->cap + slot) * (ND + 4);
#pragma unroll
          for (int d = 0; d < ND; ++d) q[d] = posn[d];
          q[ND] = momn[0]; q[ND + 1] = momn[1]; q[ND + 2] = momn[2]; q[ND + 3] = part_weight;
        }
      }
    }
  }
  if (ks != nullptr && ks->on) {
    uint32_t k = ks->nbucket - 1;
#ifdef EB_FASTKEY
    int ch = 0;
    {
      int ck[3] = {1, 1, 1};
#pragma unroll
      for (int d = 0; d < ND; ++d) {
        double f3;
        ck[d] = cell_round(cell_coord<SHAPE>(px[d] + delta[d], p.idx[d]), f3);
        const BcP& b = ks->bc.b[d];
        if (posn[d] < b.min_local) { posn[d] += b.shift; ck[d] += g->n[d]; }
        if (posn[d] >= b.max_local) { posn[d] -= b.shift; ck[d] -= g->n[d]; }
      }
      k = cell_key<ND>(*g, ck);
    }
#else
    const int ch = bc_and_key<ND, SHAPE>(*g, p, ks->bc, posn, momn, k);
#endif
    if (ch < 0) k = ks->nbucket - 1;
    if (ks->mig_axis >= 0 && ch >= 0) {
      const double xm = ks->mig_axis == 0 ? posn[0] : (ks->mig_axis == 1 ? posn[1] : posn[2]);
      const int side = xm < ks->mig_min ? 0 : (xm >= ks->mig_max ? 1 : -1);
      if (side >= 0) {
        const unsigned long long slot = atomicAdd(ks->mig_cnt + side, 1ULL);
        if (slot < (unsigned long long)ks->mig_cap) ks->mig_list[side][slot] = (uint32_t)ip;
      }
    }
    ks->key[ip] = ch < 0 ? KEY_INVALID : k;
    const unsigned peers = __match_any_sync(__activemask(), k);
    if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(ks->count + k, (uint32_t)__popc(peers));
  } else if (GATHER) {
    ks->key[ip] = 0u;        // fresh slots: no stale KEY_INVALID mark may survive
  }
  if (GATHER) a.w[ip] = part_weight;
#pragma unroll
  for (int d = 0; d < ND; ++d) a.x[d][ip] = posn[d];
  a.m[0][ip] = momn[0]; a.m[1][ip] = momn[1]; a.m[2][ip] = momn[2];
  if (p.zero_current) return false;

  int pack = 0;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    px[d] = px[d] + delta[d];
    double f3;
    const int cell3 = cell_round(cell_coord<SHAPE>(px[d], p.idx[d]), f3);
    const int ld = d == 0 ? lc % T::TX : (d == 1 ? (lc / T::TX) % T::TY : lc / (T::TX * T::TY));
    int dc = (cell3 - c0[d]) - ld;
    dc = dc < -1 ? -1 : (dc > 1 ? 1 : dc);          // |dcell| <= 1 by the CFL condition (particles.F90:506-507)
    pack |= (dc + 1) << (2 * d);
    rec[d] = f1[d];
    rec[ND + d] = f3;
  }
  rec[2 * ND] = part_weight;
  if (ND == 2) rec[2 * ND + 1] = part_vz;
  if (ND == 1) { rec[3] = part_vy; rec[4] = part_vz; }
  rec_pack = pack | (lc << 8);
  return true;
}

}  // namespace

namespace {

// phase 2b: the warp adds the contributions of one cell-changing particle that fall
// outside its home box (at most one layer per axis), lane per (component, cell)
template <int ND, int SHAPE, class R, bool FX = false>
__device__ __forceinline__ void mover_overflow(const PushP& p, const R& rec, double* __restrict__ sJ, int lane,
                                               int jn = TL<ND, SHAPE>::JN, int sub_off = 0, const FxP fx = FxP{0, 0.0}) {
  using T = TL<ND, SHAPE>;
  constexpr int NW = T::NW, NS = NW + 2;
  const int pk = rec.pack();
  const int lc = pk >> 8;
  int l[3] = {lc % T::TX, (lc / T::TX) % T::TY, lc / (T::TX * T::TY)};
  l[ND - 1] -= sub_off;
  const double w = rec.get(2 * ND);
  // full-window arrays per axis: index t = 0 (slot below the box), 1..NW (box), NW+1 (slot above)
  double g5[ND][NS], h5[ND][NS], c5[ND][NS];
  int dc[ND], wd[ND], tlo[ND], tov[ND], ov[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    dc[d] = ((pk >> (2 * d)) & 3) - 1;
    double g[NW], hb[NW], hlo, hhi;
    axis_weights<SHAPE>(rec.get(d), rec.get(ND + d), dc[d], g, hb, hlo, hhi);
    g5[d][0] = 0.0; g5[d][NS - 1] = 0.0; h5[d][0] = hlo; h5[d][NS - 1] = hhi;
#pragma unroll
    for (int i = 0; i < NW; ++i) { g5[d][i + 1] = g[i]; h5[d][i + 1] = hb[i]; }
    double c = 0.0;
#pragma unroll
    for (int t = 0; t < NS; ++t) { c += h5[d][t]; c5[d][t] = c; }
    ov[d] = dc[d] != 0;
    wd[d] = NW + ov[d];
    tlo[d] = dc[d] < 0 ? 0 : 1;
    tov[d] = dc[d] < 0 ? 0 : NW + 1;
  }
  const double third = 1.0 / 3.0;
  int nov;
  if (ND == 3) nov = ov[0] * wd[1] * wd[2] + NW * ov[1] * wd[2] + NW * NW * ov[2];
  else if (ND == 2) nov = ov[0] * wd[1] + NW * ov[1];
  else nov = ov[0];
  for (int it = lane; it < 3 * nov; it += 32) {
    const int comp = it / nov;
    int q = it - comp * nov;
    int t[3] = {1, 1, 1};
    if (ND == 3) {
      const int n0 = ov[0] * wd[1] * wd[2], n1 = NW * ov[1] * wd[2];
      if (q < n0) { t[0] = tov[0]; t[1] = tlo[1] + q % wd[1]; t[2] = tlo[2] + q / wd[1]; }
      else if (q < n0 + n1) { q -= n0; t[0] = 1 + q % NW; t[1] = tov[1]; t[2] = tlo[2] + q / NW; }
      else { q -= n0 + n1; t[0] = 1 + q % NW; t[1] = 1 + (q / NW) % NW; t[2] = tov[2]; }
    } else if (ND == 2) {
      const int n0 = ov[0] * wd[1];
      if (q < n0) { t[0] = tov[0]; t[1] = tlo[1] + q; }
      else { q -= n0; t[0] = 1 + q; t[1] = tov[1]; }
    } else {
      t[0] = tov[0];
    }
    double gv[ND], hv[ND], cv[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) {
      gv[d] = pick<NS>(g5[d], t[d]); hv[d] = pick<NS>(h5[d], t[d]); cv[d] = pick<NS>(c5[d], t[d]);
    }
    double val;
    if (ND == 3) {
      // component A: -fj * C_A * (g_P * (g_Q + h_Q/2) + h_P * (h_Q/3 + g_Q/2)),  P = A+1, Q = A+2
      const int A = comp, P = (comp + 1) % 3, Q = (comp + 2) % 3;
      const double gA = comp == 0 ? cv[0] : (comp == 1 ? cv[1] : cv[2]);
      const double gP = P == 0 ? gv[0] : (P == 1 ? gv[1] : gv[2]);
      const double hP = P == 0 ? hv[0] : (P == 1 ? hv[1] : hv[2]);
      const double gQ = Q == 0 ? gv[0] : (Q == 1 ? gv[1] : gv[2]);
      const double hQ = Q == 0 ? hv[0] : (Q == 1 ? hv[1] : hv[2]);
      const double fc = A == 0 ? p.fc[0] : (A == 1 ? p.fc[1] : p.fc[2]);
      val = -(fc * w * p.q) * gA * (gP * (gQ + 0.5 * hQ) + hP * (third * hQ + 0.5 * gQ));
    } else if (ND == 2) {
      const double xf1 = gv[0] + 0.5 * hv[0];
      const double yf1 = gv[ND - 1] + 0.5 * hv[ND - 1], yf2 = third * hv[ND - 1] + 0.5 * gv[ND - 1];
      if (comp == 0) val = -(p.fc[0] * w * p.q) * cv[0] * yf1;
      else if (comp == 1) val = -(p.fc[1] * w * p.q) * cv[ND - 1] * xf1;
      else val = (p.fc[2] * w * p.q * rec.get(2 * ND + 1)) * (gv[0] * yf1 + hv[0] * yf2);
    } else {
      // epoch1d/src/particles.F90:437-457: jx is the running sum, jy / jz carry the velocity
      const double xf1 = gv[0] + 0.5 * hv[0];
      if (comp == 0) val = -(p.fc[0] * w * p.q) * cv[0];
      else val = (p.fc[1] * w * p.q * rec.get(comp == 1 ? 3 : 4)) * xf1;
    }
    const int idx = ((ND >= 3 ? (l[2] + t[2]) : 0) * T::JY + (ND >= 2 ? (l[1] + t[1]) : 0)) * T::JX + l[0] + t[0];
    ovf_add<FX>(sJ + comp * jn + idx, val, fx);
  }
}

}  // namespace

namespace {

// Box contributions of one particle.  ALL = false: component `comp` into acc[NACC]
// (fused kernel, thread per (cell, component)); ALL = true: the three components into
// acc[3 * NACC] (split deposit kernel, thread per cell; the weights are computed once).
template <int ND, int SHAPE, bool ALL, class R>
__device__ __forceinline__ void deposit_box(const PushP& p, const R& rec, int comp, double* acc) {
  using T = TL<ND, SHAPE>;
  constexpr int NW = T::NW, NA = T::NACC;
  const int pk = rec.pack();
  double gw[ND][NW], hb[ND][NW], hlo[ND], hhi;
#pragma unroll
  for (int d = 0; d < ND; ++d)
    axis_weights<SHAPE>(rec.get(d), rec.get(ND + d), ((pk >> (2 * d)) & 3) - 1, gw[d], hb[d], hlo[d], hhi);
  const double wq = rec.get(2 * ND) * p.q;
  if (ND == 3) {
    if (ALL || comp == 0) box3<0, NW>(acc, gw, hb, hlo, -(p.fc[0] * wq));
    if (ALL || comp == 1) box3<1, NW>(acc + (ALL ? NA : 0), gw, hb, hlo, -(p.fc[1] * wq));
    if (ALL || comp == 2) box3<2, NW>(acc + (ALL ? 2 * NA : 0), gw, hb, hlo, -(p.fc[2] * wq));
  } else if (ND == 1) {
    // epoch1d/src/particles.F90:437-457
    if (ALL || comp == 0) {
      double c = hlo[0];
#pragma unroll
      for (int i = 0; i < NW; ++i) { c += hb[0][i]; acc[i] += -(p.fc[0] * wq) * c; }
    }
#pragma unroll
    for (int k = 1; k < 3; ++k) {
      if (!(ALL || comp == k)) continue;
      double* ak = acc + (ALL ? k * NA : 0);
      const double fj = p.fc[1] * wq * rec.get(2 + k);      // fcy * q * part_vy | part_vz
#pragma unroll
      for (int i = 0; i < NW; ++i) ak[i] += fj * (gw[0][i] + 0.5 * hb[0][i]);
    }
  } else {
    const double third = 1.0 / 3.0;
    if (ALL || comp == 0) {          // Jx = -fjx * cumsum_x(hx) * (gy + hy/2)
      double c = hlo[0];
#pragma unroll
      for (int i = 0; i < NW; ++i) {
        c += hb[0][i];
        const double ca = -(p.fc[0] * wq) * c;
#pragma unroll
        for (int j = 0; j < NW; ++j) acc[j * NW + i] += ca * (gw[ND - 1][j] + 0.5 * hb[ND - 1][j]);
      }
    }
    if (ALL || comp == 1) {          // Jy = -fjy * cumsum_y(hy) * (gx + hx/2)
      double* ay = acc + (ALL ? NA : 0);
      double c = hlo[ND - 1];
#pragma unroll
      for (int j = 0; j < NW; ++j) {
        c += hb[ND - 1][j];
        const double ca = -(p.fc[1] * wq) * c;
#pragma unroll
        for (int i = 0; i < NW; ++i) ay[j * NW + i] += ca * (gw[0][i] + 0.5 * hb[0][i]);
      }
    }
    if (ALL || comp == 2) {          // Jz = fjz * (gx*(gy + hy/2) + hx*(hy/3 + gy/2))
      double* az = acc + (ALL ? 2 * NA : 0);
      const double fjz = p.fc[2] * wq * rec.get(2 * ND + 1);
#pragma unroll
      for (int j = 0; j < NW; ++j) {
        const double yf1 = gw[ND - 1][j] + 0.5 * hb[ND - 1][j], yf2 = third * hb[ND - 1][j] + 0.5 * gw[ND - 1][j];
#pragma unroll
        for (int i = 0; i < NW; ++i) az[j * NW + i] += fjz * (gw[0][i] * yf1 + hb[0][i] * yf2);
      }
    }
  }
}

// work items of a 3D particle that changed cell: one per moving axis
__device__ __forceinline__ int mover_item_count(int pack) {
  int n = 0;
#pragma unroll
  for (int m = 0; m < 3; ++m) n += ((pack >> (2 * m)) & 3) != 1 ? 1 : 0;
  return n;
}

}  // namespace

// =====================================================================================
// Split variant: push and deposit as two kernels over the same tiles, the deposit
// records passing through a global scratch (64 B per particle, L2/HBM).  The push
// kernel then carries no accumulators (more resident warps to hide the gather and
// Boris latencies) and the deposit kernel can give one thread ALL THREE components of a
// cell (81 register accumulators in 3D), so the shape weights of a particle are
// evaluated once instead of once per component.  Tiles are processed in slabs whose
// particles fit the scratch.
// =====================================================================================
namespace {

#ifndef EB_PUSH_CHUNK
#define EB_PUSH_CHUNK 16384
#endif
constexpr int PUSH_CHUNK = EB_PUSH_CHUNK;   // particles per push CTA (a multiple of the CTA size)
#ifndef EB_STAGE_HALF
#define EB_STAGE_HALF 1
#endif
#ifndef EB_SMEM_PAD
#define EB_SMEM_PAD 0
#endif
#ifndef EB_NTA
#define EB_NTA 128
#endif
constexpr int NTA = EB_NTA;      // push kernel threads
#ifndef EB_MCAPB
#define EB_MCAPB 2048
#endif
#ifndef EB_ONEC
#define EB_ONEC 0                // 1: one pass per current component for every shape (fewer registers, more CTAs per SM)
#endif
constexpr int MCAPB = EB_MCAPB;  // overflow work items per sub-tile in the deposit kernel (more are deposited inline)
#ifndef EB_MINB_B
#define EB_MINB_B 2
#endif
#ifndef EB_RING
#define EB_RING 3
#endif
constexpr int RING = EB_RING;    // record slots per thread of the deposit kernel's cp.async ring

// GATHER (deferred permute, particles.cu): the particle of sorted index q is read from slot src[q] of the
// arrays `a` (the pre-sort order) and stored to slot q of `o` (the other state of the same allocations)
template <int ND, int SHAPE, bool GATHER, bool PROBE>
__global__ void __launch_bounds__(NTA, SHAPE == SHAPE_BSPLINE3 ? 2 : EB_MINB_A) k_push_split(GridP g, FieldP F, PushP p, PArr a, PArr o, PArr side,
                                                              const uint32_t* __restrict__ src, uint32_t n_main,
                                                              const uint32_t* __restrict__ cell_start, long long tile0,
                                                              double* __restrict__ scratch, unsigned long long* counters,
                                                              KeySink ks, const __grid_constant__ CUtensorMap tmap, int use_tma,
                                                              const ProbeSet* __restrict__ pr) {
  using T = TL<ND, SHAPE>;
  constexpr int SFM = T::SFM;
  extern __shared__ __align__(128) double smem_a[];
  double* smem = smem_a;
  double* sF = smem;
  __shared__ __align__(8) unsigned long long mbar;
  const int tid = threadIdx.x;
  const long long tile = tile0 + blockIdx.x;
  const long long kbase = tile * T::CPT;
  // a tile's particles are shared out in chunks of PUSH_CHUNK among the CTAs (blockIdx.x, 0 .. gridDim.y - 1): one CTA
  // per tile for a uniform plasma, many for the dense tiles of a foil (each stages the tile's fields itself)
  const uint32_t tb = cell_start[kbase], te = cell_start[kbase + T::CPT];
  if ((unsigned long long)tb + (unsigned long long)blockIdx.y * PUSH_CHUNK >= (unsigned long long)te) return;
  const uint32_t pb = tb + blockIdx.y * (uint32_t)PUSH_CHUNK;
  const uint32_t pe = te - pb > (uint32_t)PUSH_CHUNK ? pb + (uint32_t)PUSH_CHUNK : te;
  const uint32_t base = cell_start[tile0 * T::CPT];      // first particle of the slab
  int tc[3];
  tc[0] = (int)(tile % g.ntile[0]);
  tc[1] = (int)((tile / g.ntile[0]) % g.ntile[1]);
  tc[2] = (int)(tile / ((long long)g.ntile[0] * g.ntile[1]));
  int c0[3], org[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    c0[d] = d < ND ? (1 - KM) + tc[d] * tile_dim(ND, d) : 1;
    org[d] = d < ND ? c0[d] + SFM - 1 : 1;
  }
#if EB_C0_SMEM
  // the tile origin lives in shared memory and is re-read per particle: as three loop-invariant registers it pushed
  // the loop over the 128-register budget and ptxas answered with a local-memory spill whose reload (LDL, long
  // scoreboard) sat at the top of every iteration -- 14 % of the kernel's stall samples in the round-1 capture
  __shared__ int s_c0[4];
  if (tid == 0) { s_c0[0] = c0[0]; s_c0[1] = c0[1]; s_c0[2] = c0[2]; }
  // (both staging branches below pass a __syncthreads before the particle loop)
#endif
  const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&mbar);
  if (use_tma) {
    // E,B of the tile + gather halo: ONE bulk tensor copy (TMA) of the box [FZ][6 components][FY][FX] out of the
    // nine-array field allocation; cells outside the allocation arrive as zeros
    if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
      const uint32_t dst = (uint32_t)__cvta_generic_to_shared(sF);
      const int x0 = org[0] + g.ngd[0] - 1, x1 = ND >= 2 ? org[1] + g.ngd[1] - 1 : 0, x2 = ND >= 3 ? org[2] + g.ngd[2] - 1 : 0;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(6 * T::FN * sizeof(double))) : "memory");
      if (ND == 3)
        asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
                     ::"r"(dst), "l"((unsigned long long)&tmap), "r"(x0), "r"(x1), "r"(EB_ZCOMP ? 0 : x2), "r"(EB_ZCOMP ? x2 : 0), "r"(bar) : "memory");
      else
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(dst), "l"((unsigned long long)&tmap), "r"(x0), "r"(x1), "r"(0), "r"(bar) : "memory");
    }
  } else {
    for (int i0 = 0; i0 < 6 * T::FN; i0 += 4 * NTA) {
      double v[4];
  #pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * NTA + tid;
        v[u] = 0.0;
        if (i < 6 * T::FN) {
          const int fx = i % T::FX, fy = (i / T::FX) % T::FY;
          const int comp = EB_ZCOMP ? (i / T::FCS) % 6 : i / T::FN, fz = EB_ZCOMP ? i / T::FZS : (i / T::FCS) % T::FZ;
          const int X = org[0] + fx, Y = org[1] + fy, Z = org[2] + fz;
          const bool in = X >= 1 - g.ngd[0] && X <= g.n[0] + g.ngd[0] && Y >= 1 - g.ngd[1] && Y <= g.n[1] + g.ngd[1] &&
                          Z >= 1 - g.ngd[2] && Z <= g.n[2] + g.ngd[2];
          if (in) v[u] = __ldg(F.f[comp] + g.idx(X, Y, Z));
        }
      }
  #pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * NTA + tid;
        if (i < 6 * T::FN) sF[i] = v[u];
      }
    }
    __syncthreads();
  }
  // warp-uniform loop over the tile's particles; the next particle is in flight while the current
  // one is pushed.  The 64 B records of a warp are transposed through shared memory (16 B units,
  // XOR-swizzled) so that the warp writes its 2 KB block of the scratch with full-line stores.
  const int lane = tid & 31, warp = tid >> 5;
  double2* stage = reinterpret_cast<double2*>(smem + 6 * T::FN) + warp * (EB_STAGE_HALF ? 64 : 128);
  double nxt[7];
  uint32_t q0 = pb + warp * 32;
  uint32_t s2 = 0;           // GATHER: slot of the particle after next
  if (GATHER) {
    if (q0 + lane < pe) load_particle_g<ND>(a, side, __ldg(src + q0 + lane), n_main, nxt);
    if (q0 + lane + NTA < pe) s2 = __ldg(src + q0 + lane + NTA);
  } else if (q0 + lane < pe) load_particle<ND>(a, (long long)(q0 + lane), nxt);
  if (use_tma) {      // the first particles are already in flight: now wait for the tile
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(bar), "r"(0) : "memory");
  }
  for (; q0 < pe; q0 += NTA) {
    const uint32_t q = q0 + lane;
    const bool act = q < pe;
    double in[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) in[k] = nxt[k];
    if (q + NTA < pe) {
      if (GATHER) load_particle_g<ND>(a, side, s2, n_main, nxt);
      else load_particle<ND>(a, (long long)(q + NTA), nxt);
    }
    if (GATHER && q + 2 * NTA < pe) s2 = __ldg(src + q + 2 * NTA);
    double rec[8];
    if (act) {
      int pack;
      rec[6] = 0.0;
#if EB_C0_SMEM
      push_one<ND, SHAPE, GATHER, PROBE, const volatile int*>(p, in, o, (long long)q, s_c0, sF, counters, rec, pack, &g, &ks, pr);
#else
      push_one<ND, SHAPE, GATHER, PROBE>(p, in, o, (long long)q, c0, sF, counters, rec, pack, &g, &ks, pr);
#endif
      rec[7] = __longlong_as_double((long long)pack);
    }
    if (p.zero_current) continue;
    double2* out = reinterpret_cast<double2*>(scratch + (size_t)(q0 - base) * 8);
    const uint32_t nrec = pe - q0 < 32u ? pe - q0 : 32u;
#if EB_STAGE_HALF
    // 16 records (1 KB) per round: the staging area stays at 1 KB per warp, which keeps four CTAs within the
    // 196 KB shared-memory carve-out (60 KB of L1 left for the stack and the particle loads)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (act && (lane >> 4) == h) {
#pragma unroll
        for (int k = 0; k < 4; ++k) stage[(lane & 15) * 4 + (k ^ ((lane >> 1) & 3))] = make_double2(rec[2 * k], rec[2 * k + 1]);
      }
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int u = j * 32 + lane, r = u >> 2;
        if ((uint32_t)(h * 16 + r) < nrec) out[h * 64 + u] = stage[r * 4 + ((u & 3) ^ ((r >> 1) & 3))];
      }
      __syncwarp();
    }
#else
    if (act) {
#pragma unroll
      for (int k = 0; k < 4; ++k) stage[lane * 4 + (k ^ ((lane >> 1) & 3))] = make_double2(rec[2 * k], rec[2 * k + 1]);
    }
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int u = j * 32 + lane, r = u >> 2;
      if ((uint32_t)r < nrec) out[u] = stage[r * 4 + ((u & 3) ^ ((r >> 1) & 3))];
    }
    __syncwarp();
#endif
  }
}

constexpr int SUB = EB_SUB;      // a deposit CTA owns 1/SUB of a tile, split along the slowest tile axis

// Lane balance of the deposit kernel (EB_BALANCE): a thread owns one cell and a warp runs as long as its fullest
// cell, so the cells are dealt to the warps by particle count.  Any permutation of cells over threads keeps the
// register-box flush valid (the cells of a step stay distinct), but the flush is conflict-free only if a warp
// holds exactly two cells of each of the 16 shared-memory bank classes (8-byte words): the cells of a class are
// ranked by count and warp w takes ranks 2w and 2w + 1 of every class.  Available when the sub-tile's cells fall
// evenly into the classes (3D triangle, 2D triangle / bspline3).
// MEASURED WITHOUT EFFECT (512^3: deposit 92.24 -> 92.23 ms, profiles/r02_variants_512.txt): a CTA still runs as long as
// its fullest cell and with two CTAs per SM the issue slots the lighter warps give up are not what limits it.  Option.
#ifndef EB_BALANCE
#define EB_BALANCE 0
#endif
template <int ND, int SHAPE>
constexpr bool uniform_bank_classes() {
  using T = TL<ND, SHAPE>;
  constexpr int NTB = T::CPT / SUB;
  int cnt[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (int c = 0; c < NTB; ++c) {
    const int l0 = c % T::TX, l1 = (c / T::TX) % T::TY, l2 = c / (T::TX * T::TY);
    cnt[(l0 + l1 * T::JX + (ND == 3 ? l2 * T::JX * T::JY : 0)) & 15]++;
  }
  for (int b = 0; b < 16; ++b)
    if (cnt[b] != NTB / 16) return false;
  return NTB % 64 == 0;
}

__device__ __forceinline__ void load_rec(double* r, const double* src) {
  const double2* s2 = reinterpret_cast<const double2*>(src);
#pragma unroll
  for (int k = 0; k < 4; ++k) { const double2 v = __ldg(s2 + k); r[2 * k] = v.x; r[2 * k + 1] = v.y; }
}

// DET (option deterministic_deposit): the overflow items accumulate in fixed point and the tile is stored to
// `tile_out` (one J tile per CTA) instead of being added to HBM with RED; k_reduce_tiles then sums the
// tiles that overlap a cell in a fixed order.  With the canonical particle order of the sort (k_canon) the
// currents are bitwise reproducible from run to run.
template <int ND, int SHAPE, bool DET>
__global__ void __launch_bounds__(TL<ND, SHAPE>::CPT / SUB, EB_MINB_B) k_deposit_split(GridP g, FieldP F, PushP p,
                                                                                const uint32_t* __restrict__ cell_start,
                                                                                long long tile0, const double* __restrict__ scratch,
                                                                                double* __restrict__ tile_out) {
  using T = TL<ND, SHAPE>;
  constexpr int NW = T::NW, SFM = T::SFM, NA = T::NACC, NTB = T::CPT / SUB;
  constexpr int SA = ND - 1;                                   // split axis
  constexpr int TS = tile_dim(ND, SA) / SUB;                   // cells of the sub-tile along it
  constexpr int JXs = ND == 1 ? TS + NW + 1 : T::JX, JYs = ND == 3 ? T::JY : (ND == 2 ? TS + NW + 1 : 1), JZs = ND == 3 ? TS + NW + 1 : 1;
  constexpr int JNs = JXs * JYs * JZs;
  extern __shared__ double smem[];
  double* sJ = smem;                          // [3][JNs] (DET: + [3][JNs] fixed-point overflow sums)
  int* sM = (int*)(sJ + (DET ? 6 : 3) * JNs); // [2 * MCAPB] work items (record, moving axis)
  int* sCnt = sM + 2 * MCAPB;
  double2* ring = reinterpret_cast<double2*>(((uintptr_t)(sCnt + 4) + 15) & ~(uintptr_t)15);   // [RING][4][NTB] 16 B units
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long tile = tile0 + blockIdx.x / SUB;
  const int sub = blockIdx.x % SUB;
  const long long kbase = tile * T::CPT;
  const uint32_t pb = cell_start[kbase], sb0 = cell_start[kbase + sub * NTB], sb1 = cell_start[kbase + (sub + 1) * NTB];
  if (sb0 == sb1) return;
  const uint32_t base = cell_start[tile0 * T::CPT];
  int tc[3];
  tc[0] = (int)(tile % g.ntile[0]);
  tc[1] = (int)((tile / g.ntile[0]) % g.ntile[1]);
  tc[2] = (int)(tile / ((long long)g.ntile[0] * g.ntile[1]));
  int org[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) org[d] = d < ND ? (1 - KM) + tc[d] * tile_dim(ND, d) + SFM - 1 + (d == SA ? sub * TS : 0) : 1;
  for (int i = tid; i < (DET ? 6 : 3) * JNs; i += NTB) sJ[i] = 0.0;      // (+0.0 is the integer 0)
  if (tid == 0) { sCnt[0] = 0; sCnt[1] = 0; sCnt[2] = 0; sCnt[3] = 0; }
  __syncthreads();
  FxP fx{0, 0.0};
  if (DET) {
    // fixed-point unit of the overflow sums: 2^-44 of the largest |fc w q| of the sub-tile (a power of two;
    // the maximum does not depend on the order it is taken in)
    double wmax = 0.0;
    for (uint32_t q = sb0 + tid; q < sb1; q += NTB) wmax = fmax(wmax, fabs(__ldg(scratch + (size_t)(q - base) * 8 + 2 * ND)));
    atomicMax(reinterpret_cast<unsigned long long*>(sCnt + 2), (unsigned long long)__double_as_longlong(wmax));
    __syncthreads();
    const double wm = __longlong_as_double((long long)*reinterpret_cast<unsigned long long*>(sCnt + 2));
    double fcm = fmax(fabs(p.fc[0]), fabs(p.fc[1]));
    fcm = fmax(fcm, fabs(p.fc[2]) * (ND == 2 ? C0 : 1.0));
    if (ND == 1) fcm = fmax(fabs(p.fc[0]), fabs(p.fc[1]) * C0);
    const double S = fmax(wm * fabs(p.q) * fcm, 1e-300);
    int ex;
    frexp(S, &ex);                                  // S < 2^ex
    fx.delta = 3 * JNs;
    fx.inv_unit = ldexp(1.0, 44 - ex);
  }

  // ---- thread per cell: all components of the cell's box in registers; the next record is
  //      fetched while the current one is deposited.  ONEC (3D bspline3: 3 x 100 accumulators):
  //      one pass over the cell's records per component instead ----
  constexpr bool ONEC = 3 * NA > 128 || (EB_ONEC && ND == 3);
  int lc = sub * NTB + tid;
  if (EB_BALANCE && uniform_bank_classes<ND, SHAPE>()) {
    constexpr int PER = NTB / 16;                 // cells per bank class
    int* cls_n = sM;                              // (the work-item list is not in use yet)
    int* cls_list = sM + 16;                      // [16][PER] count << 8 | cell
    int* cell_of = sM + 16 + NTB;                 // [16][PER] cell of (class, rank)
    if (tid < 16) cls_n[tid] = 0;
    __syncthreads();
    const int n0 = (int)min(cell_start[kbase + lc + 1] - cell_start[kbase + lc], 0x7fffffu);
    const int m0 = tid % T::TX, m1 = (tid / T::TX) % T::TY, m2 = tid / (T::TX * T::TY);
    const int cls = (m0 + m1 * JXs + (ND == 3 ? m2 * JXs * JYs : 0)) & 15;
    cls_list[cls * PER + atomicAdd(cls_n + cls, 1)] = n0 << 8 | tid;
    __syncthreads();
    int r = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int e = cls_list[cls * PER + i];
      r += ((e >> 8) > n0 || ((e >> 8) == n0 && (e & 255) < tid)) ? 1 : 0;
    }
    cell_of[cls * PER + r] = tid;
    __syncthreads();
    // fullest cells to warp (0 + rot): neighbouring CTAs rotate so that an SM sub-partition sees a mix
    const int w = (warp + (int)(blockIdx.x & 3)) % (NTB / 32);
    lc = sub * NTB + cell_of[(lane & 15) * PER + (PER / (NTB / 32)) * w + (lane >> 4)];
    __syncthreads();
  }
  const uint32_t my_b = cell_start[kbase + lc], my_e = cell_start[kbase + lc + 1];
  int l[3] = {lc % T::TX, (lc / T::TX) % T::TY, lc / (T::TX * T::TY)};
  l[SA] -= sub * TS;
  const uint32_t ring0 = (uint32_t)__cvta_generic_to_shared(ring) + (uint32_t)tid * 16u;
  auto issue = [&](uint32_t q, int slot) {
    const double* src = scratch + (size_t)(q - base) * 8;
#pragma unroll
    for (int k = 0; k < 4; ++k)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring0 + (uint32_t)((slot * 4 + k) * NTB) * 16u), "l"(src + 2 * k) : "memory");
  };
#pragma unroll
  for (int pass = 0; pass < (ONEC ? 3 : 1); ++pass) {
  double acc[ONEC ? NA : 3 * NA];
#pragma unroll
  for (int i = 0; i < (ONEC ? NA : 3 * NA); ++i) acc[i] = 0.0;
  // the thread's records stream through a private shared-memory ring filled by cp.async
  // (RING - 1 records in flight, none of them held in registers)
  double cur[8];
#pragma unroll
  for (int s = 0; s < RING - 1; ++s) {
    if (my_b + s < my_e) issue(my_b + s, s);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  int slot = 0;
  for (uint32_t q = my_b; q < my_e; ++q) {
    asm volatile("cp.async.wait_group %0;" ::"n"(RING - 2) : "memory");
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const double2 v = ring[(slot * 4 + k) * NTB + tid];
      cur[2 * k] = v.x; cur[2 * k + 1] = v.y;
    }
    // refill the slot that the previous iteration consumed
    if (q + (RING - 1) < my_e) issue(q + (RING - 1), slot == 0 ? RING - 1 : slot - 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    slot = slot == RING - 1 ? 0 : slot + 1;
    deposit_box<ND, SHAPE, !ONEC>(p, RecR{cur}, pass, acc);
    const int pack = (int)__double_as_longlong(cur[7]);
    if (pass == 0 && (pack & 0xff) != (ND == 3 ? 0x15 : (ND == 2 ? 0x5 : 0x1))) {
      const RecG rec{scratch + (size_t)(q - base) * 8};
      const int r = (int)(q - pb);
      if (ND <= 2) {
        const int slot = atomicAdd(sCnt, 1);
        if (slot < MCAPB) sM[2 * slot] = r;
        else for (int ln = 0; ln < 32; ++ln) mover_overflow<ND, SHAPE, RecG, DET>(p, rec, sJ, ln, JNs, sub * TS, fx);   // list full
      } else {
        int slot = atomicAdd(sCnt, mover_item_count(pack));
#pragma unroll
        for (int m = 0; m < 3; ++m) {
          if (((pack >> (2 * m)) & 3) == 1) continue;
          if (slot < MCAPB) { sM[2 * slot] = r; sM[2 * slot + 1] = m; }
          else mover_item3<SHAPE, RecG, DET>(p, rec, sJ, m, JNs, sub * TS, fx);
          ++slot;
        }
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  // ---- register boxes -> shared J tile: every thread applies the same stencil offset to its own
  //      cell, so the addresses of a step are distinct; steps are separated by barriers ----
  // With a sub-tile two cells thick along z and NW = 3, the offsets dz = 0 and dz = 2 touch disjoint
  // planes and share a step.
  constexpr bool ZPAIR = ND == 3 && TS == 2 && NW == 3;
  auto box_step = [&](int dx, int dy, int dz) {
    const int off = (l[0] + 1 + dx) + (ND >= 2 ? (l[1] + 1 + dy) * JXs : 0) + (ND == 3 ? (l[2] + 1 + dz) * JXs * JYs : 0);
    if (ND == 1) {
      sJ[off] += acc[dx];
      sJ[JNs + off] += acc[NA + dx];
      sJ[2 * JNs + off] += acc[2 * NA + dx];
    } else if (ND == 3) {
      // component A: acc[(o[Q] * NW + o[P]) * (NW - 1) + o[A]], P = A + 1, Q = A + 2, o[A] < NW - 1
      if ((!ONEC || pass == 0) && dx < NW - 1) sJ[off] += acc[(dz * NW + dy) * (NW - 1) + dx];
      if ((!ONEC || pass == 1) && dy < NW - 1) sJ[JNs + off] += acc[(ONEC ? 0 : NA) + (dx * NW + dz) * (NW - 1) + dy];
      if ((!ONEC || pass == 2) && dz < NW - 1) sJ[2 * JNs + off] += acc[(ONEC ? 0 : 2 * NA) + (dy * NW + dx) * (NW - 1) + dz];
    } else {
      sJ[off] += acc[dy * NW + dx];
      sJ[JNs + off] += acc[NA + dy * NW + dx];
      sJ[2 * JNs + off] += acc[2 * NA + dy * NW + dx];
    }
  };
#pragma unroll
  for (int dz = 0; dz < (ND == 3 ? (ZPAIR ? 2 : NW) : 1); ++dz)
#pragma unroll
    for (int dy = 0; dy < (ND >= 2 ? NW : 1); ++dy)
#pragma unroll
      for (int dx = 0; dx < NW; ++dx) {
        box_step(dx, dy, dz);
        if (ZPAIR && dz == 0) box_step(dx, dy, 2);
        __syncthreads();
      }
  }   // pass
  // ---- particles that changed cell ----
  const int nm = sCnt[0] < MCAPB ? sCnt[0] : MCAPB;
  if (ND == 3) {
    for (int m = tid; m < nm; m += NTB) {
      const RecG rec{scratch + (size_t)(pb - base + (uint32_t)sM[2 * m]) * 8};
      mover_item3<SHAPE, RecG, DET>(p, rec, sJ, sM[2 * m + 1], JNs, sub * TS, fx);
    }
  } else {
    for (int m = warp; m < nm; m += NTB / 32) {
      const RecG rec{scratch + (size_t)(pb - base + (uint32_t)sM[2 * m]) * 8};
      mover_overflow<ND, SHAPE, RecG, DET>(p, rec, sJ, lane, JNs, sub * TS, fx);
    }
  }
  __syncthreads();
  if (DET) {
    // fold the fixed-point overflow sums into the tile and store it: k_reduce_tiles adds the tiles up
    const double unit = 1.0 / fx.inv_unit;
    double* out = tile_out + (size_t)blockIdx.x * (3 * JNs);
    for (int i = tid; i < 3 * JNs; i += NTB)
      out[i] = sJ[i] + (double)__double_as_longlong(sJ[3 * JNs + i]) * unit;
    return;
  }
  // tile J -> HBM (RED.ADD.F64: tile halos overlap): thread per (component, y, z) row, walking x (1D: thread per value)
  if (ND == 1) {
    for (int i = tid; i < 3 * JXs; i += NTB) {
      const int cj = i / JXs, jx = i - cj * JXs;
      const int X = org[0] + jx;
      const double v = sJ[cj * JNs + jx];
      if (v != 0.0 && X >= 1 - g.ngd[0] && X <= g.n[0] + g.ngd[0]) atomicAdd(F.f[JX + cj] + g.idx(X, 1, 1), v);
    }
    return;
  }
  for (int row = tid; row < 3 * JYs * JZs; row += NTB) {
    const int cj = row / (JYs * JZs), r = row - cj * (JYs * JZs);
    const int jy = r % JYs, jz = r / JYs;
    const int Y = org[1] + jy, Z = org[2] + jz;
    if (Y < 1 - g.ngd[1] || Y > g.n[1] + g.ngd[1] || Z < 1 - g.ngd[2] || Z > g.n[2] + g.ngd[2]) continue;
    const double* src = sJ + cj * JNs + (jz * JYs + jy) * JXs;
    double* dst = F.f[JX + cj] + g.idx(org[0], Y, Z);
    const int x_lo = 1 - g.ngd[0] - org[0], x_hi = g.n[0] + g.ngd[0] - org[0];
#pragma unroll
    for (int jx = 0; jx < JXs; ++jx) {
      const double v = src[jx];
      if (v != 0.0 && jx >= x_lo && jx <= x_hi) atomicAdd(dst + jx, v);
    }
  }
}

}  // namespace

namespace {
// Deterministic mode: J(cell) += the entries of all sub-tile J tiles of this launch that cover the cell, taken
// in ascending (z, y, x) sub-tile order.  Thread per cell of the planes [lo, hi] along the slowest axis.
template <int ND, int SHAPE>
__global__ void k_reduce_tiles(GridP g, FieldP F, const uint32_t* __restrict__ cell_start, long long t0, long long t1,
                               const double* __restrict__ tiles, int lo, int hi) {
  using T = TL<ND, SHAPE>;
  constexpr int NW = T::NW, SFM = T::SFM, NTB = T::CPT / SUB, SA = ND - 1;
  constexpr int TS = tile_dim(ND, SA) / SUB;
  constexpr int JXs = ND == 1 ? TS + NW + 1 : T::JX, JYs = ND == 3 ? T::JY : (ND == 2 ? TS + NW + 1 : 1), JZs = ND == 3 ? TS + NW + 1 : 1;
  constexpr int JNs = JXs * JYs * JZs;
  const int ex = g.n[0] + 2 * g.ngd[0], ey = ND >= 2 ? g.n[1] + 2 * g.ngd[1] : 1;
  const long long planes = (long long)(hi - lo + 1);
  const long long per_plane = ND == 3 ? (long long)ex * ey : (ND == 2 ? ex : 1);
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= planes * per_plane) return;
  int X[3] = {1, 1, 1};
  X[0] = ND == 1 ? lo + (int)i : 1 - g.ngd[0] + (int)(i % ex);
  if (ND == 1) {
  } else if (ND == 3) { X[1] = 1 - g.ngd[1] + (int)((i / ex) % ey); X[2] = lo + (int)(i / per_plane); }
  else X[1] = lo + (int)(i / per_plane);
  // candidate (sub-)tile ranges per axis: origin(u) = b0 + u * stride <= X <= origin(u) + extent - 1
  int u_lo[3] = {0, 0, 0}, u_hi[3] = {0, 0, 0};
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    const int b0 = (1 - KM) + SFM - 1;
    const int stride = d == SA ? TS : tile_dim(ND, d);
    const int extent = stride + NW + 1;
    const int nu = g.ntile[d] * (d == SA ? SUB : 1);
    const int rel = X[d] - b0;
    int hi_u = rel >= 0 ? rel / stride : -1;
    int lo_u = rel - extent + 1 > 0 ? (rel - extent + 1 + stride - 1) / stride : 0;
    if (hi_u > nu - 1) hi_u = nu - 1;
    u_lo[d] = lo_u; u_hi[d] = hi_u;
  }
  double sum[3] = {0.0, 0.0, 0.0};
  for (int uz = u_lo[2]; uz <= u_hi[2]; ++uz)
    for (int uy = u_lo[1]; uy <= u_hi[1]; ++uy)
      for (int ux = u_lo[0]; ux <= u_hi[0]; ++ux) {
        const int u[3] = {ux, uy, uz};
        const int sub = u[SA] % SUB;
        int tcd[3] = {ux, uy, uz};
        tcd[SA] = u[SA] / SUB;
        const long long tile = ((long long)(ND == 3 ? tcd[2] : 0) * g.ntile[1] + (ND >= 2 ? tcd[1] : 0)) * g.ntile[0] + tcd[0];
        if (tile < t0 || tile >= t1) continue;
        const long long kbase = tile * T::CPT;
        if (cell_start[kbase + sub * NTB] == cell_start[kbase + (sub + 1) * NTB]) continue;   // the CTA returned early
        int j[3] = {0, 0, 0};
#pragma unroll
        for (int d = 0; d < ND; ++d) j[d] = X[d] - ((1 - KM) + SFM - 1 + u[d] * (d == SA ? TS : tile_dim(ND, d)));
        const double* tl = tiles + (size_t)((tile - t0) * SUB + sub) * (3 * JNs) + (j[2] * JYs + j[1]) * JXs + j[0];
#pragma unroll
        for (int cj = 0; cj < 3; ++cj) sum[cj] += tl[cj * JNs];
      }
  const size_t gi = g.idx(X[0], X[1], X[2]);
#pragma unroll
  for (int cj = 0; cj < 3; ++cj)
    if (sum[cj] != 0.0) F.f[JX + cj][gi] += sum[cj];
}
}  // namespace

size_t keysink_size() { return sizeof(KeySink); }
void keysink_fill(void* dst, const BcP3& bc, uint32_t* key, uint32_t* count, uint32_t nbucket) {
  KeySink ks{};
  ks.on = 1; ks.bc = bc; ks.key = key; ks.count = count; ks.nbucket = nbucket;
  ks.mig_axis = -1;
  *(KeySink*)dst = ks;
}
void keysink_set_migration(void* dst, int axis, double mn, double mx, uint32_t* list_lo, uint32_t* list_hi,
                           unsigned long long* cnt, uint32_t cap) {
  KeySink& ks = *(KeySink*)dst;
  ks.mig_axis = axis; ks.mig_min = mn; ks.mig_max = mx; ks.mig_list[0] = list_lo; ks.mig_list[1] = list_hi;
  ks.mig_cnt = cnt; ks.mig_cap = cap;
}

// One tensor map over E,B as a fp64 tensor with the dimensions (x, y, component, z) -- the six arrays are
// consecutive in the field allocation -- box = the tile's gather region in the kernel's shared layout.  Needs an even x extent (16-byte strides); else, or when the
// driver lacks the encoder, the kernel stages the tile with plain loads.
static void ensure_tmap(eb200_ctx* c) {
  const long long sig[6] = {(long long)(uintptr_t)c->fbase, c->g.n[0], c->g.n[1], c->g.n[2], c->shape, c->ndims};
  if (c->tmap_valid && std::equal(sig, sig + 6, c->tmap_sig)) return;
  std::copy(sig, sig + 6, c->tmap_sig);
  c->tmap_valid = true; c->tmap_ok = false;
  if (!c->use_tma || c->ndims < 2) return;
  const int nd = c->ndims;
  const cuuint64_t ex = (cuuint64_t)(c->g.n[0] + 2 * c->g.ngd[0]), ey = (cuuint64_t)(c->g.n[1] + 2 * c->g.ngd[1]),
                   ez = (cuuint64_t)(nd == 3 ? c->g.n[2] + 2 * c->g.ngd[2] : 1);
  const int fx3[3] = {TL<3, 0>::FX, TL<3, 1>::FX, TL<3, 2>::FX}, fx2[3] = {TL<2, 0>::FX, TL<2, 1>::FX, TL<2, 2>::FX};
  if (fx3[c->shape] % 2 != 0 || fx2[c->shape] % 2 != 0) return;
  if (ex % 2 != 0 || ((uintptr_t)c->fbase & 15) != 0 || (c->g.total * sizeof(double)) % 16 != 0) return;
  typedef CUresult (*Encode)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qr;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn ||
      qr != cudaDriverEntryPointSuccess) { cudaGetLastError(); return; }
  cuuint32_t box3[4], box2[3];
  const int fy3[3] = {TL<3, 0>::FY, TL<3, 1>::FY, TL<3, 2>::FY}, fz3[3] = {TL<3, 0>::FZ, TL<3, 1>::FZ, TL<3, 2>::FZ};
  const int fy2[3] = {TL<2, 0>::FY, TL<2, 1>::FY, TL<2, 2>::FY};
  box3[0] = fx3[c->shape]; box3[1] = fy3[c->shape]; box3[2] = fz3[c->shape]; box2[0] = fx2[c->shape]; box2[1] = fy2[c->shape];
  box2[2] = 6;
  if (EB_ZCOMP) { box3[3] = box3[2]; box3[2] = 6; }     // 3D tensor order: x, y, component, z
  else box3[3] = 6;
  const cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r;
  if (nd == 3) {
    const cuuint64_t dims[4] = {ex, ey, EB_ZCOMP ? 6 : ez, EB_ZCOMP ? ez : 6};
    const cuuint64_t strides[3] = {ex * 8, EB_ZCOMP ? (cuuint64_t)c->g.total * 8 : ex * ey * 8,
                                   EB_ZCOMP ? ex * ey * 8 : (cuuint64_t)c->g.total * 8};
    r = ((Encode)fn)((CUtensorMap*)c->tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, c->fbase, dims, strides, box3, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    const cuuint64_t dims[3] = {ex, ey, 6};
    const cuuint64_t strides[2] = {ex * 8, (cuuint64_t)c->g.total * 8};
    r = ((Encode)fn)((CUtensorMap*)c->tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, c->fbase, dims, strides, box2, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  c->tmap_ok = r == CUDA_SUCCESS;
}

int32_t launch_push_split(eb200_ctx* c, SpeciesD& sp, const PushP& p, bool* handled, const void* keysink) {
  *handled = false;
  KeySink ks;
  if (keysink) ks = *(const KeySink*)keysink; else { ks.on = 0; ks.count = nullptr; ks.nbucket = 0; ks.mig_axis = -1; }
  if (!sp.sorted || !sp.h_tile_start) return 0;
  // deferred permute: read the pre-sort order through src, write the sorted order into the other state
  const bool gather = sp.deferred;
  bool fwd = true;
  long long off_new = sp.off, room = 1LL << 62;
  if (gather) {
    fwd = sp.off >= sp.gap + sp.win;
    off_new = fwd ? sp.gap : sp.gap + sp.win;
    const long long W = fwd ? sp.off - off_new : off_new - sp.off;
    room = W - (fwd ? sp.disp_fwd : sp.disp_bwd);       // particles per launch (see defer_plan, particles.cu)
  }
  PArr a, o, side;
  for (int d = 0; d < 3; ++d) {
    a.x[d] = sp.a[d]; a.m[d] = sp.a[3 + d];
    o.x[d] = sp.base[d] ? sp.base[d] + off_new : nullptr; o.m[d] = sp.base[3 + d] + off_new;
    side.x[d] = sp.far ? sp.far + (size_t)d * sp.far_cap : nullptr; side.m[d] = sp.far ? sp.far + (size_t)(3 + d) * sp.far_cap : nullptr;
  }
  a.w = sp.a[6]; o.w = sp.base[6] + off_new; side.w = sp.far ? sp.far + (size_t)6 * sp.far_cap : nullptr;
  // a species with particle probes takes the kernel variant with the crossing test; the probe set sits in device memory
  const bool probe = sp.probes.n > 0;
  const ProbeSet* d_probes = nullptr;
  if (probe) {
    if (!c->d_probeset) EB_CUDA(c, cudaMalloc(&c->d_probeset, sizeof(ProbeSet)));
    EB_CUDA(c, cudaMemcpyAsync(c->d_probeset, &sp.probes, sizeof(ProbeSet), cudaMemcpyHostToDevice, c->stream));
    d_probes = (const ProbeSet*)c->d_probeset;
  }
  const uint32_t n_main = gather ? (uint32_t)sp.n_old : 0u;
  ks.key = sp.key_base + off_new;
  ensure_tmap(c);
  const CUtensorMap tmap = *(const CUtensorMap*)c->tmap;
  const int use_tma = c->tmap_ok ? 1 : 0;
  const long long ntiles = (long long)c->g.ntile[0] * c->g.ntile[1] * c->g.ntile[2];
  const uint32_t* ts = sp.h_tile_start;      // ntiles + 1 entries, valid since the sort's synchronise
  // scratch: 64 B per particle of a slab of tiles
  long long want = sp.n < c->rec_scratch_max ? sp.n : c->rec_scratch_max;
  long long biggest = 0;
  for (long long t = 0; t < ntiles; ++t) biggest = biggest > (long long)(ts[t + 1] - ts[t]) ? biggest : (long long)(ts[t + 1] - ts[t]);
  if (want < biggest) want = biggest;       // a slab is at least one tile
  if (want < 1) return 0;
  if (gather && room < biggest) return fail(c, EB200_E_STATE, "deferred permute: window smaller than a tile");
  if (c->rec_scratch_cap < want) {
    if (c->rec_scratch) { EB_CUDA(c, cudaStreamSynchronize(c->stream)); cudaFree(c->rec_scratch); }
    c->rec_scratch = nullptr; c->rec_scratch_cap = 0;
    EB_CUDA(c, cudaMalloc(&c->rec_scratch, (size_t)want * 8 * sizeof(double)));
    c->rec_scratch_cap = want;
  }
  // deterministic mode: one J tile per deposit CTA (1 GB of them at most per launch)
  const bool det = c->deterministic && !p.zero_current;
  // option push_overlap: two halves of the scratch, the deposit of a slab on a second stream beside the next slab's push
  const bool overlap = c->push_overlap && !det && !p.zero_current && c->rec_scratch_cap / 2 >= biggest;
  if (overlap && !c->stream2) {
    EB_CUDA(c, cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking));
    for (int k = 0; k < 2; ++k) {
      EB_CUDA(c, cudaEventCreateWithFlags(&c->ev_push[k], cudaEventDisableTiming));
      EB_CUDA(c, cudaEventCreateWithFlags(&c->ev_dep[k], cudaEventDisableTiming));
    }
  }
  const long long scratch_slab = overlap ? c->rec_scratch_cap / 2 : c->rec_scratch_cap;
  const long long cap = scratch_slab < room ? scratch_slab : room;
  const int jn3[3] = {TL<3, 0>::JN, TL<3, 1>::JN, TL<3, 2>::JN}, jn2[3] = {TL<2, 0>::JN, TL<2, 1>::JN, TL<2, 2>::JN};
  const int jn1[3] = {TL<1, 0>::JN, TL<1, 1>::JN, TL<1, 2>::JN};
  const size_t tile_doubles = 3 * (size_t)(c->ndims == 3 ? jn3[c->shape] : (c->ndims == 2 ? jn2[c->shape] : jn1[c->shape]));   // >= the sub-tile's
  const long long max_tiles = det ? std::max<long long>(1, (long long)(((size_t)1 << 30) / (tile_doubles * 8 * SUB))) : (1LL << 62);
  if (det && c->det_tiles_cap < (size_t)std::min<long long>(max_tiles, ntiles) * SUB * tile_doubles) {
    if (c->det_tiles) { EB_CUDA(c, cudaStreamSynchronize(c->stream)); cudaFree(c->det_tiles); }
    c->det_tiles = nullptr; c->det_tiles_cap = 0;
    const size_t want_t = (size_t)std::min<long long>(max_tiles, ntiles) * SUB * tile_doubles;
    EB_CUDA(c, cudaMalloc(&c->det_tiles, want_t * sizeof(double)));
    c->det_tiles_cap = want_t;
  }
  // slabs of consecutive tiles, run upwards (or downwards: deferred permute towards higher slots)
  std::vector<std::pair<long long, long long>> slabs;
  for (long long t0 = 0; t0 < ntiles;) {
    if (ts[t0 + 1] == ts[t0]) { ++t0; continue; }               // skip leading empty tiles
    long long t1 = t0 + 1;
    while (t1 < ntiles && (long long)(ts[t1 + 1] - ts[t0]) <= cap && t1 - t0 < max_tiles) ++t1;
    slabs.emplace_back(t0, t1);
    t0 = t1;
  }
  for (size_t is = 0; is < slabs.size(); ++is) {
    const long long t0 = slabs[fwd ? is : slabs.size() - 1 - is].first, t1 = slabs[fwd ? is : slabs.size() - 1 - is].second;
    const unsigned nb = (unsigned)(t1 - t0);
    long long big = 1;
    for (long long t = t0; t < t1; ++t) big = std::max<long long>(big, (long long)(ts[t + 1] - ts[t]));
    const dim3 ga(nb, (unsigned)((big + PUSH_CHUNK - 1) / PUSH_CHUNK));
    double* const scr = c->rec_scratch + (overlap ? (size_t)(is & 1) * (size_t)scratch_slab * 8 : 0);
    cudaStream_t sdep = overlap ? c->stream2 : c->stream;
    if (overlap && is >= 2) EB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_dep[is & 1], 0));   // the half is free again
#define L(ND, SH)                                                                                          \
    if (c->ndims == ND && c->shape == SH) {                                                                \
      auto ka = probe ? (gather ? k_push_split<ND, SH, true, true> : k_push_split<ND, SH, false, true>)    \
                      : (gather ? k_push_split<ND, SH, true, false> : k_push_split<ND, SH, false, false>);  \
      auto kb = det ? k_deposit_split<ND, SH, true> : k_deposit_split<ND, SH, false>;                      \
      const size_t sa = sizeof(double) * (6 * TL<ND, SH>::FN + (NTA / 32) * (EB_STAGE_HALF ? 128 : 256)) + EB_SMEM_PAD;                                               \
      const size_t sb = sizeof(double) * (det ? 6 : 3) * TL<ND, SH>::JNS + sizeof(int) * (2 * MCAPB + 4)                     \
                        + (size_t)RING * 64 * (TL<ND, SH>::CPT / SUB) + 16;                                      \
      EB_CUDA(c, cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sa));          \
      EB_CUDA(c, cudaFuncSetAttribute(kb, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sb));          \
      { ScopedPhase pa(c, "push_A");                                                                       \
        ka<<<ga, NTA, sa, c->stream>>>(c->g, c->f, p, a, o, side, sp.dest, n_main, sp.cell_start, t0, scr, c->d_counters, ks, tmap, use_tma, d_probes); } \
      if (overlap) { cudaEventRecord(c->ev_push[is & 1], c->stream); cudaStreamWaitEvent(sdep, c->ev_push[is & 1], 0); } \
      if (!p.zero_current) { ScopedPhase pb(c, "push_B");   /* (push_overlap: issued on the second stream, not timed here) */ \
        kb<<<nb * SUB, TL<ND, SH>::CPT / SUB, sb, sdep>>>(c->g, c->f, p, sp.cell_start, t0, scr, c->det_tiles); \
        if (overlap) cudaEventRecord(c->ev_dep[is & 1], sdep);                                             \
        if (det) {                                                                                         \
          /* planes of the slowest axis that the tiles [t0, t1) can reach */                              \
          const long long per_layer = ND == 3 ? (long long)c->g.ntile[0] * c->g.ntile[1] : (ND == 2 ? c->g.ntile[0] : 1);  \
          const int TDS = tile_dim(ND, ND - 1);                                                            \
          int lo = (1 - KM) + (int)(t0 / per_layer) * TDS + TL<ND, SH>::SFM - 1;                           \
          int hi = (1 - KM) + (int)((t1 - 1) / per_layer) * TDS + TL<ND, SH>::SFM - 1 + TDS + TL<ND, SH>::NW; \
          lo = std::max(lo, 1 - c->g.ngd[ND - 1]); hi = std::min(hi, c->g.n[ND - 1] + c->g.ngd[ND - 1]);   \
          if (hi >= lo) {                                                                                  \
            const long long cells = (long long)(hi - lo + 1) * (ND >= 2 ? c->g.n[0] + 2 * c->g.ngd[0] : 1) * \
                                    (ND == 3 ? (c->g.n[1] + 2 * c->g.ngd[1]) : 1);                         \
            k_reduce_tiles<ND, SH><<<(unsigned)((cells + 255) / 256), 256, 0, c->stream>>>(c->g, c->f, sp.cell_start, t0, t1, \
                                                                                           c->det_tiles, lo, hi);           \
            c->launches++;                                                                                 \
          }                                                                                                \
        } }                                                                                                \
    }
    L(1, 0) L(1, 1) L(1, 2) L(2, 0) L(2, 1) L(2, 2) L(3, 0) L(3, 1) L(3, 2)
#undef L
    c->launches += p.zero_current ? 1 : 2;
  }
  if (overlap)      // join: everything after the push waits for the last deposits
    for (size_t k = 0; k < std::min<size_t>(2, slabs.size()); ++k) EB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_dep[k], 0));
  EB_CUDA(c, cudaGetLastError());
  if (gather) { set_view(sp, off_new); sp.deferred = false; }
  *handled = true;
  return 0;
}

}  // namespace eb
