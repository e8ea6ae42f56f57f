// oracle/orc_fields.cpp -- CPU ORACLE (test infrastructure only).
// Restates epochNd/src/fields.f90 (update_e_field, update_b_field,
// update_eb_fields_half/final), the field parts of epochNd/src/boundary.F90
// (field_bc, field_clamp_zero, field_zero_gradient, efield_bcs, bfield_bcs,
// bfield_final_bcs, processor_summation_bcs) and outflow_bcs_* of laser.f90.
#include "orc_state.hpp"

namespace orc {

// stagger(dir, comp): setup.F90:134-144
static inline bool stagger(int dir, int comp) {
  switch (comp) {
    case EX: case JX: return dir == 0;
    case EY: case JY: return dir == 1;
    case EZ: case JZ: return dir == 2;
    case BX: return dir != 0;
    case BY: return dir != 1;
    case BZ: return dir != 2;
  }
  return false;
}

// visit every line along `axis` (all positions of the other dims, ghosts included);
// fn(base) where element with Fortran index c along axis is base + (c+ng-1)*stride
template <class F>
static void for_each_line(const State& s, int axis, F fn) {
  int a1 = (axis + 1) % 3, a2 = (axis + 2) % 3;
  int e1 = s.n[a1] + 2 * s.ngd[a1], e2 = s.n[a2] + 2 * s.ngd[a2];
  for (int q = 0; q < e2; ++q)
    for (int p = 0; p < e1; ++p) fn((int64_t)p * s.st[a1] + (int64_t)q * s.st[a2]);
}

void set_half_step_scalars(State& s) {
  // fields.f90:821-830
  s.hdt = 0.5 * s.dt;
  for (int d = 0; d < 3; ++d) {
    s.hdtd[d] = s.hdt / s.d[d];
    s.cn[d] = s.hdtd[d] * (C * C);
  }
  s.fac = s.hdt / EPS0;
}

// fields.f90:183-424 (non-CPML branches; 1D epoch1d/src/fields.f90:76, 2D epoch2d:104)
void update_e_field(State& s) {
  const int nd = s.ndims;
  double* ex = s.f[EX].data(); double* ey = s.f[EY].data(); double* ez = s.f[EZ].data();
  const double* bx = s.f[BX].data(); const double* by = s.f[BY].data(); const double* bz = s.f[BZ].data();
  const double* jx = s.f[JX].data(); const double* jy = s.f[JY].data(); const double* jz = s.f[JZ].data();
  const int64_t sx = s.st[0], sy = s.st[1], sz = s.st[2];
  const double fac = s.fac;
  double c1 = 1, c2 = 0, c3 = 0;
  if (s.field_order == 4) { c1 = 9.0 / 8.0; c2 = -1.0 / 24.0; }
  else if (s.field_order == 6) { c1 = 75.0 / 64.0; c2 = -25.0 / 384.0; c3 = 3.0 / 640.0; }
  const int ord = s.field_order;
  double cx1 = ord == 2 ? s.cn[0] : c1 * s.cn[0], cx2 = c2 * s.cn[0], cx3 = c3 * s.cn[0];
  double cy1 = ord == 2 ? s.cn[1] : c1 * s.cn[1], cy2 = c2 * s.cn[1], cy3 = c3 * s.cn[1];
  double cz1 = ord == 2 ? s.cn[2] : c1 * s.cn[2], cz2 = c2 * s.cn[2], cz3 = c3 * s.cn[2];
  // CPML branches (fields.f90:192-310): every coefficient is divided by cpml_kappa_e*(index)
  auto kap = [&](int d, int idx) { return s.cpml_prof[d][0][idx + s.ng - 1]; };
  for (int iz = s.lo0(2); iz <= s.hi(2); ++iz)
    for (int iy = s.lo0(1); iy <= s.hi(1); ++iy)
      for (int ix = 0; ix <= s.n[0]; ++ix) {
        const int64_t i = s.idx(ix, iy, iz);
        if (s.cpml_on) {
          const double qx = s.cn[0] / kap(0, ix);
          cx1 = ord == 2 ? qx : c1 * qx; cx2 = c2 * qx; cx3 = c3 * qx;
          if (nd >= 2) { const double qy = s.cn[1] / kap(1, iy); cy1 = ord == 2 ? qy : c1 * qy; cy2 = c2 * qy; cy3 = c3 * qy; }
          if (nd >= 3) { const double qz = s.cn[2] / kap(2, iz); cz1 = ord == 2 ? qz : c1 * qz; cz2 = c2 * qz; cz3 = c3 * qz; }
        }
        // the reference evaluates left to right: e + A - B - fac*j
        double vx = ex[i], vy = ey[i], vz = ez[i];
        // ex: + cy*(d_y bz) - cz*(d_z by)
        if (nd >= 2) {
          vx = vx + cy1 * (bz[i] - bz[i - sy]);
          if (ord >= 4) vx = vx + cy2 * (bz[i + sy] - bz[i - 2 * sy]);
          if (ord >= 6) vx = vx + cy3 * (bz[i + 2 * sy] - bz[i - 3 * sy]);
        }
        if (nd >= 3) {
          vx = vx - cz1 * (by[i] - by[i - sz]);
          if (ord >= 4) vx = vx - cz2 * (by[i + sz] - by[i - 2 * sz]);
          if (ord >= 6) vx = vx - cz3 * (by[i + 2 * sz] - by[i - 3 * sz]);
        }
        vx = vx - fac * jx[i];
        // ey: + cz*(d_z bx) - cx*(d_x bz)
        if (nd >= 3) {
          vy = vy + cz1 * (bx[i] - bx[i - sz]);
          if (ord >= 4) vy = vy + cz2 * (bx[i + sz] - bx[i - 2 * sz]);
          if (ord >= 6) vy = vy + cz3 * (bx[i + 2 * sz] - bx[i - 3 * sz]);
        }
        vy = vy - cx1 * (bz[i] - bz[i - sx]);
        if (ord >= 4) vy = vy - cx2 * (bz[i + sx] - bz[i - 2 * sx]);
        if (ord >= 6) vy = vy - cx3 * (bz[i + 2 * sx] - bz[i - 3 * sx]);
        vy = vy - fac * jy[i];
        // ez: + cx*(d_x by) - cy*(d_y bx)
        vz = vz + cx1 * (by[i] - by[i - sx]);
        if (ord >= 4) vz = vz + cx2 * (by[i + sx] - by[i - 2 * sx]);
        if (ord >= 6) vz = vz + cx3 * (by[i + 2 * sx] - by[i - 3 * sx]);
        if (nd >= 2) {
          vz = vz - cy1 * (bx[i] - bx[i - sy]);
          if (ord >= 4) vz = vz - cy2 * (bx[i + sy] - bx[i - 2 * sy]);
          if (ord >= 6) vz = vz - cy3 * (bx[i + 2 * sy] - bx[i - 3 * sy]);
        }
        vz = vz - fac * jz[i];
        ex[i] = vx; ey[i] = vy; ez[i] = vz;
      }
  if (s.cpml_on) cpml_advance_currents(s, true, s.hdt);
}

// difference along axis a of the extended stencils (fields.f90:655-732; 2D :426-470; 1D :296-306),
// terms and summation order as written there; so1 / so2 = strides of the other axes in index
// order, h1 / h2 = whether they exist
static inline double ext_diff(const double* f, int64_t i, int64_t sa, int64_t so1, int64_t so2, bool h1, bool h2,
                              double alpha, double beta1, double beta2, double gamma, double delta) {
  double r = alpha * (f[i + sa] - f[i]);
  if (h1) r = r + beta1 * (f[i + sa + so1] - f[i + so1] + f[i + sa - so1] - f[i - so1]);
  if (h2) r = r + beta2 * (f[i + sa + so2] - f[i + so2] + f[i + sa - so2] - f[i - so2]);
  if (h1 && h2)
    r = r + gamma * (f[i + sa + so1 - so2] - f[i + so1 - so2] + f[i + sa - so1 - so2] - f[i - so1 - so2]
                   + f[i + sa + so1 + so2] - f[i + so1 + so2] + f[i + sa - so1 + so2] - f[i - so1 + so2]);
  r = r + delta * (f[i + 2 * sa] - f[i - sa]);
  return r;
}

// update_b_field, second order, maxwell_solver /= yee (fields.f90:655-732)
static void update_b_field_ext(State& s) {
  const int nd = s.ndims;
  const double* ex = s.f[EX].data(); const double* ey = s.f[EY].data(); const double* ez = s.f[EZ].data();
  double* bx = s.f[BX].data(); double* by = s.f[BY].data(); double* bz = s.f[BZ].data();
  const int64_t sx = s.st[0], sy = s.st[1], sz = s.st[2];
  const bool hy = nd >= 2, hz = nd >= 3;
  const double* q = s.coeffs;   // alpha x,y,z | beta xy,xz,yx,yz,zx,zy | gamma x,y,z | delta x,y,z
  double cx1 = s.hdtd[0], cy1 = s.hdtd[1], cz1 = s.hdtd[2];
  for (int iz = s.lo0(2); iz <= s.hi(2); ++iz)
    for (int iy = s.lo0(1); iy <= s.hi(1); ++iy)
      for (int ix = 0; ix <= s.n[0]; ++ix) {
        const int64_t i = s.idx(ix, iy, iz);
        if (s.cpml_on) {            // cx1 = hdtx / cpml_kappa_bx(ix) (2D fields.f90:338-341)
          cx1 = s.hdtd[0] / s.cpml_prof[0][1][ix + s.ng - 1];
          if (hy) cy1 = s.hdtd[1] / s.cpml_prof[1][1][iy + s.ng - 1];
          if (hz) cz1 = s.hdtd[2] / s.cpml_prof[2][1][iz + s.ng - 1];
        }
        double vx = bx[i], vy = by[i], vz = bz[i];
        if (hy) vx = vx - cy1 * ext_diff(ez, i, sy, sx, sz, true, hz, q[1], q[5], q[6], q[10], q[13]);
        if (hz) vx = vx + cz1 * ext_diff(ey, i, sz, sx, sy, true, true, q[2], q[7], q[8], q[11], q[14]);
        if (hz) vy = vy - cz1 * ext_diff(ex, i, sz, sx, sy, true, true, q[2], q[7], q[8], q[11], q[14]);
        vy = vy + cx1 * ext_diff(ez, i, sx, sy, sz, hy, hz, q[0], q[3], q[4], q[9], q[12]);
        vz = vz - cx1 * ext_diff(ey, i, sx, sy, sz, hy, hz, q[0], q[3], q[4], q[9], q[12]);
        if (hy) vz = vz + cy1 * ext_diff(ex, i, sy, sx, sz, true, hz, q[1], q[5], q[6], q[10], q[13]);
        bx[i] = vx; by[i] = vy; bz[i] = vz;
      }
  if (s.cpml_on) cpml_advance_currents(s, false, s.hdt);
}

// fields.f90:428-815 (Yee branches :637-654, orders 4/6 :734-812)
void update_b_field(State& s) {
  if (s.maxwell_solver != 0 && s.field_order == 2) { update_b_field_ext(s); return; }
  const int nd = s.ndims;
  const double* ex = s.f[EX].data(); const double* ey = s.f[EY].data(); const double* ez = s.f[EZ].data();
  double* bx = s.f[BX].data(); double* by = s.f[BY].data(); double* bz = s.f[BZ].data();
  const int64_t sx = s.st[0], sy = s.st[1], sz = s.st[2];
  double c1 = 1, c2 = 0, c3 = 0;
  if (s.field_order == 4) { c1 = 9.0 / 8.0; c2 = -1.0 / 24.0; }
  else if (s.field_order == 6) { c1 = 75.0 / 64.0; c2 = -25.0 / 384.0; c3 = 3.0 / 640.0; }
  const int ord = s.field_order;
  double cx1 = ord == 2 ? s.hdtd[0] : c1 * s.hdtd[0], cx2 = c2 * s.hdtd[0], cx3 = c3 * s.hdtd[0];
  double cy1 = ord == 2 ? s.hdtd[1] : c1 * s.hdtd[1], cy2 = c2 * s.hdtd[1], cy3 = c3 * s.hdtd[1];
  double cz1 = ord == 2 ? s.hdtd[2] : c1 * s.hdtd[2], cz2 = c2 * s.hdtd[2], cz3 = c3 * s.hdtd[2];
  auto kap = [&](int d, int idx) { return s.cpml_prof[d][1][idx + s.ng - 1]; };
  for (int iz = s.lo0(2); iz <= s.hi(2); ++iz)
    for (int iy = s.lo0(1); iy <= s.hi(1); ++iy)
      for (int ix = 0; ix <= s.n[0]; ++ix) {
        const int64_t i = s.idx(ix, iy, iz);
        if (s.cpml_on) {            // fields.f90:437-630
          const double qx = s.hdtd[0] / kap(0, ix);
          cx1 = ord == 2 ? qx : c1 * qx; cx2 = c2 * qx; cx3 = c3 * qx;
          if (nd >= 2) { const double qy = s.hdtd[1] / kap(1, iy); cy1 = ord == 2 ? qy : c1 * qy; cy2 = c2 * qy; cy3 = c3 * qy; }
          if (nd >= 3) { const double qz = s.hdtd[2] / kap(2, iz); cz1 = ord == 2 ? qz : c1 * qz; cz2 = c2 * qz; cz3 = c3 * qz; }
        }
        double vx = bx[i], vy = by[i], vz = bz[i];
        // bx: - cy*(d_y ez) + cz*(d_z ey)
        if (nd >= 2) {
          vx = vx - cy1 * (ez[i + sy] - ez[i]);
          if (ord >= 4) vx = vx - cy2 * (ez[i + 2 * sy] - ez[i - sy]);
          if (ord >= 6) vx = vx - cy3 * (ez[i + 3 * sy] - ez[i - 2 * sy]);
        }
        if (nd >= 3) {
          vx = vx + cz1 * (ey[i + sz] - ey[i]);
          if (ord >= 4) vx = vx + cz2 * (ey[i + 2 * sz] - ey[i - sz]);
          if (ord >= 6) vx = vx + cz3 * (ey[i + 3 * sz] - ey[i - 2 * sz]);
        }
        // by: - cz*(d_z ex) + cx*(d_x ez)
        if (nd >= 3) {
          vy = vy - cz1 * (ex[i + sz] - ex[i]);
          if (ord >= 4) vy = vy - cz2 * (ex[i + 2 * sz] - ex[i - sz]);
          if (ord >= 6) vy = vy - cz3 * (ex[i + 3 * sz] - ex[i - 2 * sz]);
        }
        vy = vy + cx1 * (ez[i + sx] - ez[i]);
        if (ord >= 4) vy = vy + cx2 * (ez[i + 2 * sx] - ez[i - sx]);
        if (ord >= 6) vy = vy + cx3 * (ez[i + 3 * sx] - ez[i - 2 * sx]);
        // bz: - cx*(d_x ey) + cy*(d_y ex)
        vz = vz - cx1 * (ey[i + sx] - ey[i]);
        if (ord >= 4) vz = vz - cx2 * (ey[i + 2 * sx] - ey[i - sx]);
        if (ord >= 6) vz = vz - cx3 * (ey[i + 3 * sx] - ey[i - 2 * sx]);
        if (nd >= 2) {
          vz = vz + cy1 * (ex[i + sy] - ex[i]);
          if (ord >= 4) vz = vz + cy2 * (ex[i + 2 * sy] - ex[i - sy]);
          if (ord >= 6) vz = vz + cy3 * (ex[i + 3 * sy] - ex[i - 2 * sy]);
        }
        bx[i] = vx; by[i] = vy; bz[i] = vz;
      }
  if (s.cpml_on) cpml_advance_currents(s, false, s.hdt);
}

// cpml_advance_e_currents / cpml_advance_b_currents (boundary.F90:2483-2815; 1D :998-1106).  The
// six blocks per routine are cyclic permutations: axis a, b = a + 1, c = a + 2,
//   E: psi_e{b}{a} <- d_a B_c, E_b -= fac psi;  psi_e{c}{a} <- d_a B_b, E_c += fac psi
//   B: psi_b{b}{a} <- d_a E_c, B_b += t psi;    psi_b{c}{a} <- d_a E_b, B_c -= t psi
// over the slab [start, end] (B on a max side: start-1 .. end-1) and 1..n of the other axes.
void cpml_advance_currents(State& s, bool efield, double tstep) {
  const double fac = tstep * (C * C);
  for (int a = 0; a < s.ndims; ++a)
    for (int side = 0; side < 2; ++side) {
      const int bnd = 2 * a + side;
      if (s.bc_field[bnd] != BC_CPML_LASER && s.bc_field[bnd] != BC_CPML_OUTFLOW) continue;
      int start = s.cpml_start[bnd], end = s.cpml_end[bnd];
      if (!efield && side == 1) { start -= 1; end -= 1; }
      const int b = (a + 1) % 3, c = (a + 2) % 3;
      const std::vector<double>& kp = s.cpml_prof[a][efield ? 0 : 1];
      const std::vector<double>& sg = s.cpml_prof[a][efield ? 2 : 3];
      const std::vector<double>& ac = s.cpml_prof[a][efield ? 4 : 5];
      double* psi0 = s.cpml_psi[a][efield ? 0 : 2].data();
      double* psi1 = s.cpml_psi[a][efield ? 1 : 3].data();
      double* F0 = s.f[(efield ? EX : BX) + b].data();
      double* F1 = s.f[(efield ? EX : BX) + c].data();
      const double* S0 = s.f[(efield ? BX : EX) + c].data();
      const double* S1 = s.f[(efield ? BX : EX) + b].data();
      const int64_t sa = s.st[a];
      int lo[3] = {1, 1, 1}, hi[3] = {1, 1, 1};
      for (int d = 0; d < s.ndims; ++d) hi[d] = s.n[d];
      lo[a] = start; hi[a] = end;
      for (int iz = lo[2]; iz <= hi[2]; ++iz)
        for (int iy = lo[1]; iy <= hi[1]; ++iy)
          for (int ix = lo[0]; ix <= hi[0]; ++ix) {
            const int cc[3] = {ix, iy, iz};
            const int ipos = cc[a];
            const int64_t i = s.idx(ix, iy, iz);
            const double kappa = kp[ipos + s.ng - 1], sigma = sg[ipos + s.ng - 1], acoeff = ac[ipos + s.ng - 1];
            const double bcoeff = std::exp(-(sigma / kappa + acoeff) * tstep);
            const double ccoeff_d = (bcoeff - 1.0) * sigma / kappa / (sigma + kappa * acoeff) / s.d[a];
            if (efield) {
              psi0[i] = bcoeff * psi0[i] + ccoeff_d * (S0[i] - S0[i - sa]);
              F0[i] = F0[i] - fac * psi0[i];
              psi1[i] = bcoeff * psi1[i] + ccoeff_d * (S1[i] - S1[i - sa]);
              F1[i] = F1[i] + fac * psi1[i];
            } else {
              psi0[i] = bcoeff * psi0[i] + ccoeff_d * (S0[i + sa] - S0[i]);
              F0[i] = F0[i] + tstep * psi0[i];
              psi1[i] = bcoeff * psi1[i] + ccoeff_d * (S1[i + sa] - S1[i]);
              F1[i] = F1[i] - tstep * psi1[i];
            }
          }
    }
}

// boundary.F90:145-152,321-468 do_field_mpi_with_lengths, single-rank form: the
// x_min neighbour of a periodic rank is itself, so A(nx+1:nx+ng) <- A(1:ng) and
// A(1-ng:0) <- A(nx+1-ng:nx); x then y then z over FULL extents of the other dims;
// on a non-periodic physical boundary the ghost cells are left untouched.
void field_bc(State& s, int comp) { field_bc_array(s, s.f[comp].data()); }
void field_bc_array(State& s, double* a) {
  const int ng = s.ng;
  for (int ax = 0; ax < s.ndims; ++ax) {
    const int nn = s.n[ax];
    const int64_t sa = s.st[ax];
    const bool per_max = s.bc_field[2 * ax + 1] == BC_PERIODIC;
    const bool per_min = s.bc_field[2 * ax] == BC_PERIODIC;
    for_each_line(s, ax, [&](int64_t base) {
      if (per_max)
        for (int i = 1; i <= ng; ++i) a[base + (nn + i + ng - 1) * sa] = a[base + (i + ng - 1) * sa];
      if (per_min)
        for (int i = 1; i <= ng; ++i) a[base + (i - ng + ng - 1) * sa] = a[base + (nn + i - ng + ng - 1) * sa];
    });
  }
}

// boundary.F90:623-698
static void field_zero_gradient(State& s, int comp, int boundary /*0..5*/) {
  if (boundary / 2 >= s.ndims) return;
  field_zero_gradient_array(s, s.f[comp].data(), stagger(boundary / 2, comp), boundary);
}
void field_zero_gradient_array(State& s, double* a, bool stg, int boundary) {
  if (s.bc_field[boundary] == BC_PERIODIC) return;
  const int ax = boundary / 2;
  if (ax >= s.ndims) return;
  const bool is_max = boundary & 1;
  if (is_max ? !s.max_bnd[ax] : !s.min_bnd[ax]) return;
  const int ng = s.ng, nn = s.n[ax];
  const int64_t sa = s.st[ax];
  for_each_line(s, ax, [&](int64_t base) {
    auto A = [&](int c) -> double& { return a[base + (int64_t)(c + ng - 1) * sa]; };
    if (!is_max) {
      if (stg) for (int i = 1; i <= ng - 1; ++i) A(i - ng) = A(ng - i);
      else for (int i = 1; i <= ng; ++i) A(i - ng) = A(ng + 1 - i);
    } else {
      if (stg) for (int i = 1; i <= ng - 1; ++i) A(nn + i) = A(nn - i);
      else for (int i = 1; i <= ng; ++i) A(nn + i) = A(nn + 1 - i);
    }
  });
}

// boundary.F90:702-783
static void field_clamp_zero(State& s, int comp, int boundary) {
  if (s.bc_field[boundary] == BC_PERIODIC) return;
  const int ax = boundary / 2;
  if (ax >= s.ndims) return;
  const bool is_max = boundary & 1;
  if (is_max ? !s.max_bnd[ax] : !s.min_bnd[ax]) return;
  double* a = s.f[comp].data();
  const int ng = s.ng, nn = s.n[ax];
  const int64_t sa = s.st[ax];
  const bool stg = stagger(ax, comp);
  for_each_line(s, ax, [&](int64_t base) {
    auto A = [&](int c) -> double& { return a[base + (int64_t)(c + ng - 1) * sa]; };
    if (!is_max) {
      if (stg) { for (int i = 1; i <= ng - 1; ++i) A(i - ng) = -A(ng - i); A(0) = 0.0; }
      else for (int i = 1; i <= ng; ++i) A(i - ng) = -A(ng + 1 - i);
    } else {
      if (stg) { A(nn) = 0.0; for (int i = 1; i <= ng - 1; ++i) A(nn + i) = -A(nn - i); }
      else for (int i = 1; i <= ng; ++i) A(nn + i) = -A(nn + 1 - i);
    }
  });
}

// boundary.F90:1148-1202
void efield_bcs(State& s) {
  field_bc(s, EX); field_bc(s, EY); field_bc(s, EZ);
  for (int ax = 0; ax < s.ndims; ++ax)
    for (int side = 0; side < 2; ++side) {
      const int b = 2 * ax + side;
      if (s.bc_field[b] == BC_CONDUCT)
        for (int c = 0; c < 3; ++c) {
          if (c == ax) field_clamp_zero(s, EX + c, b);
          else field_zero_gradient(s, EX + c, b);
        }
    }
  for (int b = 0; b < 2 * s.ndims; ++b) {
    const int bc = s.bc_field[b];
    if (bc == BC_CLAMP || bc == BC_SIMPLE_LASER || bc == BC_SIMPLE_OUTFLOW)
      for (int c = 0; c < 3; ++c) field_clamp_zero(s, EX + c, b);
    if (bc == BC_ZERO_GRADIENT || bc == BC_CPML_LASER || bc == BC_CPML_OUTFLOW)
      for (int c = 0; c < 3; ++c) field_zero_gradient(s, EX + c, b);
  }
}

// boundary.F90:1206-1263
void bfield_bcs(State& s, bool mpi_only) {
  field_bc(s, BX); field_bc(s, BY); field_bc(s, BZ);
  if (mpi_only) return;
  for (int ax = 0; ax < s.ndims; ++ax)
    for (int side = 0; side < 2; ++side) {
      const int b = 2 * ax + side;
      if (s.bc_field[b] == BC_CONDUCT)
        for (int c = 0; c < 3; ++c) {
          if (c == ax) field_zero_gradient(s, BX + c, b);
          else field_clamp_zero(s, BX + c, b);
        }
    }
  for (int b = 0; b < 2 * s.ndims; ++b) {
    const int bc = s.bc_field[b];
    if (bc == BC_CLAMP || bc == BC_SIMPLE_LASER || bc == BC_SIMPLE_OUTFLOW)
      for (int c = 0; c < 3; ++c) field_clamp_zero(s, BX + c, b);
    if (bc == BC_ZERO_GRADIENT || bc == BC_CPML_LASER || bc == BC_CPML_OUTFLOW)
      for (int c = 0; c < 3; ++c) field_zero_gradient(s, BX + c, b);
  }
}

// laser.f90:509-973 outflow_bcs_{x,y,z}_{min,max} (1D epoch1d/src/laser.f90:303,
// 2D epoch2d/src/laser.f90:411).  The six routines are cyclic permutations of one
// another: normal axis a, tangential t1=a+1, t2=a+2.  The boundary planes ex_x_min ... (the fields the run started
// with, setup_field_boundaries) enter as in laser.f90:536,559-572.

// index of grid point c[] in a plane normal to axis a (other axes incl. ghosts, lower axis fastest)
static inline size_t plane_index(const State& s, int a, const int* c) {
  const int lo = a == 0 ? 1 : 0, hi = a == 2 ? 1 : 2;
  const size_t ext_lo = (size_t)(s.n[lo] + 2 * s.ngd[lo]);
  return (size_t)(c[lo] + s.ngd[lo] - 1) + ext_lo * (size_t)(c[hi] + s.ngd[hi] - 1);
}

// housekeeping/setup.F90:426-510: e?_x_min = the field at nx0 (tangential E, normal B) or the mean of the
// planes nx0 and nx0 - 1 (normal E, tangential B); nx0 = 1 / nx, or one beyond the laser plane of a cpml_laser side
void setup_field_boundaries(State& s) {
  for (int b = 0; b < 2 * s.ndims; ++b) {
    const int a = b / 2;
    const bool is_max = b & 1;
    int n0 = is_max ? s.n[a] : 1;
    if (s.bc_field[b] == BC_CPML_LASER) n0 = is_max ? s.cpml_laser_idx[b] + 1 : s.cpml_laser_idx[b] - 1;
    const int lo = a == 0 ? 1 : 0, hi = a == 2 ? 1 : 2;
    const size_t cnt = (size_t)(s.n[lo] + 2 * s.ngd[lo]) * (size_t)(s.n[hi] + 2 * s.ngd[hi]);
    for (int k = 0; k < 6; ++k) s.bnd_planes[b][k].assign(cnt, 0.0);
    for (int q = 1 - s.ngd[hi]; q <= s.n[hi] + s.ngd[hi]; ++q)
      for (int p = 1 - s.ngd[lo]; p <= s.n[lo] + s.ngd[lo]; ++p) {
        int c[3]; c[a] = n0; c[lo] = p; c[hi] = q;
        const int64_t i = s.idx(c[0], c[1], c[2]);
        const size_t l = plane_index(s, a, c);
        for (int k = 0; k < 3; ++k) {
          const double* E = s.f[EX + k].data(); const double* B = s.f[BX + k].data();
          s.bnd_planes[b][EX + k][l] = k == a ? 0.5 * (E[i] + E[i - s.st[a]]) : E[i];
          s.bnd_planes[b][BX + k][l] = k == a ? B[i] : 0.5 * (B[i] + B[i - s.st[a]]);
        }
      }
  }
}

static void outflow_bcs(State& s, int boundary) {
  const int a = boundary / 2, t1 = (a + 1) % 3, t2 = (a + 2) % 3;
  const bool is_max = boundary & 1;
  int lp = is_max ? s.n[a] : 1;
  if (s.bc_field[boundary] == BC_CPML_LASER) lp = s.cpml_laser_idx[boundary];   // laser.f90:529-531
  const double dtc2 = s.dt * (C * C);
  const double la = dtc2 / s.d[a];
  const double l1 = t1 < s.ndims ? dtc2 / s.d[t1] : 0.0;
  const double l2 = t2 < s.ndims ? dtc2 / s.d[t2] : 0.0;
  const double sum = 1.0 / (la + C), diff = la - C, dt_eps = s.dt / EPS0;
  double* Ba = s.f[BX + a].data(); double* B1 = s.f[BX + t1].data(); double* B2 = s.f[BX + t2].data();
  const double* E1 = s.f[EX + t1].data(); const double* E2 = s.f[EX + t2].data();
  const double* J1 = s.f[JX + t1].data(); const double* J2 = s.f[JX + t2].data();
  // plane index ranges 0..n over existing tangential dims; plane storage is
  // x-fastest in (lower axis, higher axis) order like the reference's (0:ny,0:nz)
  const int lo_axis = t1 < t2 ? t1 : t2, hi_axis = t1 < t2 ? t2 : t1;
  const int nlo = lo_axis < s.ndims ? s.n[lo_axis] + 1 : 1;
  const int nhi = hi_axis < s.ndims ? s.n[hi_axis] + 1 : 1;
  const int64_t sa = s.st[a];
  for (int q = 0; q < nhi; ++q)
    for (int p = 0; p < nlo; ++p) {
      int c[3];
      c[a] = lp;
      c[lo_axis] = lo_axis < s.ndims ? p : 1;
      c[hi_axis] = hi_axis < s.ndims ? q : 1;
      const int64_t i = s.idx(c[0], c[1], c[2]);
      double source1 = 0.0, source2 = 0.0;
      if (s.add_laser[boundary]) {
        for (const Laser& L : s.lasers[boundary]) {
          if (s.time >= L.t_start && s.time <= L.t_end) {
            const size_t k = (size_t)q * nlo + p;
            const double prof = L.profile.empty() ? 1.0 : L.profile[k];
            const double ph = L.phase.empty() ? 0.0 : L.phase[k];
            const double t_env = L.t_env * L.amp;
            const double base = t_env * prof * std::sin(L.integral_phase + ph);
            source1 = source1 + base * std::cos(L.pol);
            source2 = source2 + base * std::sin(L.pol);
          }
        }
      }
      const bool have = !s.bnd_planes[boundary][0].empty();
      const size_t pl = have ? plane_index(s, a, c) : 0;
      const double PBa = have ? s.bnd_planes[boundary][BX + a][pl] : 0.0;
      const double PE1 = have ? s.bnd_planes[boundary][EX + t1][pl] : 0.0, PE2 = have ? s.bnd_planes[boundary][EX + t2][pl] : 0.0;
      const double PB1 = have ? s.bnd_planes[boundary][BX + t1][pl] : 0.0, PB2 = have ? s.bnd_planes[boundary][BX + t2][pl] : 0.0;
      if (!is_max) {
        Ba[i - sa] = PBa;  // bx(laserpos-1) = bx_x_min
        double v2 = 4.0 * source1 + 2.0 * (PE1 + C * PB2) - 2.0 * E1[i];
        if (t2 < s.ndims) v2 = v2 - l2 * (Ba[i] - Ba[i - s.st[t2]]);
        v2 = v2 + dt_eps * J1[i] + diff * B2[i];
        B2[i - sa] = sum * v2;
        double v1 = -4.0 * source2 - 2.0 * (PE2 - C * PB1) + 2.0 * E2[i];
        if (t1 < s.ndims) v1 = v1 - l1 * (Ba[i] - Ba[i - s.st[t1]]);
        v1 = v1 - dt_eps * J2[i] + diff * B1[i];
        B1[i - sa] = sum * v1;
      } else {
        Ba[i + sa] = PBa;  // bx(laserpos+1) = bx_x_max
        double v2 = -4.0 * source1 - 2.0 * (PE1 - C * PB2) + 2.0 * E1[i];
        if (t2 < s.ndims) v2 = v2 + l2 * (Ba[i] - Ba[i - s.st[t2]]);
        v2 = v2 - dt_eps * J1[i] + diff * B2[i - sa];
        B2[i] = sum * v2;
        double v1 = 4.0 * source2 + 2.0 * (PE2 + C * PB1) - 2.0 * E2[i];
        if (t1 < s.ndims) v1 = v1 + l1 * (Ba[i] - Ba[i - s.st[t1]]);
        v1 = v1 + dt_eps * J2[i] + diff * B1[i - sa];
        B1[i] = sum * v1;
      }
    }
}

// boundary.F90:1267-1312
void bfield_final_bcs(State& s) {
  bfield_bcs(s, false);
  for (int b = 0; b < 2 * s.ndims; ++b) {
    const int ax = b / 2;
    const bool on_bnd = (b & 1) ? s.max_bnd[ax] : s.min_bnd[ax];
    if (!on_bnd) continue;
    if (s.add_laser[b] || s.bc_field[b] == BC_SIMPLE_OUTFLOW) outflow_bcs(s, b);
  }
  bfield_bcs(s, true);
}

// fields.f90:819-847
void update_eb_fields_half(State& s) {
  set_half_step_scalars(s);
  update_e_field(s);
  efield_bcs(s);
  update_b_field(s);
  bfield_bcs(s, true);
}

// fields.f90:851-872
void update_eb_fields_final(State& s) {
  set_half_step_scalars(s);
  update_b_field(s);
  bfield_final_bcs(s);
  update_e_field(s);
  efield_bcs(s);
}

// bc_allspecies (deck_species_block.F90:231-245): common particle BC or mixed
int bc_allspecies_of(const State& s, int b) {
  int bc = -1;
  for (const Species& sp : s.species) {
    if (!sp.defined) continue;
    if (bc == -1) bc = sp.bc[b];
    else if (bc != sp.bc[b]) return BC_MIXED;
  }
  return bc == -1 ? BC_PERIODIC : bc;
}

// boundary.F90:787-919 particle_reflection_bcs + :923-1084 particle_periodic_bcs
static void processor_summation_bcs(State& s, int comp, int flip_dir /*0..2*/) {
  processor_summation_array(s, s.f[comp].data(), flip_dir);
}
bool bc_mixed(const State& s) {
  for (int b = 0; b < 2 * s.ndims; ++b) if (bc_allspecies_of(s, b) == BC_MIXED) return true;
  return false;
}
void processor_summation_array(State& s, double* a, int flip_dir, const int* bc6) {
  const int ng = s.ng;
  auto bc_allspecies = [&](const State& st, int b) { return bc6 ? bc6[b] : ::orc::bc_allspecies_of(st, b); };
  // reflection first
  for (int ax = 0; ax < s.ndims; ++ax) {
    const int nn = s.n[ax];
    const int64_t sa = s.st[ax];
    const bool rmin = s.min_bnd[ax] && bc_allspecies(s, 2 * ax) == BC_REFLECT;
    const bool rmax = s.max_bnd[ax] && bc_allspecies(s, 2 * ax + 1) == BC_REFLECT;
    if (!rmin && !rmax) continue;
    for_each_line(s, ax, [&](int64_t base) {
      auto A = [&](int c) -> double& { return a[base + (int64_t)(c + ng - 1) * sa]; };
      if (rmin) {
        if (flip_dir == ax) for (int i = 1; i <= ng - 1; ++i) { A(i) = A(i) - A(-i); A(-i) = 0.0; }
        else for (int i = 1; i <= ng - 1; ++i) { A(i) = A(i) + A(1 - i); A(1 - i) = 0.0; }
      }
      if (rmax) {
        if (flip_dir == ax) for (int i = 1; i <= ng; ++i) { A(nn - i) = A(nn - i) - A(nn + i); A(nn + i) = 0.0; }
        else for (int i = 1; i <= ng; ++i) { A(nn + 1 - i) = A(nn + 1 - i) + A(nn + i); A(nn + i) = 0.0; }
      }
    });
  }
  // periodic / subdomain sums, x then y then z (single rank: neighbour is self if
  // the species boundary is periodic, else MPI_PROC_NULL -> temp stays 0)
  for (int ax = 0; ax < s.ndims; ++ax) {
    const int nn = s.n[ax];
    const int64_t sa = s.st[ax];
    const bool pmin = !(s.min_bnd[ax] && bc_allspecies(s, 2 * ax) != BC_PERIODIC);
    const bool pmax = !(s.max_bnd[ax] && bc_allspecies(s, 2 * ax + 1) != BC_PERIODIC);
    std::vector<double> temp(ng);
    for_each_line(s, ax, [&](int64_t base) {
      auto A = [&](int c) -> double& { return a[base + (int64_t)(c + ng - 1) * sa]; };
      // send A(nn+1:nn+ng) to +1 neighbour, receive from -1 neighbour, add to A(1:ng)
      // single rank: receive own slab iff both the send (+1) and recv (-1) sides exist
      if (pmax && pmin) {
        for (int i = 1; i <= ng; ++i) temp[i - 1] = A(nn + i);
        for (int i = 1; i <= ng; ++i) A(i) = A(i) + temp[i - 1];
        for (int i = 1; i <= ng; ++i) temp[i - 1] = A(i - ng);
        for (int i = 1; i <= ng; ++i) A(nn - ng + i) = A(nn - ng + i) + temp[i - 1];
      }
    });
  }
  if (bc6) {      // particle_clear_bcs: every ghost zone
    for (int ax = 0; ax < s.ndims; ++ax) {
      const int nn = s.n[ax];
      const int64_t sa = s.st[ax];
      for_each_line(s, ax, [&](int64_t base) {
        for (int i = 1; i <= ng; ++i) { a[base + (int64_t)(ng - i) * sa] = 0.0; a[base + (int64_t)(nn + i + ng - 1) * sa] = 0.0; }
      });
    }
  }
}

// smooth_array, housekeeping/current_smooth.F90:61-151 (default build: strided compensated
// binomial filter).  wk holds the array with ghosts; every pass refreshes wk's ghosts with
// field_bc, writes alpha * wk + beta * (neighbours at +-stride) into the array's interior and
// copies that back into wk.  beta is computed once; alpha changes only after a pass with
// it > its -- both exactly as the reference does.
static void smooth_array(State& s, int comp) {
  std::vector<double>& arr = s.f[comp];
  std::vector<double> wk = arr;            // ng_l = MAX(sng, jng) = jng for strides <= ng
  std::vector<int> strides = s.smooth_strides;
  if (strides.empty()) strides.push_back(1);
  double alpha = 0.5;
  const double beta = (1.0 - alpha) / (2.0 * s.ndims);
  for (int it = 1; it <= s.smooth_its + s.smooth_comp_its; ++it) {
    for (int cs : strides) {
      std::swap(arr, wk);                  // field_bc works on s.f[comp]
      field_bc(s, comp);
      std::swap(arr, wk);
      for (int iz = (s.ndims >= 3 ? 1 : 1); iz <= (s.ndims >= 3 ? s.n[2] : 1); ++iz)
        for (int iy = 1; iy <= (s.ndims >= 2 ? s.n[1] : 1); ++iy)
          for (int ix = 1; ix <= s.n[0]; ++ix) {
            const int64_t i = s.idx(ix, iy, iz);
            double nb = wk[i - cs * s.st[0]] + wk[i + cs * s.st[0]];
            if (s.ndims >= 2) { nb = nb + wk[i - cs * s.st[1]]; nb = nb + wk[i + cs * s.st[1]]; }
            if (s.ndims >= 3) { nb = nb + wk[i - cs * s.st[2]]; nb = nb + wk[i + cs * s.st[2]]; }
            arr[i] = alpha * wk[i] + nb * beta;
          }
      for (int iz = 1; iz <= (s.ndims >= 3 ? s.n[2] : 1); ++iz)
        for (int iy = 1; iy <= (s.ndims >= 2 ? s.n[1] : 1); ++iy)
          for (int ix = 1; ix <= s.n[0]; ++ix) { const int64_t i = s.idx(ix, iy, iz); wk[i] = arr[i]; }
    }
    if (it > s.smooth_its) alpha = (double)s.smooth_its * 0.5 + 1.0;
  }
}

// housekeeping/current_smooth.F90:29-45
void current_finish(State& s) {
  if (!bc_mixed(s)) {       // (mixed: done species by species inside push_particles, boundary.F90:1130-1136)
    processor_summation_bcs(s, JX, 0);
    processor_summation_bcs(s, JY, 1);
    processor_summation_bcs(s, JZ, 2);
  }
  field_bc(s, JX); field_bc(s, JY); field_bc(s, JZ);
  if (s.smooth) { smooth_array(s, JX); smooth_array(s, JY); smooth_array(s, JZ); }
  if (s.use_current_correction)
    for (int k = 0; k < 3; ++k)
      for (double& v : s.f[JX + k]) v = v - s.initial_j[k];
}

// One iteration of shift_window's loop without insert_particles (housekeeping/window.F90:62-194,
// :362-384): grid scalars of setup_grid_x, remove_particles, shift_fields.
void shift_window(State& s, double grid_min_local, double min_local, double max_local, double gmin, double gmax) {
  s.grid_min_local[0] = grid_min_local; s.min_local[0] = min_local; s.max_local[0] = max_local;
  s.gmin[0] = gmin; s.gmax[0] = gmax;
  if (s.min_bnd[0]) {
    for (Species& sp : s.species) {
      if (!sp.defined) continue;
      size_t w = 0;
      for (size_t i = 0; i < sp.count(); ++i) {
        if (sp.pos[0][i] < gmin) continue;               // current%part_pos(1) < x_min: destroyed
        if (w != i) {
          for (int d = 0; d < 3; ++d) { if (d < s.ndims) sp.pos[d][w] = sp.pos[d][i]; sp.p[d][w] = sp.p[d][i]; }
          sp.w[w] = sp.w[i];
        }
        ++w;
      }
      for (int d = 0; d < 3; ++d) { if (d < s.ndims) sp.pos[d].resize(w); sp.p[d].resize(w); }
      sp.w.resize(w);
    }
  }
  const int len = s.n[0] + 2 * s.ngd[0];
  const int64_t nlines = s.total / len;
  for (int f = 0; f < NFIELD; ++f) {
    double* a = s.f[f].data();
    for (int64_t l = 0; l < nlines; ++l)
      for (int i = 0; i < len - 1; ++i) a[l * len + i] = a[l * len + i + 1];
    field_bc(s, f);
  }
  // the CPML psi arrays move with the fields (window.F90:114-129) ...
  for (int ax = 0; ax < s.ndims; ++ax)
    for (int k = 0; k < 4; ++k) {
      if (s.cpml_psi[ax][k].empty()) continue;
      double* a = s.cpml_psi[ax][k].data();
      for (int64_t l = 0; l < nlines; ++l)
        for (int i = 0; i < len - 1; ++i) a[l * len + i] = a[l * len + i + 1];
      field_bc_array(s, a);
      // ... and psi(nx+1) = psi(nx) on the x_max boundary (window.F90:151-169)
      if (s.max_bnd[0])
        for (int64_t l = 0; l < nlines; ++l) a[l * len + s.ngd[0] + s.n[0]] = a[l * len + s.ngd[0] - 1 + s.n[0]];
    }
  if (s.max_bnd[0]) {
    double *ex = s.f[EX].data(), *ey = s.f[EY].data(), *ez = s.f[EZ].data();
    double *bx = s.f[BX].data(), *by = s.f[BY].data(), *bz = s.f[BZ].data();
    const int nx = s.n[0];
    for (int64_t l = 0; l < nlines; ++l) {
      const int64_t o = l * len + s.ngd[0] - 1;          // o + i = element with x index i
      auto pl = [&](int k) { return s.win_planes[k].empty() ? 0.0 : s.win_planes[k][l]; };
      ex[o + nx] = pl(0);
      ex[o + nx + 1] = pl(0);
      ey[o + nx + 1] = pl(1);
      ez[o + nx + 1] = pl(2);
      ex[o + nx - 1] = 0.5 * (ex[o + nx - 2] + ex[o + nx]);
      ey[o + nx] = 0.5 * (ey[o + nx - 1] + ey[o + nx + 1]);
      ez[o + nx] = 0.5 * (ez[o + nx - 1] + ez[o + nx + 1]);
      bx[o + nx + 1] = pl(3);
      by[o + nx] = pl(4);
      bz[o + nx] = pl(5);
      bx[o + nx] = 0.5 * (bx[o + nx - 1] + bx[o + nx + 1]);
      by[o + nx - 1] = 0.5 * (by[o + nx - 2] + by[o + nx]);
      bz[o + nx - 1] = 0.5 * (bz[o + nx - 2] + bz[o + nx]);
    }
  }
}

}  // namespace orc
