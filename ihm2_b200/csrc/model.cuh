// model.cuh -- fkin6 controller model with hand-derived Jacobian, curvature lookup.
//
// Follows the reference definitions (paths relative to the reference tree):
//   fkin6 ODE .......... python/models.py:232-307
//   smooth_sgn ......... python/utils.py:23-24   (tanh(10 x))
//   constants .......... python/constants.py:43-111 (become the per-instance vector p[NP])
//   kappa interpolant .. python/models.py:290-295 (linear, lookup_mode "exact", extrapolating)
#pragma once
#include "common.cuh"

namespace ihm2b {

struct TrackView {
  int n;
  const double* s;
  const double* k;
};

// Largest i with s[i] <= sq, clamped to [0, n-2] (what the binary search of the "exact"
// lookup returns); found from a uniform-grid guess plus a local walk, O(1) on the
// reference's tables (three copies of a uniform lap grid, motion_planning.py:402-409).
template <typename Real>
__device__ __forceinline__ void kappa_lookup(const TrackView& tr, Real sq, Real& kap,
                                             Real& dkap) {
  const int n = tr.n;
  const double s0 = __ldg(tr.s), s1 = __ldg(tr.s + n - 1);
  double g = ((double)sq - s0) / (s1 - s0) * (double)(n - 1);
  int i = (g > 0.0) ? ((g < (double)(n - 2)) ? (int)g : n - 2) : 0;  // NaN -> 0
  while (i > 0 && __ldg(tr.s + i) > sq) i--;
  while (i < n - 2 && __ldg(tr.s + i + 1) <= sq) i++;
  const Real sa = __ldg(tr.s + i), sb = __ldg(tr.s + i + 1);
  const Real ka = __ldg(tr.k + i), kb = __ldg(tr.k + i + 1);
  dkap = (kb - ka) / (sb - sa);
  kap = ka + dkap * (sq - sa);
}

// Index of the table segment that contains sq (the same clamped "exact" lookup), for tables other than kappa.
__device__ __forceinline__ int table_segment(const TrackView& tr, double sq) {
  const int n = tr.n;
  const double s0 = __ldg(tr.s), s1 = __ldg(tr.s + n - 1);
  double g = (sq - s0) / (s1 - s0) * (double)(n - 1);
  int i = (g > 0.0) ? ((g < (double)(n - 2)) ? (int)g : n - 2) : 0;
  while (i > 0 && __ldg(tr.s + i) > sq) i--;
  while (i < n - 2 && __ldg(tr.s + i + 1) <= sq) i++;
  return i;
}

// Sparsity patterns (bit j = column j of (x0..x7, u0, u1)).
//   JMASK[i]: structural non-zeros of row i of d f / d (x,u)
//   SMASK[i]: structural non-zeros of row i of any stage sensitivity d k / d (x0,u) and of
//             d x_next / d (x0,u) minus the identity (closed under the RK recursions)
__device__ constexpr unsigned JMASK[NX] = {0x01F, 0x01C, 0x03F, 0x2D8, 0x2C8, 0x2C8, 0x140, 0x280};
__device__ constexpr unsigned SMASK[NX] = {0x3FF, 0x3FF, 0x3FF, 0x3D8, 0x3D8, 0x3D8, 0x140, 0x280};

// f[8] and, if JAC, J[8][10] (only the JMASK entries are written).  Real: double, or float for the mixed-precision
// lineariser (ihm2b_config::linearize_fp32).
template <bool JAC, typename Real>
__device__ __forceinline__ void fkin6(const Real* __restrict__ x, const Real u0,
                                      const Real u1, const Real* __restrict__ p,
                                      const TrackView& tr, Real* __restrict__ f,
                                      Real (*J)[NZ]) {
  const Real n = x[1], psi = x[2], vx = x[3], vy = x[4], r = x[5], T = x[6], dl = x[7];
  const Real m = p[IHM2B_P_M], lR = p[IHM2B_P_LR], lF = p[IHM2B_P_LF];
  const Real C = lR / (lR + lF);
  const Real itT = Real(1.0) / p[IHM2B_P_TT], itD = Real(1.0) / p[IHM2B_P_TD], im = Real(1.0) / m;

  const Real ddot = (u1 - dl) * itD;
  const Real Tdot = (u0 - T) * itT;

  const Real Fm = p[IHM2B_P_CM0] * T;
  const Real th = tanh(Real(10.0) * vx);
  const Real poly = p[IHM2B_P_CR0] + p[IHM2B_P_CR1] * vx + p[IHM2B_P_CR2] * vx * vx;
  const Real Fd = -poly * th;
  const Real FRx = Real(0.5) * Fm + Fd, FFx = Real(0.5) * Fm;

  // beta = atan(C tan(delta)): only sin/cos of beta are needed
  Real sd, cd;
  sincos(dl, &sd, &cd);
  const Real td = sd / cd;
  const Real Ctd = C * td;
  const Real den = Real(1.0) + Ctd * Ctd;
  const Real cb = rsqrt(den), sb = Ctd * cb;
  const Real sec2 = Real(1.0) + td * td;
  const Real G = C * sec2 / den;  // d beta / d delta
  const Real bdot = G * ddot;
  const Real cdb = cd * cb + sd * sb, sdb = sd * cb - cd * sb;  // cos / sin (delta - beta)
  const Real vdot = (FRx * cb + FFx * cdb) * im;

  Real kap, dkap;
  kappa_lookup(tr, x[0], kap, dkap);
  Real sp, cp;
  sincos(psi, &sp, &cp);
  const Real iD = Real(1.0) / (Real(1.0) + kap * n);
  const Real num = vx * cp - vy * sp;
  const Real sdot = num * iD;
  const Real vydot = vdot * sb + bdot * vx;

  f[0] = sdot;
  f[1] = vx * sp + vy * cp;
  f[2] = r - kap * sdot;
  f[3] = vdot * cb - bdot * vy;
  f[4] = vydot;
  f[5] = lR * vydot - bdot;  // literal, python/models.py:304
  f[6] = Tdot;
  f[7] = ddot;
  if (!JAC) return;

  const Real ds_ds = -sdot * dkap * n * iD;
  const Real ds_dn = -sdot * kap * iD;
  const Real ds_dpsi = -(vx * sp + vy * cp) * iD;
  const Real ds_dvx = cp * iD, ds_dvy = -sp * iD;
  J[0][0] = ds_ds; J[0][1] = ds_dn; J[0][2] = ds_dpsi; J[0][3] = ds_dvx; J[0][4] = ds_dvy;
  J[1][2] = num; J[1][3] = sp; J[1][4] = cp;
  J[2][0] = -dkap * sdot - kap * ds_ds;
  J[2][1] = -kap * ds_dn;
  J[2][2] = -kap * ds_dpsi;
  J[2][3] = -kap * ds_dvx;
  J[2][4] = -kap * ds_dvy;
  J[2][5] = Real(1.0);
  const Real dFd_dvx =
      -(p[IHM2B_P_CR1] + Real(2.0) * p[IHM2B_P_CR2] * vx) * th - poly * Real(10.0) * (Real(1.0) - th * th);
  const Real dG_ddl = Real(2.0) * C * (Real(1.0) - C * C) * td * sec2 / (den * den);
  const Real dbdot_ddl = dG_ddl * ddot - G * itD;
  const Real dbdot_du = G * itD;
  const Real dvdot_dvx = dFd_dvx * cb * im;
  const Real dvdot_dT = Real(0.5) * p[IHM2B_P_CM0] * (cb + cdb) * im;
  const Real dvdot_ddl = (-FRx * sb * G - FFx * sdb * (Real(1.0) - G)) * im;
  J[3][3] = dvdot_dvx * cb;
  J[3][4] = -bdot;
  J[3][6] = dvdot_dT * cb;
  J[3][7] = dvdot_ddl * cb - vdot * sb * G - dbdot_ddl * vy;
  J[3][9] = -dbdot_du * vy;
  J[4][3] = dvdot_dvx * sb + bdot;
  J[4][6] = dvdot_dT * sb;
  J[4][7] = dvdot_ddl * sb + vdot * cb * G + dbdot_ddl * vx;
  J[4][9] = dbdot_du * vx;
  J[5][3] = lR * J[4][3];
  J[5][6] = lR * J[4][6];
  J[5][7] = lR * J[4][7] - dbdot_ddl;
  J[5][9] = lR * J[4][9] - dbdot_du;
  J[6][6] = -itT; J[6][8] = itT;
  J[7][7] = -itD; J[7][9] = itD;
}

// Nonlinear path constraints h(x, u) of the gen-1 OCP (old/generate_acaods_interface.py:312-325 for the 8-state
// Frenet model; bounds and slack sets :351-449) on the quantities of the controller model above:
//   h0 =  n - L/2 sin|psi| + W/2 cos psi - right_width(s),   h1 = -n + L/2 sin|psi| + W/2 cos psi - left_width(s)
//   h2 = (u_T - T) / t_T,   h3 = (u_delta - delta) / t_delta
//   h4 = a_lat = (-F_Rx sin beta + F_Fx sin(delta - beta)) / m + (v_x^2 + v_y^2) sin(beta) / l_R
// (width tables: linear interpolants on s_ref, old/generate.py:44-46).  Writes the 16 doubles of Work::HC for one
// stage: the structurally non-zero Jacobian entries and e_i = h_i - J_i z (row value = J_i z + e_i).
// term: only h0, h1 exist at the terminal stage (con_h_expr_e), the other entries are zero.
__device__ __forceinline__ void con_h_rows(const double* __restrict__ x, const double u0, const double u1,
                                           const double* __restrict__ p, const TrackView& tr,
                                           const double* __restrict__ wr, const double* __restrict__ wl,
                                           const double half_L, const double half_W, const bool term,
                                           double* __restrict__ out) {
  const double n = x[1], psi = x[2], vx = x[3], vy = x[4], T = x[6], dl = x[7];
  const int i = table_segment(tr, x[0]);
  const double sa = __ldg(tr.s + i), ids = 1.0 / (__ldg(tr.s + i + 1) - sa);
  const double ra = __ldg(wr + i), la = __ldg(wl + i);
  const double dwr = (__ldg(wr + i + 1) - ra) * ids, dwl = (__ldg(wl + i + 1) - la) * ids;
  const double wrv = ra + dwr * (x[0] - sa), wlv = la + dwl * (x[0] - sa);
  double sp, cp, sa_, ca_;
  sincos(psi, &sp, &cp);
  sincos(fabs(psi), &sa_, &ca_);
  const double sg = (psi > 0.0) ? 1.0 : (psi < 0.0) ? -1.0 : 0.0;   // d |psi| / d psi (CasADi: sign)
  const double dabs = ca_ * sg;
  const double h0 = n - half_L * sa_ + half_W * cp - wrv;
  const double h1 = -n + half_L * sa_ + half_W * cp - wlv;
  const double j02 = -half_L * dabs - half_W * sp, j12 = half_L * dabs - half_W * sp;
  out[0] = -dwr; out[1] = 1.0; out[2] = j02;
  out[3] = h0 - (-dwr) * x[0] - n - j02 * psi;
  out[4] = -dwl; out[5] = -1.0; out[6] = j12;
  out[7] = h1 - (-dwl) * x[0] + n - j12 * psi;
  if (term) {
#pragma unroll
    for (int j = 8; j < 16; j++) out[j] = 0.0;
    return;
  }
  const double m = p[IHM2B_P_M], lR = p[IHM2B_P_LR], lF = p[IHM2B_P_LF];
  const double C = lR / (lR + lF), im = 1.0 / m;
  const double itT = 1.0 / p[IHM2B_P_TT], itD = 1.0 / p[IHM2B_P_TD];
  const double h2 = (u0 - T) / p[IHM2B_P_TT], h3 = (u1 - dl) / p[IHM2B_P_TD];
  const double Fm = p[IHM2B_P_CM0] * T;
  const double th = tanh(10.0 * vx);
  const double poly = p[IHM2B_P_CR0] + p[IHM2B_P_CR1] * vx + p[IHM2B_P_CR2] * vx * vx;
  const double FRx = 0.5 * Fm - poly * th, FFx = 0.5 * Fm;
  double sd, cd;
  sincos(dl, &sd, &cd);
  const double td = sd / cd, Ctd = C * td, den = 1.0 + Ctd * Ctd;
  const double cb = rsqrt(den), sb = Ctd * cb;
  const double G = C * (1.0 + td * td) / den;
  const double cdb = cd * cb + sd * sb, sdb = sd * cb - cd * sb;
  const double v2 = vx * vx + vy * vy;
  const double h4 = (-FRx * sb + FFx * sdb) * im + v2 * sb / lR;
  const double dFd = -(p[IHM2B_P_CR1] + 2.0 * p[IHM2B_P_CR2] * vx) * th - poly * 10.0 * (1.0 - th * th);
  const double j43 = -dFd * sb * im + 2.0 * vx * sb / lR;
  const double j44 = 2.0 * vy * sb / lR;
  const double j46 = 0.5 * p[IHM2B_P_CM0] * (sdb - sb) * im;
  const double j47 = (-FRx * cb * G + FFx * cdb * (1.0 - G)) * im + v2 * cb * G / lR;
  out[8] = j43; out[9] = j44; out[10] = j46; out[11] = j47;
  out[12] = h4 - j43 * vx - j44 * vy - j46 * T - j47 * dl;
  out[13] = h2 - (-itT) * T - itT * u0;
  out[14] = h3 - (-itD) * dl - itD * u1;
  out[15] = 0.0;
}

}  // namespace ihm2b
