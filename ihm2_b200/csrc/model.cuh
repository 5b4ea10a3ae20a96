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
__device__ __forceinline__ void kappa_lookup(const TrackView& tr, double sq, double& kap,
                                             double& dkap) {
  const int n = tr.n;
  const double s0 = __ldg(tr.s), s1 = __ldg(tr.s + n - 1);
  double g = (sq - s0) / (s1 - s0) * (double)(n - 1);
  int i = (g > 0.0) ? ((g < (double)(n - 2)) ? (int)g : n - 2) : 0;  // NaN -> 0
  while (i > 0 && __ldg(tr.s + i) > sq) i--;
  while (i < n - 2 && __ldg(tr.s + i + 1) <= sq) i++;
  const double sa = __ldg(tr.s + i), sb = __ldg(tr.s + i + 1);
  const double ka = __ldg(tr.k + i), kb = __ldg(tr.k + i + 1);
  dkap = (kb - ka) / (sb - sa);
  kap = ka + dkap * (sq - sa);
}

// Sparsity patterns (bit j = column j of (x0..x7, u0, u1)).
//   JMASK[i]: structural non-zeros of row i of d f / d (x,u)
//   SMASK[i]: structural non-zeros of row i of any stage sensitivity d k / d (x0,u) and of
//             d x_next / d (x0,u) minus the identity (closed under the RK recursions)
__device__ constexpr unsigned JMASK[NX] = {0x01F, 0x01C, 0x03F, 0x2D8, 0x2C8, 0x2C8, 0x140, 0x280};
__device__ constexpr unsigned SMASK[NX] = {0x3FF, 0x3FF, 0x3FF, 0x3D8, 0x3D8, 0x3D8, 0x140, 0x280};

// f[8] and, if JAC, J[8][10] (only the JMASK entries are written).
template <bool JAC>
__device__ __forceinline__ void fkin6(const double* __restrict__ x, const double u0,
                                      const double u1, const double* __restrict__ p,
                                      const TrackView& tr, double* __restrict__ f,
                                      double (*J)[NZ]) {
  const double n = x[1], psi = x[2], vx = x[3], vy = x[4], r = x[5], T = x[6], dl = x[7];
  const double m = p[IHM2B_P_M], lR = p[IHM2B_P_LR], lF = p[IHM2B_P_LF];
  const double C = lR / (lR + lF);
  const double itT = 1.0 / p[IHM2B_P_TT], itD = 1.0 / p[IHM2B_P_TD], im = 1.0 / m;

  const double ddot = (u1 - dl) * itD;
  const double Tdot = (u0 - T) * itT;

  const double Fm = p[IHM2B_P_CM0] * T;
  const double th = tanh(10.0 * vx);
  const double poly = p[IHM2B_P_CR0] + p[IHM2B_P_CR1] * vx + p[IHM2B_P_CR2] * vx * vx;
  const double Fd = -poly * th;
  const double FRx = 0.5 * Fm + Fd, FFx = 0.5 * Fm;

  // beta = atan(C tan(delta)): only sin/cos of beta are needed
  double sd, cd;
  sincos(dl, &sd, &cd);
  const double td = sd / cd;
  const double Ctd = C * td;
  const double den = 1.0 + Ctd * Ctd;
  const double cb = rsqrt(den), sb = Ctd * cb;
  const double sec2 = 1.0 + td * td;
  const double G = C * sec2 / den;  // d beta / d delta
  const double bdot = G * ddot;
  const double cdb = cd * cb + sd * sb, sdb = sd * cb - cd * sb;  // cos / sin (delta - beta)
  const double vdot = (FRx * cb + FFx * cdb) * im;

  double kap, dkap;
  kappa_lookup(tr, x[0], kap, dkap);
  double sp, cp;
  sincos(psi, &sp, &cp);
  const double iD = 1.0 / (1.0 + kap * n);
  const double num = vx * cp - vy * sp;
  const double sdot = num * iD;
  const double vydot = vdot * sb + bdot * vx;

  f[0] = sdot;
  f[1] = vx * sp + vy * cp;
  f[2] = r - kap * sdot;
  f[3] = vdot * cb - bdot * vy;
  f[4] = vydot;
  f[5] = lR * vydot - bdot;  // literal, python/models.py:304
  f[6] = Tdot;
  f[7] = ddot;
  if (!JAC) return;

  const double ds_ds = -sdot * dkap * n * iD;
  const double ds_dn = -sdot * kap * iD;
  const double ds_dpsi = -(vx * sp + vy * cp) * iD;
  const double ds_dvx = cp * iD, ds_dvy = -sp * iD;
  J[0][0] = ds_ds; J[0][1] = ds_dn; J[0][2] = ds_dpsi; J[0][3] = ds_dvx; J[0][4] = ds_dvy;
  J[1][2] = num; J[1][3] = sp; J[1][4] = cp;
  J[2][0] = -dkap * sdot - kap * ds_ds;
  J[2][1] = -kap * ds_dn;
  J[2][2] = -kap * ds_dpsi;
  J[2][3] = -kap * ds_dvx;
  J[2][4] = -kap * ds_dvy;
  J[2][5] = 1.0;
  const double dFd_dvx =
      -(p[IHM2B_P_CR1] + 2.0 * p[IHM2B_P_CR2] * vx) * th - poly * 10.0 * (1.0 - th * th);
  const double dG_ddl = 2.0 * C * (1.0 - C * C) * td * sec2 / (den * den);
  const double dbdot_ddl = dG_ddl * ddot - G * itD;
  const double dbdot_du = G * itD;
  const double dvdot_dvx = dFd_dvx * cb * im;
  const double dvdot_dT = 0.5 * p[IHM2B_P_CM0] * (cb + cdb) * im;
  const double dvdot_ddl = (-FRx * sb * G - FFx * sdb * (1.0 - G)) * im;
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

}  // namespace ihm2b
