// ihm2_oracle.cpp -- CPU fp64 restatement of the ihm2 NMPC hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing in ihm2_b200/ may import, link or call this file.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs use it, as the checker and the timed CPU baseline.
//
// PARITY UNPINNED: the reference (tudoroancea/ihm2) ships no tests, fixtures or golden
// vectors, and the arithmetic of this path lives in acados / HPIPM / BLASFEO / CasADi,
// which are neither vendored nor version-pinned (README.md:54-57, env.yml:17-19) and are
// absent from this image.  This file restates (i) the reference's own model/OCP
// definitions and (ii) the published algorithms of those libraries (ERK4 / Gauss-Legendre
// and Radau-IIA collocation with forward sensitivities, Gauss-Newton SQP(-RTI),
// Mehrotra primal-dual IPM with a Riccati KKT solve).  It is pinned only by the
// independent mathematical checks in tests/ (finite differences, scipy ODE solutions,
// KKT optimality certificates).
//
// Everything here is deliberately generic and dense (plain loops over small dense
// matrices); the CUDA product path is a separate, structured implementation.
//
// Reference sites restated (paths relative to the reference tree):
//   fkin6 ODE ......................... python/models.py:232-307
//   smooth_sgn ........................ python/utils.py:23-24
//   constants ......................... python/constants.py:43-111
//   kappa interpolant (exact, linear) . python/models.py:290-295
//   OCP cost / bounds / C,D ........... python/mpc.py:29-101
//   weights, yref, warm start, status . new_python/controller.py:313-405
//   solver options .................... new_python/controller.py:298-309, old/generate.py:18-26
//   fdyn6 plant ....................... python/models.py:455-606
//   sim options / switch / lap ........ python/main.py:395-400, 475-517, 561-570
//   nonlinear h(x,u) + slacks ......... old/generate_acaods_interface.py:190-212, 351-449
//   DPC: kin4 ODE, RK4 step ........... dpc/main.py:56-97, constants dpc/main.py:23-37
//   DPC: IHM1.5 tracking cost ......... dpc/main.py:374-409, doc/dpc.typ:44-56
//   Cartesian pose -> Frenet (s, n, psi)  src/ihm2/src/common/tracks.cpp:183-288 (Track::project),
//                                        src/ihm2/src/common/math.cpp:19-25, src/ihm2/src/mpc_control_node.cpp:142-157
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr int NX = 8, NU = 2, NZ = 10;

// ---------------------------------------------------------------------------------------
// vehicle parameters (python/constants.py:43-111).  Index layout shared with the tests.
// ---------------------------------------------------------------------------------------
enum {
  P_M = 0, P_IZ, P_LR, P_LF, P_CM0, P_CR0, P_CR1, P_CR2, P_TT, P_TD,
  P_ZCG, P_TRACKW, P_CDOWN, P_KTV, P_BA, P_CA, P_DA, P_EA, P_G, P_CROSS,
  P_RW, P_IW, P_KD, P_KS, P_BS, P_CS, P_DS, P_ES, NP
};

struct Track {
  int n;              // table length (3 laps x 500 in the reference)
  const double* s;    // s_ref[n], strictly increasing
  const double* k;    // kappa_ref[n]
  const double* wr = nullptr;   // right_width[n], left_width[n] (may be null: no track constraint)
  const double* wl = nullptr;
};

// CasADi interpolant("linear", lookup_mode "exact"): binary search for the interval that
// contains s, clamped to [0, n-2] so that values outside the grid extrapolate linearly
// (python/models.py:290-295).  Returns kappa and the slope of the active segment.
inline void table_lookup(const Track& tr, const double* tab, double s, double& val, double& slope) {
  int lo = 0, hi = tr.n;  // upper_bound
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (tr.s[mid] <= s) lo = mid + 1; else hi = mid;
  }
  int i = lo - 1;
  if (i < 0) i = 0;
  if (i > tr.n - 2) i = tr.n - 2;
  slope = (tab[i + 1] - tab[i]) / (tr.s[i + 1] - tr.s[i]);
  val = tab[i] + slope * (s - tr.s[i]);
}
inline void kappa_lookup(const Track& tr, double s, double& kap, double& dkap) { table_lookup(tr, tr.k, s, kap, dkap); }

// Nonlinear path constraints h(x, u) of the gen-1 OCP (old/generate_acaods_interface.py:312-325 for the
// 8-state Frenet model; bounds and slacks :351-449), on the quantities of the current controller model
// (python/models.py:232-307: F_Rx, F_Fx, beta as in fkin6):
//   h0 =  n - L/2 sin|psi| + W/2 cos psi - right_width(s)      (literal signs)
//   h1 = -n + L/2 sin|psi| + W/2 cos psi - left_width(s)
//   h2 = T_dot = (u_T - T) / t_T,   h3 = delta_dot = (u_delta - delta) / t_delta
//   h4 = a_lat = (-F_Rx sin(beta) + F_Fx sin(delta - beta)) / m + (v_x^2 + v_y^2) sin(beta) / l_R
// width interpolants are "linear" CasADi interpolants on s_ref (old/generate.py:44-46).  Jh [5][10] in (x, u) order.
constexpr int NH = 5, NH_E = 2;
void con_h(const double* x, const double* u, const double* p, const Track& tr, double half_L, double half_W,
           double* h, double (*Jh)[10]) {
  const double n = x[1], psi = x[2], vx = x[3], vy = x[4], T = x[6], dl = x[7];
  const double m = p[P_M], lR = p[P_LR], lF = p[P_LF], C = lR / (lR + lF);
  double wr = 0, dwr = 0, wl = 0, dwl = 0;
  if (tr.wr) table_lookup(tr, tr.wr, x[0], wr, dwr);
  if (tr.wl) table_lookup(tr, tr.wl, x[0], wl, dwl);
  const double sp = std::sin(std::fabs(psi)), cp = std::cos(psi);
  const double sg = (psi > 0) - (psi < 0);   // d|psi|/dpsi (CasADi: sign)
  h[0] = n - half_L * sp + half_W * cp - wr;
  h[1] = -n + half_L * sp + half_W * cp - wl;
  h[2] = (u[0] - T) / p[P_TT];
  h[3] = (u[1] - dl) / p[P_TD];
  const double Fm = p[P_CM0] * T;
  const double th = std::tanh(10.0 * vx);
  const double poly = p[P_CR0] + p[P_CR1] * vx + p[P_CR2] * vx * vx;
  const double FRx = 0.5 * Fm - poly * th, FFx = 0.5 * Fm;
  const double td = std::tan(dl), beta = std::atan(C * td);
  const double sb = std::sin(beta), cb = std::cos(beta), sdb = std::sin(dl - beta), cdb = std::cos(dl - beta);
  const double G = C * (1.0 + td * td) / (1.0 + C * C * td * td);
  const double v2 = vx * vx + vy * vy;
  h[4] = (-FRx * sb + FFx * sdb) / m + v2 * sb / lR;
  if (!Jh) return;
  for (int i = 0; i < NH; i++)
    for (int j = 0; j < 10; j++) Jh[i][j] = 0.0;
  const double dabs = std::cos(std::fabs(psi)) * sg;   // d sin|psi| / d psi
  Jh[0][0] = -dwr; Jh[0][1] = 1.0;  Jh[0][2] = -half_L * dabs - half_W * std::sin(psi);
  Jh[1][0] = -dwl; Jh[1][1] = -1.0; Jh[1][2] = half_L * dabs - half_W * std::sin(psi);
  Jh[2][6] = -1.0 / p[P_TT]; Jh[2][8] = 1.0 / p[P_TT];
  Jh[3][7] = -1.0 / p[P_TD]; Jh[3][9] = 1.0 / p[P_TD];
  const double dFd = -(p[P_CR1] + 2.0 * p[P_CR2] * vx) * th - poly * 10.0 * (1.0 - th * th);
  Jh[4][3] = -dFd * sb / m + 2.0 * vx * sb / lR;
  Jh[4][4] = 2.0 * vy * sb / lR;
  Jh[4][6] = 0.5 * p[P_CM0] * (sdb - sb) / m;
  Jh[4][7] = (-FRx * cb * G + FFx * cdb * (1.0 - G)) / m + v2 * cb * G / lR;
}

// fkin6 continuous dynamics and (optionally) its Jacobian J[8][10] w.r.t. (x,u),
// columns ordered (x0..x7, u0, u1).  python/models.py:232-307.
void fkin6(const double* x, const double* u, const double* p, const Track& tr, double* f,
           double (*J)[NX + NU]) {
  const double n = x[1], psi = x[2], vx = x[3], vy = x[4], r = x[5], T = x[6], dl = x[7];
  const double m = p[P_M], lR = p[P_LR], lF = p[P_LF];
  const double C = lR / (lR + lF);  // rear_weight_distribution
  const double tT = p[P_TT], tD = p[P_TD];

  const double ddot = (u[1] - dl) / tD;
  const double Tdot = (u[0] - T) / tT;

  const double Fm = p[P_CM0] * T;
  const double th = std::tanh(10.0 * vx);
  const double poly = p[P_CR0] + p[P_CR1] * vx + p[P_CR2] * vx * vx;
  const double Fd = -poly * th;
  const double FRx = 0.5 * Fm + Fd, FFx = 0.5 * Fm;

  const double td = std::tan(dl);
  const double beta = std::atan(C * td);
  const double cb = std::cos(beta), sb = std::sin(beta);
  const double sec2 = 1.0 + td * td, den = 1.0 + C * C * td * td;
  const double G = C * sec2 / den;  // d beta / d delta
  const double bdot = G * ddot;
  const double cdb = std::cos(dl - beta), sdb = std::sin(dl - beta);
  const double vdot = (FRx * cb + FFx * cdb) / m;

  double kap, dkap;
  kappa_lookup(tr, x[0], kap, dkap);
  const double cp = std::cos(psi), sp = std::sin(psi);
  const double D = 1.0 + kap * n;
  const double num = vx * cp - vy * sp;
  const double sdot = num / D;
  const double vydot = vdot * sb + bdot * vx;

  f[0] = sdot;
  f[1] = vx * sp + vy * cp;
  f[2] = r - kap * sdot;
  f[3] = vdot * cb - bdot * vy;
  f[4] = vydot;
  f[5] = lR * vydot - bdot;
  f[6] = Tdot;
  f[7] = ddot;
  if (!J) return;

  for (int i = 0; i < NX; i++)
    for (int j = 0; j < NX + NU; j++) J[i][j] = 0.0;
  // row 0: sdot
  const double ds_ds = -num * dkap * n / (D * D);
  const double ds_dn = -num * kap / (D * D);
  const double ds_dpsi = (-vx * sp - vy * cp) / D;
  const double ds_dvx = cp / D, ds_dvy = -sp / D;
  J[0][0] = ds_ds; J[0][1] = ds_dn; J[0][2] = ds_dpsi; J[0][3] = ds_dvx; J[0][4] = ds_dvy;
  // row 1: ndot
  J[1][2] = num; J[1][3] = sp; J[1][4] = cp;
  // row 2: psidot = r - kappa * sdot
  J[2][0] = -dkap * sdot - kap * ds_ds;
  J[2][1] = -kap * ds_dn;
  J[2][2] = -kap * ds_dpsi;
  J[2][3] = -kap * ds_dvx;
  J[2][4] = -kap * ds_dvy;
  J[2][5] = 1.0;
  // derivatives of the longitudinal / lateral block
  const double dFd_dvx = -(p[P_CR1] + 2.0 * p[P_CR2] * vx) * th - poly * 10.0 * (1.0 - th * th);
  const double dG_ddl = 2.0 * C * (1.0 - C * C) * td * sec2 / (den * den);
  const double dbdot_ddl = dG_ddl * ddot - G / tD;
  const double dbdot_du = G / tD;
  const double dvdot_dvx = dFd_dvx * cb / m;
  const double dvdot_dT = 0.5 * p[P_CM0] * (cb + cdb) / m;
  const double dvdot_ddl = (-FRx * sb * G - FFx * sdb * (1.0 - G)) / m;
  // row 3: vxdot = vdot cb - bdot vy
  J[3][3] = dvdot_dvx * cb;
  J[3][4] = -bdot;
  J[3][6] = dvdot_dT * cb;
  J[3][7] = dvdot_ddl * cb - vdot * sb * G - dbdot_ddl * vy;
  J[3][9] = -dbdot_du * vy;
  // row 4: vydot = vdot sb + bdot vx
  J[4][3] = dvdot_dvx * sb + bdot;
  J[4][6] = dvdot_dT * sb;
  J[4][7] = dvdot_ddl * sb + vdot * cb * G + dbdot_ddl * vx;
  J[4][9] = dbdot_du * vx;
  // row 5: rdot = lR vydot - bdot   (literal, python/models.py:304)
  for (int j = 0; j < NX + NU; j++) J[5][j] = lR * J[4][j];
  J[5][7] -= dbdot_ddl;
  J[5][9] -= dbdot_du;
  // rows 6, 7: actuators
  J[6][6] = -1.0 / tT; J[6][8] = 1.0 / tT;
  J[7][7] = -1.0 / tD; J[7][9] = 1.0 / tD;
}

// ---------------------------------------------------------------------------------------
// fdyn6 plant (python/models.py:455-606).  The reference states it implicitly,
// F(xdot, x, u) = 0; F is affine in (a_x, a_y, rdot), so the explicit right-hand side is
// obtained from three evaluations of the force balance (exact for an affine map).
// ---------------------------------------------------------------------------------------
inline double smooth_abs_nonzero(double v) {  // python/utils.py:27-32 (min_val 1e-6)
  return std::tanh(10.0 * v) * v + 1e-6 * std::exp(-v * v);
}
inline double lat_pacejka(const double* p, double a) {  // python/models.py:69-74
  const double Ba = p[P_BA] * a;
  return p[P_DA] * std::sin(p[P_CA] * std::atan(Ba - p[P_EA] * (Ba - std::atan(Ba))));
}

// total body-frame forces / yaw moment for given accelerations (a_x, a_y)
// cart: the Cartesian dyn6 of python/models.py:310-452 (slip angles :376-379 wheel by wheel without the smooth
// absolute value, drag :399 with tanh(1000 v_x), r_kin :409 from v_x alone)
void fdyn6_forces(const double* x, const double* p, double ax, double ay, double* FxFyMz, bool cart = false) {
  const double vx = x[3], vy = x[4], r = x[5], T = x[6], dl = x[7];
  const double m = p[P_M], g = p[P_G], lR = p[P_LR], lF = p[P_LF], wb = lR + lF;
  const double tw = p[P_TRACKW];
  const double Fdown = 0.5 * p[P_CDOWN] * vx * vx;
  const double stat = 0.5 * m * g * lF / wb;
  const double lonT = 0.5 * m * ax * p[P_ZCG] / wb;
  const double latT = 0.5 * m * ay * p[P_ZCG] / tw;
  const double FzFL = -(stat - lonT + latT + 0.25 * Fdown);
  const double FzFR = -(stat - lonT - latT + 0.25 * Fdown);
  const double FzRL = -(stat + lonT + latT + 0.25 * Fdown);
  const double FzRR = -(stat + lonT - latT + 0.25 * Fdown);
  const double cd = std::cos(dl), sd = std::sin(dl);
  const double vxFL = vx - 0.5 * tw * r, vxFR = vx + 0.5 * tw * r;
  const double vyF = vy + lF * r;
  const double vlonFL = cd * vxFL + sd * vyF, vlonFR = cd * vxFR + sd * vyF;
  const double vlatFL = -sd * vxFL + cd * vyF, vlatFR = -sd * vxFR + cd * vyF;
  const double vlonRL = vx - 0.5 * tw * r, vlonRR = vx + 0.5 * tw * r;
  const double vlatR = vy - lR * r;
  const double aFL = cart ? std::atan2(vyF, vxFL) - dl : std::atan2(vlatFL, smooth_abs_nonzero(vlonFL));
  const double aFR = cart ? std::atan2(vyF, vxFR) - dl : std::atan2(vlatFR, smooth_abs_nonzero(vlonFR));
  const double aRL = cart ? std::atan2(vlatR, vlonRL) : std::atan2(vlatR, smooth_abs_nonzero(vlonRL));
  const double aRR = cart ? std::atan2(vlatR, vlonRR) : std::atan2(vlatR, smooth_abs_nonzero(vlonRR));
  const bool cross = !cart && p[P_CROSS] != 0.0;  // literal quirk python/models.py:543-546
  const double FlatFL = FzFL * lat_pacejka(p, cross ? aRR : aFL);
  const double FlatFR = FzFR * lat_pacejka(p, cross ? aRL : aFR);
  const double FlatRL = FzRL * lat_pacejka(p, cross ? aFR : aRL);
  const double FlatRR = FzRR * lat_pacejka(p, cross ? aFL : aRR);
  const double th = std::tanh((cart ? 1000.0 : 10.0) * vx);
  const double Fdrag = -(p[P_CR0] + p[P_CR1] * vx + p[P_CR2] * vx * vx) * th;
  const double C = lR / wb;
  const double beta = std::atan(C * std::tan(dl));
  const double rkin = (cart ? vx : std::hypot(vx, vy)) * std::sin(beta) / lR;
  const double dtau = p[P_KTV] * (rkin - r);
  const double dn = -m * g - 0.25 * Fdown;
  const double FlonFL = p[P_CM0] * (T - dtau) * FzFL / dn;
  const double FlonFR = p[P_CM0] * (T + dtau) * FzFR / dn;
  const double FlonRL = p[P_CM0] * (T - dtau) * FzRL / dn;
  const double FlonRR = p[P_CM0] * (T + dtau) * FzRR / dn;
  FxFyMz[0] = (FlonFR + FlonFL) * cd - (FlatFR + FlatFL) * sd + FlonRR + FlonRL + Fdrag;
  FxFyMz[1] = (FlonFR + FlonFL) * sd + (FlatFR + FlatFL) * cd + FlatRR + FlatRL;
  FxFyMz[2] = (FlonFR * cd - FlatFR * sd) * tw / 2 + (FlonFR * sd + FlatFR * cd) * lF -
              (FlonFL * cd - FlatFL * sd) * tw / 2 + (FlonFL * sd + FlatFL * cd) * lF +
              FlonRR * tw / 2 - FlatRR * lR - FlonRL * tw / 2 - FlatRL * lR;
}

void fdyn6_any(const double* x, const double* u, const double* p, const Track& tr, double* f, bool cart) {
  const double n = x[1], psi = x[2], vx = x[3], vy = x[4], r = x[5];
  const double m = p[P_M];
  double F0[3], F1[3], F2[3];
  fdyn6_forces(x, p, 0.0, 0.0, F0, cart);
  fdyn6_forces(x, p, 1.0, 0.0, F1, cart);
  fdyn6_forces(x, p, 0.0, 1.0, F2, cart);
  // m a = F0 + Gx a_x + Gy a_y  (first two rows)  ->  2x2 solve
  const double a11 = m - (F1[0] - F0[0]), a12 = -(F2[0] - F0[0]);
  const double a21 = -(F1[1] - F0[1]), a22 = m - (F2[1] - F0[1]);
  const double det = a11 * a22 - a12 * a21;
  const double ax = (F0[0] * a22 - a12 * F0[1]) / det;
  const double ay = (a11 * F0[1] - a21 * F0[0]) / det;
  const double Mz = F0[2] + (F1[2] - F0[2]) * ax + (F2[2] - F0[2]) * ay;
  const double cp = std::cos(psi), sp = std::sin(psi);
  if (cart) {  // python/models.py:420-422
    f[0] = vx * cp - vy * sp;
    f[1] = vx * sp + vy * cp;
    f[2] = r;
  } else {
    double kap, dkap;
    kappa_lookup(tr, x[0], kap, dkap);
    const double sdot = (vx * cp - vy * sp) / (1.0 + kap * n);
    f[0] = sdot;
    f[1] = vx * sp + vy * cp;
    f[2] = r - kap * sdot;
  }
  f[3] = ax + vy * r;  // a_x = vxdot - vy r
  f[4] = ay - vx * r;  // a_y = vydot + vx r
  f[5] = Mz / p[P_IZ];
  f[6] = (u[0] - x[6]) / p[P_TT];
  f[7] = (u[1] - x[7]) / p[P_TD];
}
void fdyn6(const double* x, const double* u, const double* p, const Track& tr, double* f) { fdyn6_any(x, u, p, tr, f, false); }
void dyn6c(const double* x, const double* u, const double* p, const Track& tr, double* f) { fdyn6_any(x, u, p, tr, f, true); }

typedef void (*rhs_fn)(const double*, const double*, const double*, const Track&, double*);
void fkin6_rhs(const double* x, const double* u, const double* p, const Track& tr, double* f) {
  fkin6(x, u, p, tr, f, nullptr);
}
// Cartesian kin6 (python/models.py:168-229): shares the v_x, v_y, T, delta rows with fkin6; pose kinematics; and
// r_dot = v_y_dot / l_R (:226) where fkin6 has l_R * v_y_dot - beta_dot
void kin6c(const double* x, const double* u, const double* p, const Track& tr, double* f) {
  double xx[NX];
  std::memcpy(xx, x, sizeof xx);
  xx[0] = xx[1] = 0.0;
  fkin6(xx, u, p, tr, f, nullptr);
  const double cp = std::cos(x[2]), sp = std::sin(x[2]);
  f[0] = x[3] * cp - x[4] * sp;
  f[1] = x[3] * sp + x[4] * cp;
  f[2] = x[5];
  f[5] = f[4] / p[P_LR];
}

// ---------------------------------------------------------------------------------------
// small dense linear algebra
// ---------------------------------------------------------------------------------------
// LU with partial pivoting, in place, n x n row-major.  Returns false if singular.
bool lu_factor(int n, double* A, int* piv) {
  for (int c = 0; c < n; c++) {
    int pr = c;
    double best = std::fabs(A[c * n + c]);
    for (int r = c + 1; r < n; r++)
      if (std::fabs(A[r * n + c]) > best) { best = std::fabs(A[r * n + c]); pr = r; }
    piv[c] = pr;
    if (!(best > 0.0)) return false;
    if (pr != c)
      for (int j = 0; j < n; j++) std::swap(A[c * n + j], A[pr * n + j]);
    const double inv = 1.0 / A[c * n + c];
    for (int r = c + 1; r < n; r++) {
      const double l = A[r * n + c] * inv;
      A[r * n + c] = l;
      for (int j = c + 1; j < n; j++) A[r * n + j] -= l * A[c * n + j];
    }
  }
  return true;
}
// solve for nrhs right-hand sides stored column-wise in Bm (n x nrhs row-major)
void lu_solve(int n, const double* A, const int* piv, double* Bm, int nrhs) {
  for (int c = 0; c < n; c++)
    if (piv[c] != c)
      for (int j = 0; j < nrhs; j++) std::swap(Bm[c * nrhs + j], Bm[piv[c] * nrhs + j]);
  for (int r = 1; r < n; r++)
    for (int c = 0; c < r; c++) {
      const double l = A[r * n + c];
      for (int j = 0; j < nrhs; j++) Bm[r * nrhs + j] -= l * Bm[c * nrhs + j];
    }
  for (int r = n - 1; r >= 0; r--) {
    for (int c = r + 1; c < n; c++) {
      const double a = A[r * n + c];
      for (int j = 0; j < nrhs; j++) Bm[r * nrhs + j] -= a * Bm[c * nrhs + j];
    }
    const double inv = 1.0 / A[r * n + r];
    for (int j = 0; j < nrhs; j++) Bm[r * nrhs + j] *= inv;
  }
}

// ---------------------------------------------------------------------------------------
// dyn10 plant: `fdyn10_model`, python/models.py:609-801.  x[15] = (s, n, psi, v_x, v_y, r, omega_FL,
// omega_FR, omega_RL, omega_RR, tau_FL, tau_FR, tau_RL, tau_RR, delta), u[5] = (u_tau_FL .. u_tau_RR, u_delta).
// The reference states it implicitly (the accelerations feed the vertical loads, :655-684); every tyre force is
// proportional to its wheel's vertical load, so the force balance is affine in (a_x, a_y): one 2 x 2 solve gives the
// explicit right-hand side.  Straight slip-angle assignment (:717-720), drag switched by tanh(1000 v_x) (:659).
// ---------------------------------------------------------------------------------------
constexpr int NX10 = 15, NU10 = 5;
void fdyn10(const double* x, const double* u, const double* p, const Track& tr, double* f) {
  const double n = x[1], psi = x[2], vx = x[3], vy = x[4], r = x[5], dl = x[14];
  const double* om = x + 6; const double* tau = x + 10;
  const double m = p[P_M], g = p[P_G], lR = p[P_LR], lF = p[P_LF], wb = lR + lF, tw = p[P_TRACKW];
  const double Fdown = 0.5 * p[P_CDOWN] * vx * vx;
  const double stat = 0.5 * m * g * lF / wb;
  const double cl = 0.5 * m * p[P_ZCG] / wb, ct = 0.5 * m * p[P_ZCG] / tw;
  const double sd = std::sin(dl), cd = std::cos(dl);
  auto sabs = [](double v) { return std::tanh(10.0 * v) * v + 1e-6 * std::exp(-v * v); };   // utils.py:27-32
  const double vxL = vx - 0.5 * tw * r, vxR = vx + 0.5 * tw * r, vyF = vy + lF * r, vlatR = vy - lR * r;
  // wheel order FL, FR, RL, RR
  const double vlon[4] = {cd * vxL + sd * vyF, cd * vxR + sd * vyF, vxL, vxR};
  const double vlat[4] = {-sd * vxL + cd * vyF, -sd * vxR + cd * vyF, vlatR, vlatR};
  double mu_lat[4], mu_lon[4];
  for (int w = 0; w < 4; w++) {
    const double va = sabs(vlon[w]);
    const double al = std::atan2(vlat[w], va);
    const double Ba = p[P_BA] * al;
    mu_lat[w] = p[P_DA] * std::sin(p[P_CA] * std::atan(Ba - p[P_EA] * (Ba - std::atan(Ba))));      // lat_pacejka, :69-74
    const double sl = om[w] * p[P_RW] / va - 1.0;                                                    // slip ratio, :723-726
    const double Bs = p[P_BS] * sl;
    mu_lon[w] = -p[P_DS] * std::sin(p[P_CS] * std::atan(Bs - p[P_ES] * (Bs - std::atan(Bs))));     // F_lon = -F_z lon_pacejka(s)
  }
  const double Fdrag = -(p[P_CR0] + p[P_CR1] * vx + p[P_CR2] * vx * vx) * std::tanh(1000.0 * vx);
  // body-frame force per unit vertical load of each wheel, and its yaw arm
  double fx[4], fy[4], mz[4];
  fx[0] = mu_lon[0] * cd - mu_lat[0] * sd; fy[0] = mu_lon[0] * sd + mu_lat[0] * cd;
  fx[1] = mu_lon[1] * cd - mu_lat[1] * sd; fy[1] = mu_lon[1] * sd + mu_lat[1] * cd;
  fx[2] = mu_lon[2]; fy[2] = mu_lat[2]; fx[3] = mu_lon[3]; fy[3] = mu_lat[3];
  mz[0] = -0.5 * tw * fx[0] + lF * fy[0]; mz[1] = 0.5 * tw * fx[1] + lF * fy[1];
  mz[2] = -0.5 * tw * fx[2] - lR * fy[2]; mz[3] = 0.5 * tw * fx[3] - lR * fy[3];
  // F_z,w = z0 + cx_w a_x + cy_w a_y  (:661-684)
  const double z0 = -(stat + 0.25 * Fdown);
  const double cx[4] = {cl, cl, -cl, -cl}, cy[4] = {-ct, ct, -ct, ct};
  double Fx0 = Fdrag, Fy0 = 0, gxx = 0, gxy = 0, gyx = 0, gyy = 0;
  for (int w = 0; w < 4; w++) {
    Fx0 += z0 * fx[w]; Fy0 += z0 * fy[w];
    gxx += cx[w] * fx[w]; gxy += cy[w] * fx[w]; gyx += cx[w] * fy[w]; gyy += cy[w] * fy[w];
  }
  const double a11 = m - gxx, a12 = -gxy, a21 = -gyx, a22 = m - gyy;
  const double det = a11 * a22 - a12 * a21;
  const double ax = (Fx0 * a22 - a12 * Fy0) / det, ay = (a11 * Fy0 - a21 * Fx0) / det;
  double Mz = 0, Fz[4];
  for (int w = 0; w < 4; w++) { Fz[w] = z0 + cx[w] * ax + cy[w] * ay; Mz += Fz[w] * mz[w]; }
  double kap, dkap;
  kappa_lookup(tr, x[0], kap, dkap);
  const double sp = std::sin(psi), cp = std::cos(psi);
  const double sdot = (vx * cp - vy * sp) / (1.0 + kap * n);
  f[0] = sdot; f[1] = vx * sp + vy * cp; f[2] = r - kap * sdot;
  f[3] = ax + vy * r; f[4] = ay - vx * r; f[5] = Mz / p[P_IZ];
  for (int w = 0; w < 4; w++) {
    f[6 + w] = (tau[w] - (p[P_KD] * om[w] + p[P_KS] + p[P_RW] * Fz[w] * mu_lon[w])) / p[P_IW];   // :781-784
    f[10 + w] = (u[w] - tau[w]) / p[P_TT];
  }
  f[14] = (u[4] - dl) / p[P_TD];
}

struct Tableau { int ns; const double* A; const double* b; };  // A row-major ns x ns

// Fully implicit collocation step of a plant of any dimension: Newton on all ns * nx stage derivatives with a
// central-difference Jacobian, converged to round-off (the AcadosSimSolver of python/main.py:395-400 on fdyn10).
template <class Rhs>
int irk_step_plant(int nx, Rhs rhs, const Tableau& tb, const double* x, double h, double* xn) {
  const int ns = tb.ns, n = ns * nx;
  std::vector<double> K(n), M(n * n), R(n), Js(ns * nx * nx), xs(nx), fs(nx), xp(nx), fp(nx), fm(nx);
  std::vector<int> piv(n);
  rhs(x, fs.data());
  for (int i = 0; i < ns; i++) for (int a = 0; a < nx; a++) K[i * nx + a] = fs[a];
  double rprev = std::numeric_limits<double>::infinity();
  bool ok = false;
  for (int it = 0; it < 60; it++) {
    double rmax = 0.0, kmax = 0.0;
    for (int i = 0; i < ns; i++) {
      for (int a = 0; a < nx; a++) {
        double v = x[a];
        for (int j = 0; j < ns; j++) v += h * tb.A[i * ns + j] * K[j * nx + a];
        xs[a] = v;
      }
      rhs(xs.data(), fs.data());
      for (int c = 0; c < nx; c++) {
        const double e = 1e-6 * (1.0 + std::fabs(xs[c]));
        xp = xs; xp[c] = xs[c] + e; rhs(xp.data(), fp.data());
        xp[c] = xs[c] - e; rhs(xp.data(), fm.data());
        for (int a = 0; a < nx; a++) Js[(i * nx + a) * nx + c] = (fp[a] - fm[a]) / (2 * e);
      }
      for (int a = 0; a < nx; a++) {
        R[i * nx + a] = K[i * nx + a] - fs[a];
        rmax = std::max(rmax, std::fabs(R[i * nx + a])); kmax = std::max(kmax, std::fabs(K[i * nx + a]));
      }
    }
    if (!(rmax == rmax)) return -1;
    if (rmax <= 4e-15 * (1.0 + kmax) || (it >= 2 && rmax >= 0.25 * rprev && rmax <= 1e-9 * (1.0 + kmax))) { ok = true; break; }
    rprev = rmax;
    for (int i = 0; i < ns; i++)
      for (int a = 0; a < nx; a++)
        for (int j = 0; j < ns; j++)
          for (int c = 0; c < nx; c++)
            M[(i * nx + a) * n + j * nx + c] = ((i == j && a == c) ? 1.0 : 0.0) - h * tb.A[i * ns + j] * Js[(i * nx + a) * nx + c];
    if (!lu_factor(n, M.data(), piv.data())) return -2;
    lu_solve(n, M.data(), piv.data(), R.data(), 1);
    for (int q = 0; q < n; q++) K[q] -= R[q];
  }
  if (!ok) return -3;
  for (int a = 0; a < nx; a++) {
    double v = x[a];
    for (int i = 0; i < ns; i++) v += h * tb.b[i] * K[i * nx + a];
    xn[a] = v;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------
// integrators with forward sensitivities.  Output: xn[8], S[8][10] = d xn / d (x,u).
// ---------------------------------------------------------------------------------------
// classic RK4, one step (acados ERK, 4 stages, 1 step: old/generate.py:24-26)
void erk4_sens(const double* x, const double* u, const double* p, const Track& tr, double h,
               double* xn, double (*S)[NZ]) {
  static const double a[4] = {0.0, 0.5, 0.5, 1.0}, bw[4] = {1.0 / 6, 1.0 / 3, 1.0 / 3, 1.0 / 6};
  double k[NX], kprev[NX] = {0}, xs[NX], J[NX][NZ];
  double dk[NX][NZ], dkprev[NX][NZ] = {{0}}, dxs[NX][NZ];
  double acc[NX] = {0}, dacc[NX][NZ] = {{0}};
  for (int st = 0; st < 4; st++) {
    for (int i = 0; i < NX; i++) {
      xs[i] = x[i] + h * a[st] * kprev[i];
      for (int j = 0; j < NZ; j++) dxs[i][j] = (i == j ? 1.0 : 0.0) + h * a[st] * dkprev[i][j];
    }
    fkin6(xs, u, p, tr, k, J);
    for (int i = 0; i < NX; i++)
      for (int j = 0; j < NZ; j++) {
        double v = (j >= NX) ? J[i][j] : 0.0;  // direct dependence on u
        for (int l = 0; l < NX; l++) v += J[i][l] * dxs[l][j];
        dk[i][j] = v;
      }
    for (int i = 0; i < NX; i++) {
      acc[i] += bw[st] * k[i];
      kprev[i] = k[i];
      for (int j = 0; j < NZ; j++) { dacc[i][j] += bw[st] * dk[i][j]; dkprev[i][j] = dk[i][j]; }
    }
  }
  for (int i = 0; i < NX; i++) {
    xn[i] = x[i] + h * acc[i];
    if (S)
      for (int j = 0; j < NZ; j++) S[i][j] = (i == j ? 1.0 : 0.0) + h * dacc[i][j];
  }
}


// Fully implicit RK (collocation) step, Newton on all ns*8 stage derivatives, converged
// to round-off; sensitivities by the implicit-function theorem on the collocation system
// (acados IRK: new_python/controller.py:305-307; sim: python/main.py:395-400).
// If analytic == false the Newton matrix comes from central differences of `rhs` and no
// sensitivities are produced (plant simulation only).  Returns Newton iterations, <0 on failure.
int irk_step(rhs_fn rhs, bool analytic, const Tableau& tb, const double* x, const double* u,
             const double* p, const Track& tr, double h, double* xn, double (*S)[NZ]) {
  const int ns = tb.ns, n = ns * NX;
  std::vector<double> K(n), M(n * n), R(n), Js(ns * NX * NZ);
  std::vector<int> piv(n);
  double f0[NX];
  rhs(x, u, p, tr, f0);
  for (int i = 0; i < ns; i++)
    for (int a = 0; a < NX; a++) K[i * NX + a] = f0[a];
  auto stage_state = [&](int i, double* xs) {
    for (int a = 0; a < NX; a++) {
      double v = x[a];
      for (int j = 0; j < ns; j++) v += h * tb.A[i * ns + j] * K[j * NX + a];
      xs[a] = v;
    }
  };
  auto jac = [&](const double* xs, double* Jout) {  // Jout[8][10]
    if (analytic) {
      double ftmp[NX];
      fkin6(xs, u, p, tr, ftmp, reinterpret_cast<double(*)[NZ]>(Jout));
    } else {
      double xp[NX], fp[NX], fm[NX];
      for (int c = 0; c < NX; c++) {
        const double e = 1e-6 * (1.0 + std::fabs(xs[c]));
        std::memcpy(xp, xs, sizeof xp);
        xp[c] = xs[c] + e; rhs(xp, u, p, tr, fp);
        xp[c] = xs[c] - e; rhs(xp, u, p, tr, fm);
        for (int a = 0; a < NX; a++) Jout[a * NZ + c] = (fp[a] - fm[a]) / (2 * e);
      }
      for (int a = 0; a < NX; a++) Jout[a * NZ + 8] = Jout[a * NZ + 9] = 0.0;
    }
  };
  int it = 0;
  bool ok = false;
  double rprev = std::numeric_limits<double>::infinity();
  for (; it < 60; it++) {
    double rmax = 0.0, kmax = 0.0;
    for (int i = 0; i < ns; i++) {
      double xs[NX], fs[NX];
      stage_state(i, xs);
      rhs(xs, u, p, tr, fs);
      jac(xs, &Js[i * NX * NZ]);
      for (int a = 0; a < NX; a++) {
        R[i * NX + a] = K[i * NX + a] - fs[a];
        rmax = std::max(rmax, std::fabs(R[i * NX + a]));
        kmax = std::max(kmax, std::fabs(K[i * NX + a]));
      }
    }
    if (!(rmax == rmax)) return -1;
    for (int i = 0; i < ns; i++)
      for (int a = 0; a < NX; a++)
        for (int j = 0; j < ns; j++)
          for (int c = 0; c < NX; c++)
            M[(i * NX + a) * n + j * NX + c] =
                ((i == j && a == c) ? 1.0 : 0.0) - h * tb.A[i * ns + j] * Js[(i * NX + a) * NZ + c];
    if (!lu_factor(n, M.data(), piv.data())) return -2;
    // converged to round-off: tiny residual, or stagnation at the floor set by the stiff
    // actuator rows (an error eps in T is amplified by 1/t_T in the residual)
    if (rmax <= 4e-15 * (1.0 + kmax) || (it >= 2 && rmax >= 0.25 * rprev && rmax <= 1e-9 * (1.0 + kmax))) {
      ok = true; break;
    }
    rprev = rmax;
    lu_solve(n, M.data(), piv.data(), R.data(), 1);
    for (int q = 0; q < n; q++) K[q] -= R[q];
  }
  if (!ok) return -3;
  for (int a = 0; a < NX; a++) {
    double v = x[a];
    for (int i = 0; i < ns; i++) v += h * tb.b[i] * K[i * NX + a];
    xn[a] = v;
  }
  if (S && analytic) {
    // M dK/dw = [J_x(X_i) | J_u(X_i)]  (M factored at the converged point above)
    std::vector<double> dK(n * NZ);
    for (int i = 0; i < ns; i++)
      for (int a = 0; a < NX; a++)
        for (int c = 0; c < NZ; c++) dK[(i * NX + a) * NZ + c] = Js[(i * NX + a) * NZ + c];
    lu_solve(n, M.data(), piv.data(), dK.data(), NZ);
    for (int a = 0; a < NX; a++)
      for (int c = 0; c < NZ; c++) {
        double v = (a == c) ? 1.0 : 0.0;
        for (int i = 0; i < ns; i++) v += h * tb.b[i] * dK[(i * NX + a) * NZ + c];
        S[a][c] = v;
      }
  }
  return it;
}

}  // namespace (model + integrators)

#include "../include/ihm2b.h"

namespace {

// ---------------------------------------------------------------------------------------
// OCP -> QP (python/mpc.py:29-101, new_python/controller.py:320-385) in absolute variables
// z_k = (u_k, x_k).  The LINEAR_LS cost is exactly quadratic and all path constraints are
// linear, so only the dynamics are linearised.
// ---------------------------------------------------------------------------------------
// One two-sided inequality row  lb <= c'z + e <= ub  of a stage (z = (u, x)).  soft: the row carries a lower and an
// upper slack, lb - sL <= c'z + e <= ub + sU, sL, sU >= 0, penalised by z_s (sL + sU) + Z_s/2 (sL^2 + sU^2)
// (acados idxsbx / idxsh with zl = zu, Zl = Zu: old/generate_acaods_interface.py:383-398, 413, 417, 438, 449).
struct Row { double c[NZ]; double e, lb, ub; int soft; };

struct Qp {
  int N;
  std::vector<double> A, B, c;        // [N][8][8], [N][8][2], [N][8]: x+ = A x + B u + c
  std::vector<double> H, g;           // [N+1][10][10], [N+1][10]
  std::vector<std::vector<Row>> rows; // [N+1]
  std::vector<int> nstatic;           // [N+1] rows that do not depend on the linearisation point
  std::vector<double> zs, Zs;         // [N+1] slack penalties of the stage (cost scaling applied)
  double x0[NX];
};

void build_cost(const ihm2b_config& cf, double s0, Qp& qp) {
  const int N = cf.N;
  qp.H.assign((N + 1) * NZ * NZ, 0.0);
  qp.g.assign((N + 1) * NZ, 0.0);
  for (int k = 0; k <= N; k++) {
    double* H = &qp.H[k * NZ * NZ];
    double* g = &qp.g[k * NZ];
    if (k < N) {
      // y = Vx x + Vu u : rows 0-7 x, rows 8-9 u, rows 10-11 (T - u_T, delta - u_delta)
      double V[IHM2B_NY][NZ] = {{0}};
      for (int i = 0; i < NX; i++) V[i][NU + i] = 1.0;
      for (int i = 0; i < NU; i++) V[NX + i][i] = 1.0;
      V[10][NU + 6] = 1.0; V[10][0] = -1.0;
      V[11][NU + 7] = 1.0; V[11][1] = -1.0;
      double yref[IHM2B_NY] = {0};
      yref[0] = s0 + cf.s_target * k / N;  // controller.py:374-382
      for (int r = 0; r < IHM2B_NY; r++) {
        const double w = cf.cost_scale_stage * cf.W[r];
        for (int a = 0; a < NZ; a++) {
          g[a] -= w * V[r][a] * yref[r];
          for (int b = 0; b < NZ; b++) H[a * NZ + b] += w * V[r][a] * V[r][b];
        }
      }
    } else {
      for (int i = 0; i < NX; i++) H[(NU + i) * NZ + NU + i] = cf.cost_scale_terminal * cf.W_e[i];
      g[NU + 0] = -cf.cost_scale_terminal * cf.W_e[0] * (s0 + cf.s_target);  // controller.py:383-385
      for (int i = 0; i < NU; i++) H[i * NZ + i] = 1.0;  // unused u_N, keeps H regular
    }
  }
}

void build_rows(const ihm2b_config& cf, Qp& qp) {
  const int N = cf.N;
  const bool nl = cf.nonlinear_constraints != 0;
  qp.rows.assign(N + 1, {});
  qp.nstatic.assign(N + 1, 0);
  qp.zs.assign(N + 1, 0.0); qp.Zs.assign(N + 1, 0.0);
  auto unit = [](int idx, double lb, double ub, int soft) {
    Row r; std::memset(&r, 0, sizeof r); r.c[idx] = 1.0; r.lb = lb; r.ub = ub; r.soft = soft; return r;
  };
  for (int k = 0; k <= N; k++) {
    auto& R = qp.rows[k];
    // acados scales the slack penalties of a stage like the rest of its cost (ocp_nlp_cost_ls scaling)
    qp.zs[k] = (k < N ? cf.cost_scale_stage : cf.cost_scale_terminal) * cf.slack_z;
    qp.Zs[k] = (k < N ? cf.cost_scale_stage : cf.cost_scale_terminal) * cf.slack_Z;
    if (k < N) {
      R.push_back(unit(0, -cf.T_max, cf.T_max, 0));          // idxbu (mpc.py:86-89)
      R.push_back(unit(1, -cf.delta_max, cf.delta_max, 0));
      Row g0; std::memset(&g0, 0, sizeof g0);             // C, D (mpc.py:91-99)
      g0.c[0] = 1.0; g0.c[NU + 6] = -1.0; g0.lb = -cf.g_T_max; g0.ub = cf.g_T_max;
      Row g1; std::memset(&g1, 0, sizeof g1);
      g1.c[1] = 1.0; g1.c[NU + 7] = -1.0; g1.lb = -cf.g_delta_max; g1.ub = cf.g_delta_max;
      // These two rows ARE the T_dot / delta_dot rows of h (rows 2, 3 of old/generate_acaods_interface.py:312-325:
      // (u - x) / t within +-x_dot_max, i.e. u - x within +-t x_dot_max = g_max): with the h rows on they are kept
      // once, in this form (python/mpc.py:91-99, hard, unit coefficients).  Both copies together are linearly
      // dependent active constraints, on which the interior point stalls at the end, and the h form has
      // coefficients of 1 / t_T = 1000 next to the O(1) rows.
      R.push_back(g0); R.push_back(g1);
      if (k > 0) {  // idxbx = [1,3,6,7] (mpc.py:78-80); stage 0 carries x0 instead.
        // With the gen-1 slack sets the bounds at positions idxsbx = [2, 3] of idxbx are soft
        // (old/generate_acaods_interface.py:413): here T and delta.
        R.push_back(unit(NU + 1, -cf.n_max, cf.n_max, 0));
        R.push_back(unit(NU + 3, 0.0, cf.v_x_max, 0));
        R.push_back(unit(NU + 6, -cf.T_max, cf.T_max, nl));
        R.push_back(unit(NU + 7, -cf.delta_max, cf.delta_max, nl));
      }
    } else {  // idxbx_e = [1,3,4,5] with the n / v_x / T / delta bounds (literal, mpc.py:81-83); idxsbx_e = [2, 3]
      R.push_back(unit(NU + 1, -cf.n_max, cf.n_max, 0));
      R.push_back(unit(NU + 3, -cf.v_x_max, cf.v_x_max, 0));
      R.push_back(unit(NU + 4, -cf.T_max, cf.T_max, nl));
      R.push_back(unit(NU + 5, -cf.delta_max, cf.delta_max, nl));
    }
    qp.nstatic[k] = (int)R.size();
  }
}

// rows of the linearised h(x, u) (all soft, idxsh = all: old/generate_acaods_interface.py:438, 449), appended
// after the static rows:  lh <= h(zbar) + Jh (z - zbar) <= uh
void build_h_rows(const ihm2b_config& cf, const double* p, const Track& tr, const double* x, const double* u, Qp& qp) {
  if (!cf.nonlinear_constraints) return;
  const int N = cf.N;
  const double lh[NH] = {-1e3, -1e3, -cf.T_dot_max, -cf.delta_dot_max, -cf.a_lat_max};
  const double uh[NH] = {0.0, 0.0, cf.T_dot_max, cf.delta_dot_max, cf.a_lat_max};
  for (int k = 0; k <= N; k++) {
    auto& R = qp.rows[k];
    R.resize(qp.nstatic[k]);
    double h[NH], Jh[NH][10];
    const double uz[NU] = {0.0, 0.0};
    const double* uk = (k < N) ? &u[k * NU] : uz;
    con_h(&x[k * NX], uk, p, tr, cf.track_half_L, cf.track_half_W, h, Jh);
    const int nh = (k < N) ? NH : NH_E;
    for (int i = 0; i < nh; i++) {
      if (i == 2 || i == 3) continue;   // T_dot, delta_dot: see build_rows
      Row r; std::memset(&r, 0, sizeof r);
      double e = h[i];
      for (int a = 0; a < NX; a++) { r.c[NU + a] = Jh[i][a]; e -= Jh[i][a] * x[k * NX + a]; }
      if (k < N) for (int a = 0; a < NU; a++) { r.c[a] = Jh[i][NX + a]; e -= Jh[i][NX + a] * uk[a]; }
      r.e = e; r.lb = lh[i]; r.ub = uh[i]; r.soft = 1;
      R.push_back(r);
    }
  }
}

struct QpSol {
  std::vector<double> z;     // [N+1][10]
  std::vector<double> pi;    // [N+1][8]  (pi[0] unused)
  std::vector<std::vector<double>> lamL, lamU, tL, tU;
  std::vector<std::vector<double>> sL, sU, lamSL, lamSU;   // slacks of the soft rows and the multipliers of s >= 0
  int iters, status;         // status: 0 ok, 1 max iter, 2 min step, 3 nan, 4 infeasible (stagnation exit)
  double mu, res_g, res_b, res_d, res_m, mag_g, mag_b, mag_d;
};

inline double row_val(const Row& R, const double* z) {
  double v = R.e;
  for (int a = 0; a < NZ; a++) v += R.c[a] * z[a];
  return v;
}

// Mehrotra predictor-corrector primal-dual IPM; KKT step by a Riccati recursion
// (what HPIPM's d_ocp_qp_ipm_solve does: SURVEY App. C).  Works on the residual (delta)
// form of the Newton system, x_0 is held fixed (stage 0 embeds the measured state).
// Soft rows: the slack pair (sL, sU) and the multipliers of sL, sU >= 0 are eliminated row by row from the Newton
// system (HPIPM does the same with its `Zs_inv` vectors), which only changes the barrier weight and the
// first-order term the row contributes to the Riccati recursion:
//   D = Z + lam/t + lamS/s,   gamma_eff = (lam/t)(Z + lamS/s)/D,   q_eff = q - (lam/t)(r_s + q + q_S)/D.
void ipm_solve(const ihm2b_config& cf, const Qp& qp, QpSol& s) {
  const int N = qp.N;
  const double tau = 0.995, thr0 = cf.qp_thr0 > 0 ? cf.qp_thr0 : 0.1;
  static const bool trace = std::getenv("IHM2O_TRACE") != nullptr;
  // ---- initialisation: primal from the caller (warm), slacks / multipliers cold ----
  for (int a = 0; a < NX; a++) s.z[NU + a] = qp.x0[a];
  s.pi.assign((N + 1) * NX, 0.0);
  auto fit = [&](std::vector<std::vector<double>>& v) { v.assign(N + 1, {}); for (int k = 0; k <= N; k++) v[k].assign(qp.rows[k].size(), 0.0); };
  fit(s.lamL); fit(s.lamU); fit(s.tL); fit(s.tU); fit(s.sL); fit(s.sU); fit(s.lamSL); fit(s.lamSU);
  int nineq = 0;
  for (int k = 0; k <= N; k++) {
    const auto& R = qp.rows[k];
    for (size_t r = 0; r < R.size(); r++) {
      const double cz = row_val(R[r], &s.z[k * NZ]);
      if (R[r].soft) {
        // the slacks start where they absorb the row's violation at the initial point (t = thr0 there): the soft rows
        // are then primal feasible from the first iterate on, whatever the warm start violates
        s.sL[k][r] = std::max(thr0, R[r].lb - cz + thr0);
        s.sU[k][r] = std::max(thr0, cz - R[r].ub + thr0);
        s.lamSL[k][r] = cf.qp_mu0 / s.sL[k][r];
        s.lamSU[k][r] = cf.qp_mu0 / s.sU[k][r];
        nineq += 2;
      }
      s.tL[k][r] = std::max(cz + s.sL[k][r] - R[r].lb, thr0);
      s.tU[k][r] = std::max(R[r].ub + s.sU[k][r] - cz, thr0);
      s.lamL[k][r] = cf.qp_mu0 / s.tL[k][r];
      s.lamU[k][r] = cf.qp_mu0 / s.tU[k][r];
      nineq += 2;
    }
  }
  std::vector<double> rg((N + 1) * NZ), rb(N * NX);
  std::vector<std::vector<double>> rdL, rdU, rmL, rmU, rmSL, rmSU, rsL, rsU;
  std::vector<std::vector<double>> dlamL, dlamU, dtL, dtU, dsL, dsU, dlamSL, dlamSU;
  fit(rdL); fit(rdU); fit(rmL); fit(rmU); fit(rmSL); fit(rmSU); fit(rsL); fit(rsU);
  fit(dlamL); fit(dlamU); fit(dtL); fit(dtU); fit(dsL); fit(dsU); fit(dlamSL); fit(dlamSU);
  std::vector<double> Ht((N + 1) * NZ * NZ), rgt((N + 1) * NZ);
  std::vector<double> P((N + 1) * NX * NX), pv((N + 1) * NX), Kfb(N * NU * NX), kff(N * NU);
  std::vector<double> Linv(N * 4), dz((N + 1) * NZ), dpi((N + 1) * NX);
  s.status = 1;
  int it = 0;
  const double mu_floor = 0.1 * cf.qp_tol_comp;
  int at_floor = 0;
  double prev_res_g = std::numeric_limits<double>::infinity();
  double mu_hist[8] = {0};
  int short_steps = 0;
  for (;; it++) {
    // ---- residuals ----
    double mg = 0, mb = 0, res_g = 0, res_b = 0, res_d = 0, res_m = 0, mu = 0;
    for (int k = 0; k <= N; k++) {
      const double* z = &s.z[k * NZ];
      for (int a = 0; a < NZ; a++) {
        double v = qp.g[k * NZ + a];
        mg = std::max(mg, std::fabs(v));
        for (int b = 0; b < NZ; b++) v += qp.H[(k * NZ + a) * NZ + b] * z[b];
        rg[k * NZ + a] = v;
      }
      if (k < N) {
        const double* A = &qp.A[k * NX * NX]; const double* B = &qp.B[k * NX * NU];
        const double* pin = &s.pi[(k + 1) * NX];
        for (int a = 0; a < NU; a++)
          for (int i = 0; i < NX; i++) rg[k * NZ + a] += B[i * NU + a] * pin[i];
        for (int a = 0; a < NX; a++)
          for (int i = 0; i < NX; i++) rg[k * NZ + NU + a] += A[i * NX + a] * pin[i];
        for (int i = 0; i < NX; i++) {
          double v = qp.c[k * NX + i];
          mb = std::max(mb, std::fabs(v));
          for (int a = 0; a < NX; a++) v += A[i * NX + a] * z[NU + a];
          for (int a = 0; a < NU; a++) v += B[i * NU + a] * z[a];
          rb[k * NX + i] = v - s.z[(k + 1) * NZ + NU + i];
          res_b = std::max(res_b, std::fabs(rb[k * NX + i]));
        }
      }
      for (int a = 0; a < NX; a++) rg[k * NZ + NU + a] -= s.pi[k * NX + a];
      const auto& R = qp.rows[k];
      for (size_t r = 0; r < R.size(); r++) {
        const double cz = row_val(R[r], z);
        for (int a = 0; a < NZ; a++) rg[k * NZ + a] -= R[r].c[a] * (s.lamL[k][r] - s.lamU[k][r]);
        rdL[k][r] = cz + s.sL[k][r] - R[r].lb - s.tL[k][r];
        rdU[k][r] = R[r].ub + s.sU[k][r] - cz - s.tU[k][r];
        // slack residuals relative to the size of the row's bounds (the rows span 0.02 ... 1e6)
        const double isc = 1.0 / (1.0 + std::min(std::fabs(R[r].lb), std::fabs(R[r].ub)));
        res_d = std::max(res_d, std::max(std::fabs(rdL[k][r]), std::fabs(rdU[k][r])) * isc);
        rmL[k][r] = s.lamL[k][r] * s.tL[k][r];
        rmU[k][r] = s.lamU[k][r] * s.tU[k][r];
        res_m = std::max(res_m, std::max(rmL[k][r], rmU[k][r]));

        mu += rmL[k][r] + rmU[k][r];
        if (R[r].soft) {
          rmSL[k][r] = s.lamSL[k][r] * s.sL[k][r];
          rmSU[k][r] = s.lamSU[k][r] * s.sU[k][r];
          res_m = std::max(res_m, std::max(rmSL[k][r], rmSU[k][r]));
          mu += rmSL[k][r] + rmSU[k][r];
          rsL[k][r] = qp.Zs[k] * s.sL[k][r] + qp.zs[k] - s.lamL[k][r] - s.lamSL[k][r];
          rsU[k][r] = qp.Zs[k] * s.sU[k][r] + qp.zs[k] - s.lamU[k][r] - s.lamSU[k][r];
          res_g = std::max(res_g, std::max(std::fabs(rsL[k][r]), std::fabs(rsU[k][r])));
        }
      }
      const int a0 = (k == 0) ? NU : NZ;  // x_0 is not a variable
      for (int a = 0; a < (k == 0 ? NU : NZ); a++) res_g = std::max(res_g, std::fabs(rg[k * NZ + a]));
      if (k == 0) for (int a = a0; a < NZ; a++) rg[a] = 0.0;
    }
    mu /= nineq;
    s.mu = mu; s.res_g = res_g; s.res_b = res_b; s.res_d = res_d; s.res_m = res_m; s.iters = it;
    s.mag_g = mg; s.mag_b = mb; s.mag_d = 0.0;
    if (trace) std::fprintf(stderr, "it %2d res_g %.2e res_b %.2e res_d %.2e res_m %.2e mu %.2e\n", it, res_g, res_b, res_d, res_m, mu);
    if (!(res_g == res_g) || !(res_b == res_b) || !(mu == mu)) { s.status = 3; return; }
    const double tol = cf.qp_tol;
    const double tol_comp = cf.qp_tol_comp;
    if (res_g <= tol * (1 + mg) && res_b <= tol * (1 + mb) && res_d <= tol && res_m <= tol_comp) {
      s.status = 0; return;
    }
    // converged as far as the arithmetic allows: feasibility and complementarity at the tight tolerances, the
    // barrier parameter at its floor, stationarity within 1e-6 relative and no longer improving (thin interiors
    // leave gamma = lam / t ~ 1e9+ and the Riccati step loses the last digits of the stationarity residual)
    const bool loose = res_g <= 1e-6 * (1 + mg) && res_b <= tol * (1 + mb) && res_d <= tol && res_m <= tol_comp;
    if (loose && mu <= 2.0 * mu_floor && res_g >= 0.1 * prev_res_g) { s.status = 0; return; }
    prev_res_g = res_g;
    // numerical breakdown: four iterations with the barrier parameter at its floor and still no acceptable point (the
    // classical Riccati recursion cannot resolve the step once lam / t ~ 1e15 sits next to an O(1) Hessian: thin
    // interiors, < 1 % of the config-2 batch).  Reported as an iteration-limit stop.
    if (mu <= 2.0 * mu_floor && ++at_floor >= 4) { s.status = 1; return; }
    if (it >= cf.qp_max_iter) { s.status = 1; return; }
    if (cf.qp_infeas_exit) {
      // primal infeasible QP: the step lengths collapse against rows that cannot be met and the multipliers of
      // those rows grow, so the barrier parameter stops decreasing.  Eight iterations in a row with a step below
      // 0.5 and no decrease of the barrier parameter over them end the solve (a feasible problem loses a factor 2+
      // per iteration once it is centred).
      if (it >= 8 && short_steps >= 8 && mu >= mu_hist[it & 7]) { s.status = 4; return; }
      mu_hist[it & 7] = mu;
    }

    auto kkt_solve = [&]() {
      // modified Hessian / gradient
      for (int k = 0; k <= N; k++) {
        for (int q = 0; q < NZ * NZ; q++) Ht[k * NZ * NZ + q] = qp.H[k * NZ * NZ + q];
        for (int a = 0; a < NZ; a++) rgt[k * NZ + a] = rg[k * NZ + a];
        const auto& R = qp.rows[k];
        for (size_t r = 0; r < R.size(); r++) {
          double gL = s.lamL[k][r] / s.tL[k][r], gU = s.lamU[k][r] / s.tU[k][r];
          double qL = (rmL[k][r] + s.lamL[k][r] * rdL[k][r]) / s.tL[k][r];
          double qU = (rmU[k][r] + s.lamU[k][r] * rdU[k][r]) / s.tU[k][r];
          if (R[r].soft) {
            const double gSL = s.lamSL[k][r] / s.sL[k][r], gSU = s.lamSU[k][r] / s.sU[k][r];
            const double DL = qp.Zs[k] + gL + gSL, DU = qp.Zs[k] + gU + gSU;
            qL -= gL * (rsL[k][r] + qL + rmSL[k][r] / s.sL[k][r]) / DL;
            qU -= gU * (rsU[k][r] + qU + rmSU[k][r] / s.sU[k][r]) / DU;
            gL *= (qp.Zs[k] + gSL) / DL;
            gU *= (qp.Zs[k] + gSU) / DU;
          }
          const double gam = gL + gU, q = qL - qU;
          for (int a = 0; a < NZ; a++) {
            rgt[k * NZ + a] += R[r].c[a] * q;
            for (int b = 0; b < NZ; b++) Ht[(k * NZ + a) * NZ + b] += gam * R[r].c[a] * R[r].c[b];
          }
        }
      }
      // backward Riccati
      for (int a = 0; a < NX; a++) {
        pv[N * NX + a] = rgt[N * NZ + NU + a];
        for (int b = 0; b < NX; b++) P[(N * NX + a) * NX + b] = Ht[(N * NZ + NU + a) * NZ + NU + b];
      }
      for (int k = N - 1; k >= 0; k--) {
        const double* A = &qp.A[k * NX * NX]; const double* B = &qp.B[k * NX * NU];
        const double* Pn = &P[(k + 1) * NX * NX];
        double BA[NX][NZ], W[NX][NZ], v[NX], M[NZ][NZ], mv[NZ];
        for (int i = 0; i < NX; i++) {
          for (int a = 0; a < NU; a++) BA[i][a] = B[i * NU + a];
          for (int a = 0; a < NX; a++) BA[i][NU + a] = A[i * NX + a];
        }
        for (int i = 0; i < NX; i++) {
          double acc = pv[(k + 1) * NX + i];
          for (int j = 0; j < NX; j++) acc += Pn[i * NX + j] * rb[k * NX + j];
          v[i] = acc;
          for (int a = 0; a < NZ; a++) {
            double w = 0;
            for (int j = 0; j < NX; j++) w += Pn[i * NX + j] * BA[j][a];
            W[i][a] = w;
          }
        }
        for (int a = 0; a < NZ; a++) {
          double acc = rgt[k * NZ + a];
          for (int i = 0; i < NX; i++) acc += BA[i][a] * v[i];
          mv[a] = acc;
          for (int b = 0; b < NZ; b++) {
            double mm = Ht[(k * NZ + a) * NZ + b];
            for (int i = 0; i < NX; i++) mm += BA[i][a] * W[i][b];
            M[a][b] = mm;
          }
        }
        // 2x2 inverse of M_uu
        const double det = M[0][0] * M[1][1] - M[0][1] * M[1][0];
        const double i00 = M[1][1] / det, i01 = -M[0][1] / det, i10 = -M[1][0] / det, i11 = M[0][0] / det;
        Linv[k * 4 + 0] = i00; Linv[k * 4 + 1] = i01; Linv[k * 4 + 2] = i10; Linv[k * 4 + 3] = i11;
        for (int b = 0; b < NX; b++) {
          Kfb[(k * NU + 0) * NX + b] = -(i00 * M[0][NU + b] + i01 * M[1][NU + b]);
          Kfb[(k * NU + 1) * NX + b] = -(i10 * M[0][NU + b] + i11 * M[1][NU + b]);
        }
        kff[k * NU + 0] = -(i00 * mv[0] + i01 * mv[1]);
        kff[k * NU + 1] = -(i10 * mv[0] + i11 * mv[1]);
        for (int a = 0; a < NX; a++) {
          pv[k * NX + a] = mv[NU + a] + M[NU + a][0] * kff[k * NU] + M[NU + a][1] * kff[k * NU + 1];
          for (int b = 0; b < NX; b++)
            P[(k * NX + a) * NX + b] = M[NU + a][NU + b] + M[NU + a][0] * Kfb[(k * NU + 0) * NX + b] +
                                       M[NU + a][1] * Kfb[(k * NU + 1) * NX + b];
        }
        for (int a = 0; a < NX; a++)
          for (int b = a + 1; b < NX; b++) {
            const double sym = 0.5 * (P[(k * NX + a) * NX + b] + P[(k * NX + b) * NX + a]);
            P[(k * NX + a) * NX + b] = P[(k * NX + b) * NX + a] = sym;
          }
      }
      // forward
      for (int a = 0; a < NX; a++) dz[NU + a] = 0.0;  // x_0 fixed
      for (int k = 0; k < N; k++) {
        const double* A = &qp.A[k * NX * NX]; const double* B = &qp.B[k * NX * NU];
        for (int a = 0; a < NU; a++) {
          double acc = kff[k * NU + a];
          for (int b = 0; b < NX; b++) acc += Kfb[(k * NU + a) * NX + b] * dz[k * NZ + NU + b];
          dz[k * NZ + a] = acc;
        }
        for (int i = 0; i < NX; i++) {
          double acc = rb[k * NX + i];
          for (int a = 0; a < NX; a++) acc += A[i * NX + a] * dz[k * NZ + NU + a];
          for (int a = 0; a < NU; a++) acc += B[i * NU + a] * dz[k * NZ + a];
          dz[(k + 1) * NZ + NU + i] = acc;
        }
        for (int i = 0; i < NX; i++) {
          double acc = pv[(k + 1) * NX + i];
          for (int j = 0; j < NX; j++) acc += P[((k + 1) * NX + i) * NX + j] * dz[(k + 1) * NZ + NU + j];
          dpi[(k + 1) * NX + i] = acc;  // costate of the delta problem = multiplier step
        }
      }
      dz[N * NZ] = dz[N * NZ + 1] = 0.0;
      // recover slack / multiplier steps
      for (int k = 0; k <= N; k++) {
        const auto& R = qp.rows[k];
        for (size_t r = 0; r < R.size(); r++) {
          double cdz = 0;
          for (int a = 0; a < NZ; a++) cdz += R[r].c[a] * dz[k * NZ + a];
          if (R[r].soft) {
            const double gL = s.lamL[k][r] / s.tL[k][r], gU = s.lamU[k][r] / s.tU[k][r];
            const double qL = (rmL[k][r] + s.lamL[k][r] * rdL[k][r]) / s.tL[k][r];
            const double qU = (rmU[k][r] + s.lamU[k][r] * rdU[k][r]) / s.tU[k][r];
            const double DL = qp.Zs[k] + gL + s.lamSL[k][r] / s.sL[k][r], DU = qp.Zs[k] + gU + s.lamSU[k][r] / s.sU[k][r];
            dsL[k][r] = -(rsL[k][r] + qL + rmSL[k][r] / s.sL[k][r] + gL * cdz) / DL;
            dsU[k][r] = -(rsU[k][r] + qU + rmSU[k][r] / s.sU[k][r] - gU * cdz) / DU;
            dlamSL[k][r] = -(rmSL[k][r] + s.lamSL[k][r] * dsL[k][r]) / s.sL[k][r];
            dlamSU[k][r] = -(rmSU[k][r] + s.lamSU[k][r] * dsU[k][r]) / s.sU[k][r];
          }
          dtL[k][r] = cdz + dsL[k][r] + rdL[k][r];
          dtU[k][r] = -cdz + dsU[k][r] + rdU[k][r];
          dlamL[k][r] = -(rmL[k][r] + s.lamL[k][r] * dtL[k][r]) / s.tL[k][r];
          dlamU[k][r] = -(rmU[k][r] + s.lamU[k][r] * dtU[k][r]) / s.tU[k][r];
        }
      }
    };
    // largest step in [0, 1] that keeps the slacks (primal) / the multipliers (dual) non-negative
    auto max_step = [&](double& al_p, double& al_d) {
      al_p = al_d = 1.0;
      for (int k = 0; k <= N; k++)
        for (size_t r = 0; r < qp.rows[k].size(); r++) {
          if (dlamL[k][r] < 0) al_d = std::min(al_d, -s.lamL[k][r] / dlamL[k][r]);
          if (dlamU[k][r] < 0) al_d = std::min(al_d, -s.lamU[k][r] / dlamU[k][r]);
          if (dtL[k][r] < 0) al_p = std::min(al_p, -s.tL[k][r] / dtL[k][r]);
          if (dtU[k][r] < 0) al_p = std::min(al_p, -s.tU[k][r] / dtU[k][r]);
          if (qp.rows[k][r].soft) {
            if (dlamSL[k][r] < 0) al_d = std::min(al_d, -s.lamSL[k][r] / dlamSL[k][r]);
            if (dlamSU[k][r] < 0) al_d = std::min(al_d, -s.lamSU[k][r] / dlamSU[k][r]);
            if (dsL[k][r] < 0) al_p = std::min(al_p, -s.sL[k][r] / dsL[k][r]);
            if (dsU[k][r] < 0) al_p = std::min(al_p, -s.sU[k][r] / dsU[k][r]);
          }
        }
      if (!cf.qp_split_step) al_p = al_d = std::min(al_p, al_d);
    };
    // ---- predictor ----
    kkt_solve();
    double ap_aff, ad_aff;
    max_step(ap_aff, ad_aff);
    double mu_aff = 0;
    for (int k = 0; k <= N; k++)
      for (size_t r = 0; r < qp.rows[k].size(); r++) {
        mu_aff += (s.lamL[k][r] + ad_aff * dlamL[k][r]) * (s.tL[k][r] + ap_aff * dtL[k][r]);
        mu_aff += (s.lamU[k][r] + ad_aff * dlamU[k][r]) * (s.tU[k][r] + ap_aff * dtU[k][r]);
        if (qp.rows[k][r].soft) {
          mu_aff += (s.lamSL[k][r] + ad_aff * dlamSL[k][r]) * (s.sL[k][r] + ap_aff * dsL[k][r]);
          mu_aff += (s.lamSU[k][r] + ad_aff * dlamSU[k][r]) * (s.sU[k][r] + ap_aff * dsU[k][r]);
        }
      }
    mu_aff /= nineq;
    const double sig = std::pow(mu_aff / mu, 3.0);
    // centring target; floored so the iterates stay on the central path at mu ~ 0.1 tol_comp
    // instead of diving to round-off, which wrecks the conditioning of the Riccati recursion
    const double mu_t = std::max(sig * mu, mu_floor);
    // ---- corrector ----
    for (int k = 0; k <= N; k++)
      for (size_t r = 0; r < qp.rows[k].size(); r++) {
        rmL[k][r] += dlamL[k][r] * dtL[k][r] - mu_t;
        rmU[k][r] += dlamU[k][r] * dtU[k][r] - mu_t;
        if (qp.rows[k][r].soft) {
          rmSL[k][r] += dlamSL[k][r] * dsL[k][r] - mu_t;
          rmSU[k][r] += dlamSU[k][r] * dsU[k][r] - mu_t;
        }
      }
    kkt_solve();
    double ap, ad;
    max_step(ap, ad);
    // fraction to the boundary: 0.995 far from the solution, -> 1 with the barrier parameter (Mehrotra), which
    // restores the fast local convergence of the last iterations
    const double tau_k = std::max(tau, 1.0 - mu_t);
    ap = std::min(1.0, tau_k * ap); ad = std::min(1.0, tau_k * ad);
    if (trace) std::fprintf(stderr, "      al_aff %.3f/%.3f sig %.2e mu_t %.2e  al %.3f/%.3f\n", ap_aff, ad_aff, sig, mu_t, ap, ad);
    if (std::min(ap, ad) < 1e-12) { s.status = 2; return; }
    short_steps = (std::min(ap, ad) < 0.5) ? short_steps + 1 : 0;
    for (int k = 0; k <= N; k++) {
      for (int a = 0; a < NZ; a++) s.z[k * NZ + a] += ap * dz[k * NZ + a];
      if (k > 0) for (int a = 0; a < NX; a++) s.pi[k * NX + a] += ad * dpi[k * NX + a];
      for (size_t r = 0; r < qp.rows[k].size(); r++) {
        s.lamL[k][r] += ad * dlamL[k][r]; s.lamU[k][r] += ad * dlamU[k][r];
        s.tL[k][r] += ap * dtL[k][r]; s.tU[k][r] += ap * dtU[k][r];
        if (qp.rows[k][r].soft) {
          s.sL[k][r] += ap * dsL[k][r]; s.sU[k][r] += ap * dsU[k][r];
          s.lamSL[k][r] += ad * dlamSL[k][r]; s.lamSU[k][r] += ad * dlamSU[k][r];
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// SQP / SQP-RTI driver for one problem (acados ocp_nlp_sqp / ocp_nlp_sqp_rti restated:
// SURVEY App. C; call site new_python/controller.py:396).
// ---------------------------------------------------------------------------------------
struct Problem {
  const ihm2b_config* cf;
  Track tr;
  const double* p;
};

int integrate_ctrl(const Problem& pb, const double* x, const double* u, double* xn, double (*S)[NZ]) {
  const ihm2b_config& cf = *pb.cf;
  if (cf.integrator == IHM2B_ERK4) { erk4_sens(x, u, pb.p, pb.tr, cf.dt, xn, S); return 0; }
  Tableau tb{4, cf.irk_A, cf.irk_b};
  return irk_step(fkin6_rhs, true, tb, x, u, pb.p, pb.tr, cf.dt, xn, S);
}

// linearise the dynamics along (x, u); fills qp.A, qp.B, qp.c and phi. Returns false on NaN/failure.
bool linearize(const Problem& pb, const double* x, const double* u, Qp& qp, double* phi) {
  const int N = pb.cf->N;
  qp.A.resize(N * NX * NX); qp.B.resize(N * NX * NU); qp.c.resize(N * NX);
  bool ok = true;
  for (int k = 0; k < N; k++) {
    double S[NX][NZ], xn[NX];
    if (integrate_ctrl(pb, &x[k * NX], &u[k * NU], xn, S) < 0) ok = false;
    for (int i = 0; i < NX; i++) {
      double c = xn[i];
      for (int a = 0; a < NX; a++) { qp.A[(k * NX + i) * NX + a] = S[i][a]; c -= S[i][a] * x[k * NX + a]; }
      for (int a = 0; a < NU; a++) { qp.B[(k * NX + i) * NU + a] = S[i][NX + a]; c -= S[i][NX + a] * u[k * NU + a]; }
      qp.c[k * NX + i] = c;
      phi[k * NX + i] = xn[i];
      if (!(xn[i] == xn[i])) ok = false;
    }
  }
  if (!qp.rows.empty()) build_h_rows(*pb.cf, pb.p, pb.tr, x, u, qp);
  return ok;
}

struct SolveOut { int status, qp_iters, sqp_iters; double mu, res_stat, res_eq, res_ineq, res_comp, cost; };

// Value of a soft row's slack penalty with the slack eliminated: the cheapest feasible slack of a row at value v is
// its violation (the penalty is increasing in s >= 0), so the NLP in (x, u, s) equals the NLP in (x, u) with
// cost + sum z_s viol + Z_s/2 viol^2.  The SQP-level quantities (cost, merit, KKT residuals) use this form.
inline double soft_penalty(double zs, double Zs, double viol) { return viol * (zs + 0.5 * Zs * viol); }

double ocp_cost(const ihm2b_config& cf, const Qp& qp, const double* z) {
  double c = 0;
  for (int k = 0; k <= cf.N; k++)
    for (int a = 0; a < NZ; a++) {
      double hz = 0;
      for (int b = 0; b < NZ; b++) hz += qp.H[(k * NZ + a) * NZ + b] * z[k * NZ + b];
      c += z[k * NZ + a] * (0.5 * hz + qp.g[k * NZ + a]);
    }
  // soft rows (the h rows are evaluated through their linearisation: exact at the linearisation point)
  for (int k = 0; k <= cf.N; k++)
    for (const Row& R : qp.rows[k])
      if (R.soft) {
        const double v = row_val(R, &z[k * NZ]);
        c += soft_penalty(qp.zs[k], qp.Zs[k], std::max(0.0, R.lb - v) + std::max(0.0, v - R.ub));
      }
  return c;
}

// L1 merit function of the SQP line search (acados MERIT_BACKTRACKING; configured by
// new_python/controller.py:308): cost + weighted 1-norm of the equality defects (x0 embedding, dynamics) +
// weighted violations of the inequality rows.  Weights follow the multipliers of the last QP:
// w = |mult| at the first SQP iteration, max(|mult|, (w + |mult|)/2) afterwards (Leineweber 1999).
struct MeritWeights {
  std::vector<double> eq;                    // [(N+1)*NX]: [0] = x0 embedding, [k+1] = dynamics of interval k
  std::vector<std::vector<double>> lo, up;   // per stage, per row
};

void merit_update_weights(const Qp& qp, const QpSol& sol, bool first, MeritWeights& mw) {
  auto upd = [first](double& w, double mult) {
    const double m = std::fabs(mult);
    w = first ? m : std::max(m, 0.5 * (w + m));
  };
  const int N = qp.N;
  if (first) {
    mw.eq.assign((N + 1) * NX, 0.0);
    mw.lo.assign(N + 1, {}); mw.up.assign(N + 1, {});
    for (int k = 0; k <= N; k++) { mw.lo[k].assign(qp.rows[k].size(), 0.0); mw.up[k].assign(qp.rows[k].size(), 0.0); }
  }
  for (int i = 0; i < (N + 1) * NX; i++) upd(mw.eq[i], sol.pi[i]);
  for (int k = 0; k <= N; k++)
    for (size_t r = 0; r < qp.rows[k].size(); r++) { upd(mw.lo[k][r], sol.lamL[k][r]); upd(mw.up[k][r], sol.lamU[k][r]); }
}

double merit_value(const ihm2b_config& cf, const Qp& qp, const MeritWeights& mw, const double* x0, const double* x,
                   const double* u, const double* phi) {
  const int N = cf.N;
  std::vector<double> z((N + 1) * NZ, 0.0);
  for (int k = 0; k <= N; k++) {
    if (k < N) { z[k * NZ] = u[k * NU]; z[k * NZ + 1] = u[k * NU + 1]; }
    for (int a = 0; a < NX; a++) z[k * NZ + NU + a] = x[k * NX + a];
  }
  double m = ocp_cost(cf, qp, z.data());
  for (int a = 0; a < NX; a++) m += mw.eq[a] * std::fabs(x[a] - x0[a]);
  for (int k = 0; k < N; k++)
    for (int i = 0; i < NX; i++) m += mw.eq[(k + 1) * NX + i] * std::fabs(phi[k * NX + i] - x[(k + 1) * NX + i]);
  for (int k = 0; k <= N; k++)
    for (size_t r = 0; r < qp.rows[k].size(); r++) {
      const Row& R = qp.rows[k][r];
      if (R.soft) continue;   // the exact penalty of a soft row is part of the cost
      const double cz = row_val(R, &z[k * NZ]);
      m += mw.lo[k][r] * std::max(0.0, R.lb - cz) + mw.up[k][r] * std::max(0.0, cz - R.ub);
    }
  return m;
}

void solve_one(const Problem& pb, const double* x0, double* x, double* u, SolveOut& out, QpSol* keep,
               Qp* keep_qp) {
  const ihm2b_config& cf = *pb.cf;
  const int N = cf.N;
  Qp qp; qp.N = N;
  build_cost(cf, x0[0], qp);
  build_rows(cf, qp);
  std::memcpy(qp.x0, x0, sizeof qp.x0);
  QpSol sol; sol.z.resize((N + 1) * NZ);
  std::vector<double> phi(N * NX);
  const int iters = (cf.nlp_solver == IHM2B_SQP_RTI) ? 1 : cf.nlp_max_iter;
  out = SolveOut{IHM2B_MAXITER, 0, 0, 0, 0, 0, 0, 0, 0};
  bool have_mult = false, lin_valid = false;
  const bool line_search = cf.nlp_solver == IHM2B_SQP && cf.globalization == IHM2B_MERIT_BACKTRACKING;
  MeritWeights mw;
  for (int it = 0; it < iters; it++) {
    // with the line search the accepted trial point has just been linearised
    if (!lin_valid && !linearize(pb, x, u, qp, phi.data())) { out.status = IHM2B_FAILURE; break; }
    if (cf.nlp_solver == IHM2B_SQP) {
      // NLP KKT residuals at the current iterate with the last QP multipliers
      double rs = 0, re = 0, ri = 0, rc = 0;
      for (int a = 0; a < NX; a++) re = std::max(re, std::fabs(x[a] - x0[a]));
      for (int k = 0; k <= N; k++) {
        double z[NZ] = {0};
        if (k < N) { z[0] = u[k * NU]; z[1] = u[k * NU + 1]; }
        for (int a = 0; a < NX; a++) z[NU + a] = x[k * NX + a];
        if (k < N)
          for (int i = 0; i < NX; i++) re = std::max(re, std::fabs(phi[k * NX + i] - x[(k + 1) * NX + i]));
        double rg[NZ];
        for (int a = 0; a < NZ; a++) {
          double v = qp.g[k * NZ + a];
          for (int b = 0; b < NZ; b++) v += qp.H[(k * NZ + a) * NZ + b] * z[b];
          rg[a] = v;
        }
        if (have_mult) {
          if (k < N)
            for (int i = 0; i < NX; i++) {
              for (int a = 0; a < NU; a++) rg[a] += qp.B[(k * NX + i) * NU + a] * sol.pi[(k + 1) * NX + i];
              for (int a = 0; a < NX; a++) rg[NU + a] += qp.A[(k * NX + i) * NX + a] * sol.pi[(k + 1) * NX + i];
            }
          for (int a = 0; a < NX; a++) rg[NU + a] -= sol.pi[k * NX + a];
        }
        for (size_t r = 0; r < qp.rows[k].size(); r++) {
          const Row& R = qp.rows[k][r];
          const double cz = row_val(R, z);
          // soft row: slack eliminated, s* = violation of the side (see soft_penalty)
          const double sL = R.soft ? std::max(0.0, R.lb - cz) : 0.0, sU = R.soft ? std::max(0.0, cz - R.ub) : 0.0;
          ri = std::max(ri, std::max(R.lb - sL - cz, cz - R.ub - sU));
          if (have_mult) {
            for (int a = 0; a < NZ; a++) rg[a] -= R.c[a] * (sol.lamL[k][r] - sol.lamU[k][r]);
            rc = std::max(rc, std::max(std::fabs(sol.lamL[k][r] * (cz + sL - R.lb)), std::fabs(sol.lamU[k][r] * (R.ub + sU - cz))));
            if (R.soft) {
              rc = std::max(rc, std::max(std::fabs(sol.lamSL[k][r] * sL), std::fabs(sol.lamSU[k][r] * sU)));
              rs = std::max(rs, std::fabs(qp.Zs[k] * sL + qp.zs[k] - sol.lamL[k][r] - sol.lamSL[k][r]));
              rs = std::max(rs, std::fabs(qp.Zs[k] * sU + qp.zs[k] - sol.lamU[k][r] - sol.lamSU[k][r]));
            }
          }
        }
        for (int a = (k == 0 ? 0 : 0); a < (k == 0 ? NU : NZ); a++) rs = std::max(rs, std::fabs(rg[a]));
        if (k == N) { /* u_N dummy */ }
      }
      ri = std::max(ri, 0.0);
      if (!have_mult) rs = rc = std::numeric_limits<double>::infinity();
      out.res_stat = rs; out.res_eq = re; out.res_ineq = ri; out.res_comp = rc;
      if (have_mult && rs <= cf.nlp_tol && re <= cf.nlp_tol && ri <= cf.nlp_tol && rc <= cf.nlp_tol) {
        out.status = IHM2B_SUCCESS; break;
      }
    }
    for (int k = 0; k <= N; k++) {
      sol.z[k * NZ] = (k < N) ? u[k * NU] : 0.0;
      sol.z[k * NZ + 1] = (k < N) ? u[k * NU + 1] : 0.0;
      for (int a = 0; a < NX; a++) sol.z[k * NZ + NU + a] = x[k * NX + a];
    }
    ipm_solve(cf, qp, sol);
    out.qp_iters += sol.iters; out.sqp_iters = it + 1; out.mu = sol.mu;
    if (sol.status == 3) { out.status = IHM2B_FAILURE; break; }
    if (sol.status != 0) {
      // accept an IPM that stalled above the tight tolerance but is converged to 1e-6
      const bool loose = sol.res_g <= 1e-6 * (1 + sol.mag_g) && sol.res_b <= 1e-6 * (1 + sol.mag_b) &&
                         sol.res_d <= 1e-6 * (1 + sol.mag_d) && sol.res_m <= 1e-6;
      // acados' SQP / SQP_RTI carry on with the iterate of a QP solver that stopped at its iteration limit; the
      // early exit for infeasible QPs (status 4, not in HPIPM) stands in for that limit
      if (!loose && !(cf.qp_tolerate_maxiter && (sol.status == 1 || sol.status == 4))) { out.status = IHM2B_QP_FAILURE; break; }
    }
    have_mult = true;
    if (line_search) {
      // backtracking on the merit function: alpha <- 0.7 alpha until the merit decreases, not below alpha_min = 0.05
      // (plain decrease: acados' sufficient-descent test is off by default); multipliers take the full step
      const double alpha_min = 0.05, reduction = 0.7;
      // multiplier of the embedding x_0 = x0 (ipm_solve eliminates x_0): stationarity in x_0 at the QP solution
      for (int a = 0; a < NX; a++) {
        double v = qp.g[NU + a];
        for (int b = 0; b < NZ; b++) v += qp.H[(NU + a) * NZ + b] * sol.z[b];
        for (int i = 0; i < NX; i++) v += qp.A[i * NX + a] * sol.pi[NX + i];
        for (size_t r = 0; r < qp.rows[0].size(); r++) v -= qp.rows[0][r].c[NU + a] * (sol.lamL[0][r] - sol.lamU[0][r]);
        sol.pi[a] = v;
      }
      merit_update_weights(qp, sol, it == 0, mw);
      const double m0 = merit_value(cf, qp, mw, x0, x, u, phi.data());
      std::vector<double> xo(x, x + (N + 1) * NX), uo(u, u + N * NU);
      double alpha = 1.0;
      bool ok = true;
      for (;;) {
        for (int k = 0; k <= N; k++) {
          if (k < N) for (int a = 0; a < NU; a++) u[k * NU + a] = uo[k * NU + a] + alpha * (sol.z[k * NZ + a] - uo[k * NU + a]);
          for (int a = 0; a < NX; a++) x[k * NX + a] = xo[k * NX + a] + alpha * (sol.z[k * NZ + NU + a] - xo[k * NX + a]);
        }
        ok = linearize(pb, x, u, qp, phi.data());
        const double m1 = ok ? merit_value(cf, qp, mw, x0, x, u, phi.data()) : std::numeric_limits<double>::quiet_NaN();
        if (m1 < m0 || alpha <= alpha_min) break;
        alpha = std::max(reduction * alpha, alpha_min);
      }
      out.mu = alpha;  // statistics slot [2] carries the accepted step length in this mode
      if (!ok) { out.status = IHM2B_FAILURE; break; }
      lin_valid = true;
      continue;
    }
    const double alpha = 1.0;  // FIXED_STEP; (RTI always takes the full step)
    for (int k = 0; k <= N; k++) {
      if (k < N) for (int a = 0; a < NU; a++) u[k * NU + a] += alpha * (sol.z[k * NZ + a] - u[k * NU + a]);
      for (int a = 0; a < NX; a++) x[k * NX + a] += alpha * (sol.z[k * NZ + NU + a] - x[k * NX + a]);
    }
    if (cf.nlp_solver == IHM2B_SQP_RTI) out.status = IHM2B_SUCCESS;
  }
  std::vector<double> z((N + 1) * NZ, 0.0);
  for (int k = 0; k <= N; k++) {
    if (k < N) { z[k * NZ] = u[k * NU]; z[k * NZ + 1] = u[k * NU + 1]; }
    for (int a = 0; a < NX; a++) z[k * NZ + NU + a] = x[k * NX + a];
  }
  out.cost = ocp_cost(cf, qp, z.data());
  if (keep) *keep = sol;
  if (keep_qp) *keep_qp = qp;
}

// ---------------------------------------------------------------------------------------
// plant step (python/main.py:395-400, 475-502): Radau-IIA collocation, sim_steps sub-steps,
// model switch on the kinematic lateral acceleration.
// ---------------------------------------------------------------------------------------
int sim_step(const Problem& pb, const double* x, const double* u, double* xn) {
  const ihm2b_config& cf = *pb.cf;
  int model = cf.sim_model;
  if (model == IHM2B_SIM_KIN6_DYN6) {
    const double C = pb.p[P_LR] / (pb.p[P_LR] + pb.p[P_LF]);
    const double beta = std::atan(C * std::tan(x[7]));
    const double v2 = x[3] * x[3] + x[4] * x[4];
    model = (v2 * std::sin(beta) / pb.p[P_LR] <= cf.sim_switch) ? IHM2B_SIM_FKIN6 : IHM2B_SIM_FDYN6;
  } else if (model == IHM2B_SIM_KIN6C_DYN6C) {  // sim_node.cpp:199
    model = (std::hypot(x[3], x[4]) < cf.sim_switch) ? IHM2B_SIM_KIN6C : IHM2B_SIM_DYN6C;
  }
  const bool cartesian = cf.sim_model >= IHM2B_SIM_KIN6C;
  Tableau tb{4, cf.sim_A, cf.sim_b};
  const double h = cf.dt / cf.sim_steps;
  double xc[NX], xt[NX];
  std::memcpy(xc, x, sizeof xc);
  for (int st = 0; st < cf.sim_steps; st++) {
    int rc = (model == IHM2B_SIM_FKIN6)   ? irk_step(fkin6_rhs, true, tb, xc, u, pb.p, pb.tr, h, xt, nullptr)
             : (model == IHM2B_SIM_FDYN6) ? irk_step(fdyn6, false, tb, xc, u, pb.p, pb.tr, h, xt, nullptr)
             : (model == IHM2B_SIM_KIN6C) ? irk_step(kin6c, false, tb, xc, u, pb.p, pb.tr, h, xt, nullptr)
                                          : irk_step(dyn6c, false, tb, xc, u, pb.p, pb.tr, h, xt, nullptr);
    if (rc < 0) {
      for (int a = 0; a < NX; a++) xn[a] = std::numeric_limits<double>::quiet_NaN();
      return model;
    }
    std::memcpy(xc, xt, sizeof xc);
  }
  // the sim node keeps the car from going backwards (sim_node.cpp:256-260)
  if (cartesian && (xc[3] < 0.0 || (xc[6] <= 0.1 && xc[3] < 0.01))) xc[3] = xc[4] = xc[5] = 0.0;
  std::memcpy(xn, xc, sizeof xc);
  return model;
}

void shift_iterate(int N, double* x, double* u) {  // controller.py:388-393 (literal)
  for (int j = 0; j < N - 1; j++) {
    std::memcpy(&x[j * NX], &x[(j + 1) * NX], NX * sizeof(double));
    std::memcpy(&u[j * NU], &u[(j + 1) * NU], NU * sizeof(double));
  }
  std::memcpy(&x[(N - 1) * NX], &x[N * NX], NX * sizeof(double));
  u[(N - 1) * NU] = u[(N - 1) * NU + 1] = 0.0;
}

}  // namespace

// =======================================================================================
// C ABI (host memory, batch-major rows).  Mirrors include/ihm2b.h so one Python harness can
// drive both sides of a parity test.
// =======================================================================================

// ---- DPC rollout ---------------------------------------------------------------
// kin4 model x = (X, Y, phi, v), u = (T, delta) of dpc/main.py:56-84, one RK4 step of dt = 1/20
// (dpc/main.py:87-97), cost of dpc/main.py:374-409 (tracking terms on stages 1..Nf-1, control terms on
// stages 0..Nf-1, terminal terms with the _f weights).  The gradient with respect to the controls is what a
// control policy trained through the rollout receives.  Dense and generic on purpose: full 4x4 / 4x2
// Jacobians and transposed products.
namespace dpc {
constexpr int DX = 4, DU = 2;
constexpr double M = 230.0, WHEELBASE = 1.5706, CM0 = 4.950, CR0 = 297.030, CR1 = 16.665, CR2 = 0.6784;  // :23-29

static void f_kin4(const double* x, const double* u, double* f, double* Jx, double* Ju) {
  const double phi = x[2], v = x[3], T = u[0], delta = u[1];
  const double beta = 0.5 * delta, lR = 0.5 * WHEELBASE;
  const double vx = v * std::cos(beta);
  const double th = std::tanh(10.0 * vx);
  const double poly = CR0 + CR1 * vx + CR2 * vx * vx;
  f[0] = v * std::cos(phi + beta);
  f[1] = v * std::sin(phi + beta);
  f[2] = v * std::sin(beta) / lR;
  f[3] = (CM0 * T - poly * th) / M;
  if (!Jx) return;
  std::fill(Jx, Jx + DX * DX, 0.0);
  std::fill(Ju, Ju + DX * DU, 0.0);
  const double dres = (CR1 + 2.0 * CR2 * vx) * th + poly * 10.0 * (1.0 - th * th);  // d(poly tanh)/d vx
  const double dvx_dv = std::cos(beta), dvx_dd = -0.5 * v * std::sin(beta);
  Jx[0 * DX + 2] = -v * std::sin(phi + beta);  Jx[0 * DX + 3] = std::cos(phi + beta);
  Jx[1 * DX + 2] = v * std::cos(phi + beta);   Jx[1 * DX + 3] = std::sin(phi + beta);
  Jx[2 * DX + 3] = std::sin(beta) / lR;
  Jx[3 * DX + 3] = -dres * dvx_dv / M;
  Ju[0 * DU + 1] = -0.5 * v * std::sin(phi + beta);
  Ju[1 * DU + 1] = 0.5 * v * std::cos(phi + beta);
  Ju[2 * DU + 1] = 0.5 * v * std::cos(beta) / lR;
  Ju[3 * DU + 0] = CM0 / M;
  Ju[3 * DU + 1] = -dres * dvx_dd / M;
}

// one RK4 step; with Sx/Su also d x+/d x [4x4] and d x+/d u [4x2] by the chain rule through the stages
static void rk4_kin4(const double* x, const double* u, double dt, double* xn, double* Sx, double* Su) {
  const double cw[4] = {0.0, 0.5, 0.5, 1.0}, bw[4] = {1.0 / 6, 1.0 / 3, 1.0 / 3, 1.0 / 6};
  double k[4][DX], dkx[4][DX * DX], dku[4][DX * DU];
  for (int s = 0; s < 4; s++) {
    double xs[DX], Jx[DX * DX], Ju[DX * DU];
    for (int a = 0; a < DX; a++) xs[a] = x[a] + (s ? cw[s] * dt * k[s - 1][a] : 0.0);
    f_kin4(xs, u, k[s], Sx ? Jx : nullptr, Ju);
    if (!Sx) continue;
    // d xs / d x = I + c dt dk_{s-1}/dx ; d xs / d u = c dt dk_{s-1}/du
    for (int a = 0; a < DX; a++) {
      for (int b = 0; b < DX; b++) {
        double acc = 0.0;
        for (int c = 0; c < DX; c++)
          acc += Jx[a * DX + c] * ((c == b ? 1.0 : 0.0) + (s ? cw[s] * dt * dkx[s - 1][c * DX + b] : 0.0));
        dkx[s][a * DX + b] = acc;
      }
      for (int b = 0; b < DU; b++) {
        double acc = Ju[a * DU + b];
        for (int c = 0; c < DX; c++) acc += Jx[a * DX + c] * (s ? cw[s] * dt * dku[s - 1][c * DU + b] : 0.0);
        dku[s][a * DU + b] = acc;
      }
    }
  }
  for (int a = 0; a < DX; a++) {
    xn[a] = x[a];
    for (int s = 0; s < 4; s++) xn[a] += dt * bw[s] * k[s][a];
  }
  if (!Sx) return;
  for (int a = 0; a < DX; a++) {
    for (int b = 0; b < DX; b++) {
      Sx[a * DX + b] = (a == b) ? 1.0 : 0.0;
      for (int s = 0; s < 4; s++) Sx[a * DX + b] += dt * bw[s] * dkx[s][a * DX + b];
    }
    for (int b = 0; b < DU; b++) {
      Su[a * DU + b] = 0.0;
      for (int s = 0; s < 4; s++) Su[a * DU + b] += dt * bw[s] * dku[s][a * DU + b];
    }
  }
}

// tracking term q_lon e_lon^2 + q_lat e_lat^2 + q_phi (phi - phi_ref)^2 + q_v (v - v_ref)^2 and its x-gradient
static double track_cost(const double* x, const double* r, const double* q, double* g) {
  const double cp = std::cos(r[2]), sp = std::sin(r[2]);
  const double elon = cp * (x[0] - r[0]) + sp * (x[1] - r[1]);
  const double elat = -sp * (x[0] - r[0]) + cp * (x[1] - r[1]);
  const double ephi = x[2] - r[2], ev = x[3] - r[3];
  g[0] = 2.0 * q[0] * elon * cp - 2.0 * q[1] * elat * sp;
  g[1] = 2.0 * q[0] * elon * sp + 2.0 * q[1] * elat * cp;
  g[2] = 2.0 * q[2] * ephi;
  g[3] = 2.0 * q[3] * ev;
  return q[0] * elon * elon + q[1] * elat * elat + q[2] * ephi * ephi + q[3] * ev * ev;
}
}  // namespace dpc

extern "C" {

int ihm2o_np(void) { return NP; }

// OpenMP team size used by the batched entry points (torchrun exports OMP_NUM_THREADS=1)
int ihm2o_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}

void ihm2o_fkin6(const double* x, const double* u, const double* p, const double* s_ref,
                 const double* kappa_ref, int n_grid, double* f, double* J /* [8][10] or NULL */) {
  Track tr{n_grid, s_ref, kappa_ref};
  fkin6(x, u, p, tr, f, reinterpret_cast<double(*)[NZ]>(J));
}

void ihm2o_fdyn6(const double* x, const double* u, const double* p, const double* s_ref,
                 const double* kappa_ref, int n_grid, double* f) {
  Track tr{n_grid, s_ref, kappa_ref};
  fdyn6(x, u, p, tr, f);
}

// kind 0: ERK4; 1: IRK with analytic Newton (fkin6); 2: IRK on fdyn6 (no sensitivities)
int ihm2o_integrate(int kind, const double* tabA, const double* tabb, double h, const double* x,
                    const double* u, const double* p, const double* s_ref, const double* kappa_ref,
                    int n_grid, double* xn, double* S /* [8][10] or NULL */) {
  Track tr{n_grid, s_ref, kappa_ref};
  Tableau tb{4, tabA, tabb};
  auto Sm = reinterpret_cast<double(*)[NZ]>(S);
  if (kind == 0) { erk4_sens(x, u, p, tr, h, xn, Sm); return 0; }
  if (kind == 1) return irk_step(fkin6_rhs, true, tb, x, u, p, tr, h, xn, Sm);
  return irk_step(fdyn6, false, tb, x, u, p, tr, h, xn, nullptr);
}

// track tables of the batched entry points: s_ref, kappa_ref and (optional) right / left widths, [n_tracks][n_grid]
struct ihm2o_tracks { const double *s_ref, *kappa_ref, *right_width, *left_width; };

static Problem make_problem(const ihm2b_config* cf, const ihm2o_tracks* tk, const int32_t* track_index,
                            const double* params, int i) {
  const size_t off = (size_t)(track_index ? track_index[i] : 0) * cf->n_grid;
  Problem pb;
  pb.cf = cf;
  pb.tr = Track{cf->n_grid, tk->s_ref + off, tk->kappa_ref + off, tk->right_width ? tk->right_width + off : nullptr,
                tk->left_width ? tk->left_width + off : nullptr};
  pb.p = params + (size_t)i * NP;
  return pb;
}

// A [B][N][8][8], Bm [B][N][8][2], phi [B][N][8]
int ihm2o_linearize(const ihm2b_config* cf, const ihm2o_tracks* tk, const int32_t* track_index, const double* params, const double* x,
                    const double* u, double* A, double* Bm, double* phi) {
  const int N = cf->N;
  int bad = 0;
#pragma omp parallel for schedule(dynamic) reduction(+ : bad)
  for (int i = 0; i < cf->batch; i++) {
    Problem pb = make_problem(cf, tk, track_index, params, i);
    Qp qp; qp.N = N;
    std::vector<double> ph(N * NX);
    if (!linearize(pb, x + (size_t)i * (N + 1) * NX, u + (size_t)i * N * NU, qp, ph.data())) bad++;
    std::memcpy(A + (size_t)i * N * NX * NX, qp.A.data(), sizeof(double) * N * NX * NX);
    std::memcpy(Bm + (size_t)i * N * NX * NU, qp.B.data(), sizeof(double) * N * NX * NU);
    std::memcpy(phi + (size_t)i * N * NX, ph.data(), sizeof(double) * N * NX);
  }
  return bad;
}

// One controller call per instance: (optional literal shift) + SQP(-RTI) solve.
// x [B][N+1][8], u [B][N][2] in/out; u0 [B][2]; status [B]; stats [B][8].
int ihm2o_solve(const ihm2b_config* cf, const ihm2o_tracks* tk, const int32_t* track_index, const double* params, const double* x0, double* x,
                double* u, double* u0, int32_t* status, double* stats, int do_shift) {
  const int N = cf->N;
#pragma omp parallel for schedule(dynamic)
  for (int i = 0; i < cf->batch; i++) {
    Problem pb = make_problem(cf, tk, track_index, params, i);
    double* xi = x + (size_t)i * (N + 1) * NX;
    double* ui = u + (size_t)i * N * NU;
    if (do_shift) shift_iterate(N, xi, ui);
    SolveOut so;
    solve_one(pb, x0 + (size_t)i * NX, xi, ui, so, nullptr, nullptr);
    status[i] = so.status;
    u0[i * NU] = ui[0]; u0[i * NU + 1] = ui[1];
    if (stats) {
      double* st = stats + (size_t)i * 8;
      st[0] = so.qp_iters; st[1] = so.sqp_iters; st[2] = so.mu; st[3] = so.res_stat;
      st[4] = so.res_eq; st[5] = so.res_ineq; st[6] = so.res_comp; st[7] = so.cost;
    }
  }
  return 0;
}

// Debug view of ONE problem's first QP and its IPM solution, for the KKT-certificate tests.
// Row storage padded to 16 rows per stage: Crow [N+1][16][10], e/lb/ub/lamL/lamU/tL/tU/sL/sU/lamSL/lamSU [N+1][16],
// soft [N+1][16], nrows [N+1], zZ [N+1][2] slack penalties (z_s, Z_s) of the stage.  z [N+1][10], pi [N+1][8].
int ihm2o_qp_debug(const ihm2b_config* cf, const ihm2o_tracks* tk, const double* params, const double* x0,
                   const double* x, const double* u, double* A, double* Bm, double* c, double* H, double* g,
                   double* Crow, double* e, double* lb, double* ub, int32_t* soft, int32_t* nrows, double* zZ,
                   double* z, double* pi, double* lamL, double* lamU, double* tL, double* tU, double* sL, double* sU,
                   double* lamSL, double* lamSU, double* info /* iters,status,mu,res_g,res_b,res_d,res_m */) {
  const int N = cf->N, PR = 16;
  ihm2b_config c1 = *cf;
  c1.batch = 1; c1.nlp_solver = IHM2B_SQP_RTI;
  Problem pb = make_problem(&c1, tk, nullptr, params, 0);
  std::vector<double> xi(x, x + (N + 1) * NX), ui(u, u + N * NU);
  SolveOut so; QpSol sol; Qp qp;
  solve_one(pb, x0, xi.data(), ui.data(), so, &sol, &qp);
  std::memcpy(A, qp.A.data(), sizeof(double) * N * NX * NX);
  std::memcpy(Bm, qp.B.data(), sizeof(double) * N * NX * NU);
  std::memcpy(c, qp.c.data(), sizeof(double) * N * NX);
  std::memcpy(H, qp.H.data(), sizeof(double) * (N + 1) * NZ * NZ);
  std::memcpy(g, qp.g.data(), sizeof(double) * (N + 1) * NZ);
  for (int k = 0; k <= N; k++) {
    nrows[k] = (int)qp.rows[k].size();
    zZ[k * 2] = qp.zs[k]; zZ[k * 2 + 1] = qp.Zs[k];
    for (int r = 0; r < nrows[k]; r++) {
      const int q = k * PR + r;
      std::memcpy(Crow + (size_t)q * NZ, qp.rows[k][r].c, sizeof(double) * NZ);
      e[q] = qp.rows[k][r].e; lb[q] = qp.rows[k][r].lb; ub[q] = qp.rows[k][r].ub; soft[q] = qp.rows[k][r].soft;
      lamL[q] = sol.lamL[k][r]; lamU[q] = sol.lamU[k][r]; tL[q] = sol.tL[k][r]; tU[q] = sol.tU[k][r];
      sL[q] = sol.sL[k][r]; sU[q] = sol.sU[k][r]; lamSL[q] = sol.lamSL[k][r]; lamSU[q] = sol.lamSU[k][r];
    }
  }
  std::memcpy(z, sol.z.data(), sizeof(double) * (N + 1) * NZ);
  std::memcpy(pi, sol.pi.data(), sizeof(double) * (N + 1) * NX);
  info[0] = sol.iters; info[1] = sol.status; info[2] = sol.mu; info[3] = sol.res_g;
  info[4] = sol.res_b; info[5] = sol.res_d; info[6] = sol.res_m;
  return so.status;
}

// h(x, u) and its Jacobian [5][10] at one point (track 0, first parameter vector)
void ihm2o_con_h(const ihm2b_config* cf, const ihm2o_tracks* tk, const double* params, const double* x,
                 const double* u, double* h, double* Jh) {
  Problem pb = make_problem(cf, tk, nullptr, params, 0);
  con_h(x, u, pb.p, pb.tr, cf->track_half_L, cf->track_half_W, h, reinterpret_cast<double(*)[10]>(Jh));
}

// x [B][8], u [B][2] -> xn [B][8], model [B]
int ihm2o_sim_step(const ihm2b_config* cf, const ihm2o_tracks* tk, const int32_t* track_index, const double* params, const double* x,
                   const double* u, double* xn, int32_t* model) {
#pragma omp parallel for schedule(dynamic)
  for (int i = 0; i < cf->batch; i++) {
    Problem pb = make_problem(cf, tk, track_index, params, i);
    const int m = sim_step(pb, x + (size_t)i * NX, u + (size_t)i * NU, xn + (size_t)i * NX);
    if (model) model[i] = m;
  }
  return 0;
}

// dyn10 plant: right-hand side at one point (track 0, first parameter vector) and the batched step
void ihm2o_fdyn10(const ihm2b_config* cf, const ihm2o_tracks* tk, const double* params, const double* x, const double* u, double* f) {
  Problem pb = make_problem(cf, tk, nullptr, params, 0);
  fdyn10(x, u, pb.p, pb.tr, f);
}

// x [B][15], u [B][5] -> xn [B][15] (NaN rows on failure): cfg.sim_steps Radau-IIA sub-steps over cfg.dt
int ihm2o_sim_step_dyn10(const ihm2b_config* cf, const ihm2o_tracks* tk, const int32_t* track_index, const double* params,
                         const double* x, const double* u, double* xn) {
#pragma omp parallel for schedule(dynamic)
  for (int i = 0; i < cf->batch; i++) {
    Problem pb = make_problem(cf, tk, track_index, params, i);
    Tableau tb{4, cf->sim_A, cf->sim_b};
    const double h = cf->dt / cf->sim_steps;
    double xc[NX10], xt[NX10];
    std::memcpy(xc, x + (size_t)i * NX10, sizeof xc);
    const double* ui = u + (size_t)i * NU10;
    bool ok = true;
    for (int st = 0; st < cf->sim_steps && ok; st++) {
      auto rhs = [&](const double* xx, double* ff) { fdyn10(xx, ui, pb.p, pb.tr, ff); };
      ok = irk_step_plant(NX10, rhs, tb, xc, h, xt) == 0;
      std::memcpy(xc, xt, sizeof xc);
    }
    for (int a = 0; a < NX10; a++) xn[(size_t)i * NX10 + a] = ok ? xc[a] : std::numeric_limits<double>::quiet_NaN();
  }
  return 0;
}

// python/main.py:448-517 with the NMPC enabled.  xs [B][8] in/out; iterate x/u in/out;
// xlog [n_steps+1][B][8], ulog [n_steps][B][2] (optional); done/fail [B] step index or -1;
// lap_time [B] (optional): python/main.py:561-570 -- dt x (second - first index at which mod(s, lap_length) jumps by
// more than 1 between consecutive states), NaN while fewer than two jumps have been seen; cost_sum [B] (optional):
// sum of the OCP cost at the solution over the control steps taken.
int ihm2o_closed_loop(const ihm2b_config* cf, const ihm2o_tracks* tk, const int32_t* track_index, const double* params,
                      const double* lap_length, int n_steps, double* xs, double* x, double* u, double* xlog,
                      double* ulog, int32_t* done, int32_t* fail, double* lap_time, double* cost_sum) {
  const int N = cf->N, B = cf->batch;
#pragma omp parallel for schedule(dynamic)
  for (int i = 0; i < B; i++) {
    Problem pb = make_problem(cf, tk, track_index, params, i);
    const int t = track_index ? track_index[i] : 0;
    double* xi = x + (size_t)i * (N + 1) * NX;
    double* ui = u + (size_t)i * N * NU;
    double* xc = xs + (size_t)i * NX;
    done[i] = fail[i] = -1;
    int jumps = 0, first_jump = -1;
    if (lap_time) lap_time[i] = std::numeric_limits<double>::quiet_NaN();
    if (cost_sum) cost_sum[i] = 0.0;
    if (xlog) std::memcpy(xlog + (size_t)i * NX, xc, NX * sizeof(double));
    for (int st = 0; st < n_steps; st++) {
      bool active = done[i] < 0 && fail[i] < 0;
      double uu[NU] = {0, 0};
      if (active) {
        shift_iterate(N, xi, ui);
        SolveOut so;
        solve_one(pb, xc, xi, ui, so, nullptr, nullptr);
        if (so.status != IHM2B_SUCCESS && so.status != IHM2B_MAXITER) { fail[i] = st; active = false; }
        else { uu[0] = ui[0]; uu[1] = ui[1]; if (cost_sum) cost_sum[i] += so.cost; }
      }
      if (active) {
        double xn[NX];
        sim_step(pb, xc, uu, xn);
        bool nan = false;
        for (int a = 0; a < NX; a++) nan |= !(xn[a] == xn[a]);
        if (nan) fail[i] = st;  // main.py:503-512: keep the last state
        else {
          const double L = lap_length[t];
          const double m0 = xc[0] - L * std::floor(xc[0] / L), m1 = xn[0] - L * std::floor(xn[0] / L);   // np.mod
          if (std::fabs(m1 - m0) > 1.0) {   // main.py:561-566
            if (jumps == 0) first_jump = st;
            else if (jumps == 1 && lap_time) lap_time[i] = cf->dt * (st - first_jump);
            jumps++;
          }
          std::memcpy(xc, xn, sizeof xn);
        }
        if (!nan && xc[0] > lap_length[t] + 1.0) done[i] = st;  // main.py:515
      }
      if (ulog) { ulog[((size_t)st * B + i) * NU] = uu[0]; ulog[((size_t)st * B + i) * NU + 1] = uu[1]; }
      if (xlog) std::memcpy(xlog + ((size_t)(st + 1) * B + i) * NX, xc, NX * sizeof(double));
    }
  }
  return 0;
}

// x0 [B][4], u [B][Nf][2], ref [B][Nf+1][4] = (X_ref, Y_ref, phi_ref, v_ref), w[10] = (q_lon, q_lat, q_phi, q_v,
// r_T, r_delta, q_lon_f, q_lat_f, q_phi_f, q_v_f) -> cost [B], grad [B][Nf][2] = d cost / d u, xtraj [B][Nf+1][4]
int ihm2o_dpc_rollout_grad(int B, int Nf, double dt, const double* x0, const double* u, const double* ref,
                           const double* w, double* cost, double* grad, double* xtraj) {
  using namespace dpc;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < B; i++) {
    std::vector<double> xs((size_t)(Nf + 1) * DX), Sx((size_t)Nf * DX * DX), Su((size_t)Nf * DX * DU);
    const double* ui = u + (size_t)i * Nf * DU;
    const double* ri = ref + (size_t)i * (Nf + 1) * DX;
    std::memcpy(xs.data(), x0 + (size_t)i * DX, DX * sizeof(double));
    double J = 0.0;
    for (int k = 0; k < Nf; k++) {
      rk4_kin4(&xs[(size_t)k * DX], ui + k * DU, dt, &xs[(size_t)(k + 1) * DX], &Sx[(size_t)k * DX * DX],
               &Su[(size_t)k * DX * DU]);
      J += w[4] * ui[k * DU] * ui[k * DU] + w[5] * ui[k * DU + 1] * ui[k * DU + 1];
    }
    double a[DX], g[DX];
    J += track_cost(&xs[(size_t)Nf * DX], ri + Nf * DX, w + 6, a);
    for (int k = Nf - 1; k >= 0; k--) {
      double* gk = grad + ((size_t)i * Nf + k) * DU;
      for (int b = 0; b < DU; b++) {
        double acc = 2.0 * w[4 + b] * ui[k * DU + b];
        for (int c = 0; c < DX; c++) acc += Su[((size_t)k * DX + c) * DU + b] * a[c];
        gk[b] = acc;
      }
      double an[DX];
      for (int b = 0; b < DX; b++) {
        an[b] = 0.0;
        for (int c = 0; c < DX; c++) an[b] += Sx[((size_t)k * DX + c) * DX + b] * a[c];
      }
      if (k > 0) {
        J += track_cost(&xs[(size_t)k * DX], ri + k * DX, w, g);
        for (int b = 0; b < DX; b++) an[b] += g[b];
      }
      std::memcpy(a, an, sizeof an);
    }
    cost[i] = J;
    if (xtraj) std::memcpy(xtraj + (size_t)i * (Nf + 1) * DX, xs.data(), xs.size() * sizeof(double));
  }
  return 0;
}

// ---- Cartesian pose -> Frenet state -------------------------------------------
// Track::project (tracks.cpp:200-288): window of table points with s within s_guess +- s_tol (one more point on
// each side), closest point, second end of the segment chosen by the larger angle seen from the car, orthogonal
// projection on that segment; phi_ref from the linear interpolant anchored at the closest point (tracks.cpp:191-198,
// 169-170).  Then the control node's n and psi (mpc_control_node.cpp:148-157).
namespace proj {
double wrap_to_pi(double x) {  // math.cpp:19-25
  double t = std::fmod(x + M_PI, 2 * M_PI);
  if (t < 0) t += 2 * M_PI;
  return t - M_PI;
}
long long locate_index(const double* v, int n, double x) { return std::upper_bound(v, v + n, x) - v - 1; }  // tracks.cpp:183-185
double angle3pt(const double* a, const double* b, const double* c) {  // tracks.cpp:187-189
  return wrap_to_pi(std::atan2(c[1] - b[1], c[0] - b[0]) - std::atan2(a[1] - b[1], a[0] - b[0]));
}
}  // namespace proj

// tables [n_tracks][n_grid]; pose [B][3] = (X, Y, phi); out frenet [B][3] = (s, n, psi), ref [B][3] = (X_ref, Y_ref,
// phi_ref) at the projection, flag [B] = 1 where |tangent x normal| < 1e-6 (the node throws there)
int ihm2o_project(int B, int n_grid, const double* s_ref, const double* X_ref, const double* Y_ref,
                  const double* phi_ref, const int32_t* track, const double* pose, const double* s_guess,
                  double s_tol, double* frenet, double* ref, int32_t* flag) {
  using namespace proj;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < B; i++) {
    const size_t off = (size_t)(track ? track[i] : 0) * n_grid;
    const double *sr = s_ref + off, *Xr = X_ref + off, *Yr = Y_ref + off, *pr = phi_ref + off;
    const double car[2] = {pose[i * 3], pose[i * 3 + 1]};
    const double s_low = std::max(s_guess[i] - s_tol, sr[0]), s_up = std::min(s_guess[i] + s_tol, sr[n_grid - 1]);
    long long id_low = locate_index(sr, n_grid, s_low), id_up = locate_index(sr, n_grid, s_up);
    if (id_low > 0) --id_low;
    if (id_up < n_grid - 1) ++id_up;
    const long long rows = id_up - id_low + 1;
    long long id_min = 0;
    double best = std::numeric_limits<double>::infinity();
    for (long long j = 0; j < rows; j++) {
      const double dx = Xr[id_low + j] - car[0], dy = Yr[id_low + j] - car[1];
      const double d = dx * dx + dy * dy;
      if (d < best) { best = d; id_min = j; }
    }
    const long long id_prev = (id_min == 0) ? rows - 1 : id_min - 1;
    const long long id_next = (id_min == rows - 1) ? 0 : id_min + 1;
    auto pt = [&](long long j, double* o) { o[0] = Xr[id_low + j]; o[1] = Yr[id_low + j]; };
    double pm[2], pp[2], pn[2];
    pt(id_min, pm); pt(id_prev, pp); pt(id_next, pn);
    const double angle_prev = std::fabs(angle3pt(pm, car, pp)), angle_next = std::fabs(angle3pt(pm, car, pn));
    const double *a, *b;
    double sa, sb;
    if (angle_prev > angle_next) { a = pp; b = pm; sa = sr[id_prev + id_low]; sb = sr[id_min + id_low]; }
    else { a = pm; b = pn; sa = sr[id_min + id_low]; sb = sr[id_next + id_low]; }
    const double dx = b[0] - a[0], dy = b[1] - a[1];
    const double lambda = ((car[0] - a[0]) * dx + (car[1] - a[1]) * dy) / (dx * dx + dy * dy);
    const double s = sa + lambda * (sb - sa);
    const double Xp = a[0] + lambda * (b[0] - a[0]), Yp = a[1] + lambda * (b[1] - a[1]);
    // the reference indexes its (n_grid - 1)-row coefficient table with id_min + id_low unclamped
    const long long ind = std::min<long long>(id_min + id_low, n_grid - 2);
    const double phir = pr[ind] + (pr[ind + 1] - pr[ind]) / (sr[ind + 1] - sr[ind]) * (s - sr[ind]);
    const double rho = wrap_to_pi(phir);
    const double psi = wrap_to_pi(pose[i * 3 + 2] - rho);
    const double e = std::hypot(Xp - car[0], Yp - car[1]);
    const double tpr = (car[1] - Yp) * std::cos(rho) - (car[0] - Xp) * std::sin(rho);
    frenet[i * 3] = s; frenet[i * 3 + 1] = e * (tpr > 0.0 ? 1.0 : -1.0); frenet[i * 3 + 2] = psi;
    if (ref) { ref[i * 3] = Xp; ref[i * 3 + 1] = Yp; ref[i * 3 + 2] = phir; }
    if (flag) flag[i] = std::fabs(tpr) < 1e-6;
  }
  return 0;
}

}  // extern "C"
