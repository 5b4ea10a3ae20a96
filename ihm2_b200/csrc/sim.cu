// sim.cu -- plant step of the closed loop (python/main.py:395-400, 475-502 of the reference):
// AcadosSimSolver with IRK, collocation_type GAUSS_RADAU_IIA, 4 stages, `sim_steps` sub-steps over
// dt, model fkin6 (python/models.py:232-307) or fdyn6 (:455-606) chosen per step by the kinematic
// lateral acceleration v^2 sin(beta) / l_R (main.py:482-489).
//
// Mapping: 4 lanes per problem, lane i evaluates the model at collocation point i, so the four
// model evaluations of a sub-step run side by side; stage derivatives are exchanged with
// shuffles inside the 4-lane group.  With h = dt/100 the collocation equations are far from
// stiff (h |J| <~ 0.1) once the two actuator lags are taken out in closed form, so they are
// solved by fixed-point iteration to round-off, warm-started from the previous sub-step, which
// converges to the same point as acados' Newton iteration.
//
// fdyn6 is stated implicitly in the reference (accelerations feed the vertical loads); the
// force balance is affine in (a_x, a_y), so the tyre terms are evaluated once per point and a
// 2 x 2 linear solve yields the explicit right-hand side.
#include "model.cuh"
#include <mutex>

namespace ihm2b {

namespace {

__device__ __forceinline__ double smooth_abs_nonzero(double v) {  // python/utils.py:27-32
  return tanh(10.0 * v) * v + 1e-6 * exp(-v * v);
}
__device__ __forceinline__ double lat_pacejka(const double* p, double al) {  // python/models.py:69-74
  const double Ba = p[IHM2B_P_BA] * al;
  return p[IHM2B_P_DA] * sin(p[IHM2B_P_CA] * atan(Ba - p[IHM2B_P_EA] * (Ba - atan(Ba))));
}

// rows 0..5 of the fdyn6 right-hand side (rows 6, 7 are the actuator lags, handled in closed form)
// cart: the Cartesian dyn6 of python/models.py:310-452 instead (pose kinematics, slip angles atan2(v_y_i, v_x_i)
// [- delta] assigned wheel by wheel, drag switched by tanh(1000 v_x), r_kin = v_x sin(beta) / l_R)
__device__ void fdyn6_rhs(const double* __restrict__ x, const double* __restrict__ p, const TrackView& tr,
                          double* __restrict__ f, const bool cart) {
  const double n = x[1], psi = x[2], vx = x[3], vy = x[4], r = x[5], T = x[6], dl = x[7];
  const double m = p[IHM2B_P_M], g = p[IHM2B_P_G], lR = p[IHM2B_P_LR], lF = p[IHM2B_P_LF], wb = lR + lF;
  const double tw = p[IHM2B_P_TRACKW];
  const double Fdown = 0.5 * p[IHM2B_P_CDOWN] * vx * vx;
  const double stat = 0.5 * m * g * lF / wb;
  const double cl = 0.5 * m * p[IHM2B_P_ZCG] / wb;   // d(lonT)/d a_x
  const double ct = 0.5 * m * p[IHM2B_P_ZCG] / tw;   // d(latT)/d a_y
  double sd, cd;
  sincos(dl, &sd, &cd);
  const double vxL = vx - 0.5 * tw * r, vxR = vx + 0.5 * tw * r;
  const double vyF = vy + lF * r, vlatR = vy - lR * r;
  const double aFL = cart ? atan2(vyF, vxL) - dl : atan2(-sd * vxL + cd * vyF, smooth_abs_nonzero(cd * vxL + sd * vyF));
  const double aFR = cart ? atan2(vyF, vxR) - dl : atan2(-sd * vxR + cd * vyF, smooth_abs_nonzero(cd * vxR + sd * vyF));
  const double aRL = atan2(vlatR, cart ? vxL : smooth_abs_nonzero(vxL));
  const double aRR = atan2(vlatR, cart ? vxR : smooth_abs_nonzero(vxR));
  const bool cross = !cart && p[IHM2B_P_CROSS] != 0.0;  // literal quirk python/models.py:543-546
  const double muFL = lat_pacejka(p, cross ? aRR : aFL), muFR = lat_pacejka(p, cross ? aRL : aFR);
  const double muRL = lat_pacejka(p, cross ? aFR : aRL), muRR = lat_pacejka(p, cross ? aFL : aRR);
  const double Fdrag = -(p[IHM2B_P_CR0] + p[IHM2B_P_CR1] * vx + p[IHM2B_P_CR2] * vx * vx) * tanh((cart ? 1000.0 : 10.0) * vx);
  const double C = lR / wb;
  const double td = sd / cd, Ctd = C * td;
  const double sbeta = Ctd * rsqrt(1.0 + Ctd * Ctd);
  const double dtau = p[IHM2B_P_KTV] * ((cart ? vx : hypot(vx, vy)) * sbeta / lR - r);
  const double idn = 1.0 / (-m * g - 0.25 * Fdown);
  const double lonL = p[IHM2B_P_CM0] * (T - dtau) * idn, lonR = p[IHM2B_P_CM0] * (T + dtau) * idn;
  // body-frame force per unit vertical load of each wheel, and its yaw arm
  const double fxFL = lonL * cd - muFL * sd, fyFL = lonL * sd + muFL * cd;
  const double fxFR = lonR * cd - muFR * sd, fyFR = lonR * sd + muFR * cd;
  const double fxRL = lonL, fyRL = muRL, fxRR = lonR, fyRR = muRR;
  const double mzFL = -0.5 * tw * fxFL + lF * fyFL, mzFR = 0.5 * tw * fxFR + lF * fyFR;
  const double mzRL = -0.5 * tw * fxRL - lR * fyRL, mzRR = 0.5 * tw * fxRR - lR * fyRR;
  // Fz_i = z0 + cx_i a_x + cy_i a_y   with (cx, cy) = FL (+cl,-ct) FR (+cl,+ct) RL (-cl,-ct) RR (-cl,+ct)
  const double z0 = -(stat + 0.25 * Fdown);
  const double Fx0 = z0 * (fxFL + fxFR + fxRL + fxRR) + Fdrag, Fy0 = z0 * (fyFL + fyFR + fyRL + fyRR);
  const double Mz0 = z0 * (mzFL + mzFR + mzRL + mzRR);
  const double gxx = cl * (fxFL + fxFR - fxRL - fxRR), gxy = ct * (-fxFL + fxFR - fxRL + fxRR);
  const double gyx = cl * (fyFL + fyFR - fyRL - fyRR), gyy = ct * (-fyFL + fyFR - fyRL + fyRR);
  const double gzx = cl * (mzFL + mzFR - mzRL - mzRR), gzy = ct * (-mzFL + mzFR - mzRL + mzRR);
  const double a11 = m - gxx, a12 = -gxy, a21 = -gyx, a22 = m - gyy;
  const double idet = 1.0 / (a11 * a22 - a12 * a21);
  const double ax = (Fx0 * a22 - a12 * Fy0) * idet, ay = (a11 * Fy0 - a21 * Fx0) * idet;
  const double Mz = Mz0 + gzx * ax + gzy * ay;
  double sp, cp;
  sincos(psi, &sp, &cp);
  if (cart) {
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
  f[3] = ax + vy * r;
  f[4] = ay - vx * r;
  f[5] = Mz / p[IHM2B_P_IZ];
}

// rows 0..5 of the plant right-hand side for the model chosen for this step
__device__ __forceinline__ void plant_rhs(const int model, const double* x, const double u0, const double u1,
                                          const double* p, const TrackView& tr, double* f) {
  if (model == IHM2B_SIM_FKIN6) {
    fkin6<false, double>(x, u0, u1, p, tr, f, nullptr);
  } else if (model == IHM2B_SIM_KIN6C) {
    // Cartesian kin6 (python/models.py:168-229): the v_x, v_y rows of fkin6, pose kinematics, and r_dot = v_y_dot / l_R
    // (fkin6 has l_R * v_y_dot - beta_dot there, :307)
    double xx[NX];
    for (int r = 0; r < NX; r++) xx[r] = x[r];
    xx[0] = 0.0; xx[1] = 0.0;
    fkin6<false, double>(xx, u0, u1, p, tr, f, nullptr);
    double sp, cp;
    sincos(x[2], &sp, &cp);
    f[0] = x[3] * cp - x[4] * sp;
    f[1] = x[3] * sp + x[4] * cp;
    f[2] = x[5];
    f[5] = f[4] / p[IHM2B_P_LR];
  } else {
    fdyn6_rhs(x, p, tr, f, model == IHM2B_SIM_DYN6C);
  }
}

constexpr int MAX_FP = 80;
// fixed-point tolerance per sub-step on the stage derivatives (relative); 100 sub-steps keep the step well
// inside the 1e-9 the parity tests ask for
#ifndef FP_TOL
#define FP_TOL 1e-13
#endif

}  // namespace

struct SimCfg {
  int B, n_grid, model, n_sub;
  double h, sw;
  double A[16], b[4];
};

// blockDim 128 = 32 problems; lane group of 4 = one problem
__global__ void __launch_bounds__(128) sim_step_kernel(const SimCfg sc, const Work w, const double* __restrict__ xin,
                                                        const double* __restrict__ uin, double* __restrict__ xout,
                                                        int32_t* __restrict__ model_out, const int32_t* __restrict__ order) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int slot = gid >> 2, i = gid & 3;
  if (slot >= sc.B) return;  // whole 4-lane groups leave together
  const int b = order ? order[slot] : slot;
  if (w.active && !w.active[b]) return;   // closed loop: instance finished its lap or stopped
  const unsigned gm = 0xFu << (threadIdx.x & 28);
  const int t = w.track[b];
  TrackView tr{sc.n_grid, w.s_ref + (size_t)t * sc.n_grid, w.kappa_ref + (size_t)t * sc.n_grid};
  double p[NP], x[NX];
  for (int j = 0; j < NP; j++) p[j] = __ldg(w.params + (size_t)b * NP + j);
  for (int j = 0; j < NX; j++) x[j] = xin[(size_t)b * NX + j];
  const double u0 = uin[(size_t)b * 2], u1 = uin[(size_t)b * 2 + 1];
  double wT[4], wD[4];
  for (int j = 0; j < 4; j++) { wT[j] = __ldg(w.actw_sim + (size_t)b * 8 + j); wD[j] = __ldg(w.actw_sim + (size_t)b * 8 + 4 + j); }
  int model = sc.model;
  if (model == IHM2B_SIM_KIN6_DYN6) {  // python/main.py:482-489
    const double C = p[IHM2B_P_LR] / (p[IHM2B_P_LR] + p[IHM2B_P_LF]);
    const double beta = atan(C * tan(x[7]));
    model = ((x[3] * x[3] + x[4] * x[4]) * sin(beta) / p[IHM2B_P_LR] <= sc.sw) ? IHM2B_SIM_FKIN6 : IHM2B_SIM_FDYN6;
  } else if (model == IHM2B_SIM_KIN6C_DYN6C) {  // sim_node.cpp:199
    model = (hypot(x[3], x[4]) < sc.sw) ? IHM2B_SIM_KIN6C : IHM2B_SIM_DYN6C;
  }
  const bool cartesian = sc.model >= IHM2B_SIM_KIN6C;
  const double h = sc.h, itT = 1.0 / p[IHM2B_P_TT], itD = 1.0 / p[IHM2B_P_TD];
  double Ai[4], bw[4];
  for (int j = 0; j < 4; j++) { Ai[j] = h * sc.A[i * 4 + j]; bw[j] = h * sc.b[j]; }
  double K[6], Kp[6];  // rows 0..5 of this lane's stage derivative; the converged one of the sub-step before
  bool ok = true, first = true, second = true;
#pragma unroll 1
  for (int sub = 0; sub < sc.n_sub; sub++) {
    // actuator lags in closed form: K_j = w_j (u - x) / t
    const double cT = (u0 - x[6]) * itT, cD = (u1 - x[7]) * itD;
    double xs[NX];
    xs[6] = x[6]; xs[7] = x[7];
    double dT = 0.0, dD = 0.0;
    for (int j = 0; j < 4; j++) { xs[6] = fma(Ai[j], wT[j] * cT, xs[6]); xs[7] = fma(Ai[j], wD[j] * cD, xs[7]); dT = fma(bw[j], wT[j] * cT, dT); dD = fma(bw[j], wD[j] * cD, dD); }
    if (first) {  // first sub-step: start from f(x); afterwards from the previous sub-step's stages
      double xx[NX], f[NX];
      for (int r = 0; r < NX; r++) xx[r] = x[r];
      plant_rhs(model, xx, u0, u1, p, tr, f);
      for (int r = 0; r < 6; r++) { K[r] = f[r]; Kp[r] = f[r]; }
      first = false;
    } else {
      // predictor: stage derivatives extrapolated linearly over the last two sub-steps (the stages of consecutive
      // sub-steps sit h apart), which starts the iteration O(h^2) instead of O(h) from its fixed point
      for (int r = 0; r < 6; r++) {
        const double kc = K[r];
        if (!second) K[r] = 2.0 * kc - Kp[r];
        Kp[r] = kc;
      }
      second = false;
    }
    int it = 0;
    double dprev = 1e300;
#pragma unroll 1
    for (;; it++) {
      for (int r = 0; r < 6; r++) {
        double v = x[r];
#pragma unroll
        for (int j = 0; j < 4; j++) v = fma(Ai[j], __shfl_sync(gm, K[r], j, 4), v);
        xs[r] = v;
      }
      double f[NX];
      plant_rhs(model, xs, u0, u1, p, tr, f);
      double dmax = 0.0, kmax = 0.0;
      for (int r = 0; r < 6; r++) { dmax = fmax(dmax, fabs(f[r] - K[r])); kmax = fmax(kmax, fabs(f[r])); K[r] = f[r]; }
      if (!(dmax == dmax)) dmax = 1e300;
      dmax = fmax(dmax, __shfl_xor_sync(gm, dmax, 1)); dmax = fmax(dmax, __shfl_xor_sync(gm, dmax, 2));
      kmax = fmax(kmax, __shfl_xor_sync(gm, kmax, 1)); kmax = fmax(kmax, __shfl_xor_sync(gm, kmax, 2));
      // converged to round-off: tiny update, or stagnation at the noise floor of the model evaluation
      if (dmax <= FP_TOL * (1.0 + kmax) || (it >= 3 && dmax >= 0.5 * dprev && dmax <= 1e-11 * (1.0 + kmax))) break;
      dprev = dmax;
      if (it >= MAX_FP || dmax >= 1e300) { ok = false; break; }
    }
    for (int r = 0; r < 6; r++) {
      double v = x[r];
#pragma unroll
      for (int j = 0; j < 4; j++) v = fma(bw[j], __shfl_sync(gm, K[r], j, 4), v);
      x[r] = v;
    }
    x[6] += dT; x[7] += dD;
    if (!ok) break;
  }
  // the sim node keeps the car from going backwards (sim_node.cpp:256-260)
  if (cartesian && (x[3] < 0.0 || (x[6] <= 0.1 && x[3] < 0.01))) { x[3] = 0.0; x[4] = 0.0; x[5] = 0.0; }
  if (i == 0) {
    for (int r = 0; r < NX; r++) xout[(size_t)b * NX + r] = ok ? x[r] : nan("");
    if (model_out) model_out[b] = model;
  }
}

// ---------------------------------------------------------------------------------------
// dyn10 plant: `fdyn10_model`, python/models.py:609-801.  x[15] = (s, n, psi, v_x, v_y, r, omega_FL,
// omega_FR, omega_RL, omega_RR, tau_FL .. tau_RR, delta), u[5] = (u_tau_FL .. u_tau_RR, u_delta).
// Rows 0..9 of the right-hand side at a point whose torques / steering angle are given (rows 10..14 are first-order
// lags, closed form).  As for fdyn6 every tyre force is proportional to its wheel's vertical load, which is affine in
// (a_x, a_y): one 2 x 2 solve.  dwdw (optional): d omega_w' / d omega_w, the stiff diagonal of the wheel-speed rows.
// ---------------------------------------------------------------------------------------
__device__ void fdyn10_rhs(const double* __restrict__ x, const double* __restrict__ tau, const double dl,
                           const double* __restrict__ p, const TrackView& tr, double* __restrict__ f, double* __restrict__ dwdw) {
  const double n = x[1], psi = x[2], vx = x[3], vy = x[4], r = x[5];
  const double m = p[IHM2B_P_M], g = p[IHM2B_P_G], lR = p[IHM2B_P_LR], lF = p[IHM2B_P_LF], wb = lR + lF, tw = p[IHM2B_P_TRACKW];
  const double Fdown = 0.5 * p[IHM2B_P_CDOWN] * vx * vx;
  const double stat = 0.5 * m * g * lF / wb;
  const double cl = 0.5 * m * p[IHM2B_P_ZCG] / wb, ct = 0.5 * m * p[IHM2B_P_ZCG] / tw;
  double sd, cd;
  sincos(dl, &sd, &cd);
  const double vxL = vx - 0.5 * tw * r, vxR = vx + 0.5 * tw * r, vyF = vy + lF * r, vlatR = vy - lR * r;
  const double vlon[4] = {cd * vxL + sd * vyF, cd * vxR + sd * vyF, vxL, vxR};
  const double vlat[4] = {-sd * vxL + cd * vyF, -sd * vxR + cd * vyF, vlatR, vlatR};
  double mu_lat[4], mu_lon[4], dmu[4];
#pragma unroll
  for (int w = 0; w < 4; w++) {
    const double va = smooth_abs_nonzero(vlon[w]);
    mu_lat[w] = lat_pacejka(p, atan2(vlat[w], va));
    const double iva = p[IHM2B_P_RW] / va;
    const double Bs = p[IHM2B_P_BS] * (x[6 + w] * iva - 1.0);         // B_s * slip ratio, models.py:723-726
    const double q = 1.0 / (1.0 + Bs * Bs);
    const double arg = Bs - p[IHM2B_P_ES] * (Bs - atan(Bs));
    double sa, ca;
    sincos(p[IHM2B_P_CS] * atan(arg), &sa, &ca);
    mu_lon[w] = -p[IHM2B_P_DS] * sa;                                  // F_lon = -F_z lon_pacejka(s), :728-732
    // d mu_lon / d omega
    dmu[w] = -p[IHM2B_P_DS] * p[IHM2B_P_CS] * ca / (1.0 + arg * arg) * (1.0 - p[IHM2B_P_ES] * (1.0 - q)) * p[IHM2B_P_BS] * iva;
  }
  const double Fdrag = -(p[IHM2B_P_CR0] + p[IHM2B_P_CR1] * vx + p[IHM2B_P_CR2] * vx * vx) * tanh(1000.0 * vx);
  double fx[4], fy[4], mz[4];
  fx[0] = mu_lon[0] * cd - mu_lat[0] * sd; fy[0] = mu_lon[0] * sd + mu_lat[0] * cd;
  fx[1] = mu_lon[1] * cd - mu_lat[1] * sd; fy[1] = mu_lon[1] * sd + mu_lat[1] * cd;
  fx[2] = mu_lon[2]; fy[2] = mu_lat[2]; fx[3] = mu_lon[3]; fy[3] = mu_lat[3];
  mz[0] = -0.5 * tw * fx[0] + lF * fy[0]; mz[1] = 0.5 * tw * fx[1] + lF * fy[1];
  mz[2] = -0.5 * tw * fx[2] - lR * fy[2]; mz[3] = 0.5 * tw * fx[3] - lR * fy[3];
  const double z0 = -(stat + 0.25 * Fdown);
  const double cx[4] = {cl, cl, -cl, -cl}, cy[4] = {-ct, ct, -ct, ct};
  double Fx0 = Fdrag, Fy0 = 0.0, gxx = 0.0, gxy = 0.0, gyx = 0.0, gyy = 0.0;
#pragma unroll
  for (int w = 0; w < 4; w++) {
    Fx0 = fma(z0, fx[w], Fx0); Fy0 = fma(z0, fy[w], Fy0);
    gxx = fma(cx[w], fx[w], gxx); gxy = fma(cy[w], fx[w], gxy); gyx = fma(cx[w], fy[w], gyx); gyy = fma(cy[w], fy[w], gyy);
  }
  const double a11 = m - gxx, a12 = -gxy, a21 = -gyx, a22 = m - gyy;
  const double idet = 1.0 / (a11 * a22 - a12 * a21);
  const double ax = (Fx0 * a22 - a12 * Fy0) * idet, ay = (a11 * Fy0 - a21 * Fx0) * idet;
  double Mz = 0.0;
  const double iIw = 1.0 / p[IHM2B_P_IW];
#pragma unroll
  for (int w = 0; w < 4; w++) {
    const double Fz = z0 + cx[w] * ax + cy[w] * ay;
    Mz = fma(Fz, mz[w], Mz);
    f[6 + w] = (tau[w] - (p[IHM2B_P_KD] * x[6 + w] + p[IHM2B_P_KS] + p[IHM2B_P_RW] * Fz * mu_lon[w])) * iIw;   // :781-784
    if (dwdw) dwdw[w] = -(p[IHM2B_P_KD] + p[IHM2B_P_RW] * Fz * dmu[w]) * iIw;
  }
  double kap, dkap, sp, cp;
  kappa_lookup(tr, x[0], kap, dkap);
  sincos(psi, &sp, &cp);
  const double sdot = (vx * cp - vy * sp) / (1.0 + kap * n);
  f[0] = sdot; f[1] = vx * sp + vy * cp; f[2] = r - kap * sdot;
  f[3] = ax + vy * r; f[4] = ay - vx * r; f[5] = Mz / p[IHM2B_P_IZ];
}

// row i of (I - z A)^-1 for the 4-stage tableau A: the stage coupling of a scalar stiff mode z = h lambda
__device__ __forceinline__ void inv_row4(const double z, const double* __restrict__ A, const int i, double* __restrict__ y) {
  double M[4][5];   // (I - z A)' y = e_i
#pragma unroll
  for (int r = 0; r < 4; r++) {
#pragma unroll
    for (int c = 0; c < 4; c++) M[r][c] = ((r == c) ? 1.0 : 0.0) - z * A[c * 4 + r];
    M[r][4] = (r == i) ? 1.0 : 0.0;
  }
#pragma unroll
  for (int c = 0; c < 4; c++) {
    const double ip = 1.0 / M[c][c];
#pragma unroll
    for (int r = c + 1; r < 4; r++) {
      const double l = M[r][c] * ip;
#pragma unroll
      for (int j = c + 1; j < 5; j++) M[r][j] = fma(-l, M[c][j], M[r][j]);
    }
  }
#pragma unroll
  for (int r = 3; r >= 0; r--) {
    double v = M[r][4];
#pragma unroll
    for (int c = r + 1; c < 4; c++) v = fma(-M[r][c], y[c], v);
    y[r] = v / M[r][r];
  }
}

// One plant step of the dyn10 model: 4 lanes per car (lane i = collocation stage i), Radau IIA, n_sub sub-steps.  The
// five actuator lags are closed form; the body rows are solved by fixed-point iteration like the other plants; the four
// wheel-speed rows are stiff through the tyre slip (|h d omega' / d omega| reaches 1 at h = 0.5 ms below ~3 m/s), so
// their update goes through the stage coupling (I - h lambda_w A)^-1 of each wheel's own stiff mode, lambda_w taken at
// the start of the sub-step (a simplified Newton iteration on the diagonal; converges to the same collocation point).
__global__ void __launch_bounds__(128) sim_step_dyn10_kernel(const SimCfg sc, const Work w, const double* __restrict__ xin,
                                                              const double* __restrict__ uin, double* __restrict__ xout) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = gid >> 2, i = gid & 3;
  if (b >= sc.B) return;
  const unsigned gm = 0xFu << (threadIdx.x & 28);
  const int t = w.track[b];
  TrackView tr{sc.n_grid, w.s_ref + (size_t)t * sc.n_grid, w.kappa_ref + (size_t)t * sc.n_grid};
  double p[NP], x[IHM2B_NX10], u[IHM2B_NU10];
  for (int j = 0; j < NP; j++) p[j] = __ldg(w.params + (size_t)b * NP + j);
  for (int j = 0; j < IHM2B_NX10; j++) x[j] = xin[(size_t)b * IHM2B_NX10 + j];
  for (int j = 0; j < IHM2B_NU10; j++) u[j] = uin[(size_t)b * IHM2B_NU10 + j];
  double wT[4], wD[4];
  for (int j = 0; j < 4; j++) { wT[j] = __ldg(w.actw_sim + (size_t)b * 8 + j); wD[j] = __ldg(w.actw_sim + (size_t)b * 8 + 4 + j); }
  const double h = sc.h, itT = 1.0 / p[IHM2B_P_TT], itD = 1.0 / p[IHM2B_P_TD];
  double Ai[4], bw[4];
  for (int j = 0; j < 4; j++) { Ai[j] = h * sc.A[i * 4 + j]; bw[j] = h * sc.b[j]; }
  double K[10], Kp[10];
  bool ok = true, first = true, second = true;
#pragma unroll 1
  for (int sub = 0; sub < sc.n_sub; sub++) {
    // actuator lags in closed form
    double tau_i[4], dtau[4], dl_i = x[14], ddl = 0.0;
    const double cD = (u[4] - x[14]) * itD;
    double sA = 0.0, sAd = 0.0, sB = 0.0, sBd = 0.0;
    for (int j = 0; j < 4; j++) { sA = fma(Ai[j], wT[j], sA); sAd = fma(Ai[j], wD[j], sAd); sB = fma(bw[j], wT[j], sB); sBd = fma(bw[j], wD[j], sBd); }
    for (int q = 0; q < 4; q++) { const double cT = (u[q] - x[10 + q]) * itT; tau_i[q] = fma(sA, cT, x[10 + q]); dtau[q] = sB * cT; }
    dl_i = fma(sAd, cD, x[14]); ddl = sBd * cD;
    // stiff diagonal of the wheel-speed rows at the start of the sub-step, this lane's row of each wheel's stage coupling
    double f[10], lam[4], inv[4][4];
    fdyn10_rhs(x, x + 10, x[14], p, tr, f, lam);
    for (int q = 0; q < 4; q++) inv_row4(h * lam[q], sc.A, i, inv[q]);
    if (first) {
      for (int r = 0; r < 10; r++) { K[r] = f[r]; Kp[r] = f[r]; }
      first = false;
    } else {
      for (int r = 0; r < 10; r++) {
        const double kc = K[r];
        if (!second) K[r] = 2.0 * kc - Kp[r];
        Kp[r] = kc;
      }
      second = false;
    }
    int it = 0;
    double dprev = 1e300;
#pragma unroll 1
    for (;; it++) {
      double xs[10];
      for (int r = 0; r < 10; r++) {
        double v = x[r];
#pragma unroll
        for (int j = 0; j < 4; j++) v = fma(Ai[j], __shfl_sync(gm, K[r], j, 4), v);
        xs[r] = v;
      }
      fdyn10_rhs(xs, tau_i, dl_i, p, tr, f, nullptr);
      double dmax = 0.0, kmax = 0.0;
      for (int r = 0; r < 6; r++) { dmax = fmax(dmax, fabs(f[r] - K[r])); kmax = fmax(kmax, fabs(f[r])); K[r] = f[r]; }
      for (int q = 0; q < 4; q++) {
        const double d = f[6 + q] - K[6 + q];
        double upd = 0.0;
#pragma unroll
        for (int j = 0; j < 4; j++) upd = fma(inv[q][j], __shfl_sync(gm, d, j, 4), upd);
        dmax = fmax(dmax, fabs(d)); kmax = fmax(kmax, fabs(f[6 + q]));
        K[6 + q] += upd;
      }
      if (!(dmax == dmax)) dmax = 1e300;
      dmax = fmax(dmax, __shfl_xor_sync(gm, dmax, 1)); dmax = fmax(dmax, __shfl_xor_sync(gm, dmax, 2));
      kmax = fmax(kmax, __shfl_xor_sync(gm, kmax, 1)); kmax = fmax(kmax, __shfl_xor_sync(gm, kmax, 2));
      if (dmax <= FP_TOL * (1.0 + kmax) || (it >= 3 && dmax >= 0.5 * dprev && dmax <= 1e-10 * (1.0 + kmax))) break;
      dprev = dmax;
      if (it >= MAX_FP || dmax >= 1e300) { ok = false; break; }
    }
    for (int r = 0; r < 10; r++) {
      double v = x[r];
#pragma unroll
      for (int j = 0; j < 4; j++) v = fma(bw[j], __shfl_sync(gm, K[r], j, 4), v);
      x[r] = v;
    }
    for (int q = 0; q < 4; q++) x[10 + q] += dtau[q];
    x[14] += ddl;
    if (!ok) break;
  }
  if (i == 0)
    for (int r = 0; r < IHM2B_NX10; r++) xout[(size_t)b * IHM2B_NX10 + r] = ok ? x[r] : nan("");
}

void launch_sim_step_dyn10(const ihm2b_config& cfg, const KCfg& cf, const Work& w, const double* x, const double* u, double* xn,
                           cudaStream_t stream) {
  SimCfg sc;
  sc.B = cf.B; sc.n_grid = cf.n_grid; sc.model = 0; sc.n_sub = cfg.sim_steps;
  sc.h = cfg.dt / cfg.sim_steps; sc.sw = 0.0;
  for (int i = 0; i < 16; i++) sc.A[i] = cfg.sim_A[i];
  for (int i = 0; i < 4; i++) sc.b[i] = cfg.sim_b[i];
  const int threads = 128, total = cf.B * 4;
  sim_step_dyn10_kernel<<<(total + threads - 1) / threads, threads, 0, stream>>>(sc, w, x, u, xn);
}

// With a model switch the eight instances of a warp would mix the kinematic and the (several times longer) dynamic
// right-hand side and every warp would run both; this puts the instances that take the kinematic model first and the
// others last (order within a class is whatever the atomics give: results do not depend on it).
__global__ void sim_partition_kernel(int B, int mode, double sw, const double* __restrict__ params, const double* __restrict__ x,
                                     const int32_t* __restrict__ active, int32_t* __restrict__ counters, int32_t* __restrict__ order) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  bool kin;
  if (mode == IHM2B_SIM_KIN6_DYN6) {
    const double lR = params[(size_t)b * NP + IHM2B_P_LR], C = lR / (lR + params[(size_t)b * NP + IHM2B_P_LF]);
    const double beta = atan(C * tan(x[(size_t)b * NX + 7]));
    const double vx = x[(size_t)b * NX + 3], vy = x[(size_t)b * NX + 4];
    kin = (vx * vx + vy * vy) * sin(beta) / lR <= sw;
  } else {
    kin = hypot(x[(size_t)b * NX + 3], x[(size_t)b * NX + 4]) < sw;
  }
  if (active && !active[b]) kin = true;   // idle instances cost nothing wherever they sit
  if (kin) order[atomicAdd(&counters[0], 1)] = b;
  else order[B - 1 - atomicAdd(&counters[1], 1)] = b;
}

void launch_sim_step(const ihm2b_config& cfg, const KCfg& cf, const Work& w, const double* x, const double* u,
                     double* xn, int32_t* model, int32_t* order_scratch, cudaStream_t stream) {
  SimCfg sc;
  sc.B = cf.B; sc.n_grid = cf.n_grid; sc.model = cfg.sim_model; sc.n_sub = cfg.sim_steps > 0 ? cfg.sim_steps : 1;
  sc.h = cfg.dt / sc.n_sub; sc.sw = cfg.sim_switch;
  for (int i = 0; i < 16; i++) sc.A[i] = cfg.sim_A[i];
  for (int i = 0; i < 4; i++) sc.b[i] = cfg.sim_b[i];
  const int threads = 128, total = cf.B * 4;
  const int32_t* order = nullptr;
  const bool switching = cfg.sim_model == IHM2B_SIM_KIN6_DYN6 || cfg.sim_model == IHM2B_SIM_KIN6C_DYN6C;
  if (switching && order_scratch && cf.B > 8) {   // order_scratch: [B] order + 2 counters
    cudaMemsetAsync(order_scratch + cf.B, 0, 2 * sizeof(int32_t), stream);
    sim_partition_kernel<<<(cf.B + 255) / 256, 256, 0, stream>>>(cf.B, cfg.sim_model, cfg.sim_switch, w.params, x, w.active,
                                                                 order_scratch + cf.B, order_scratch);
    order = order_scratch;
  }
  sim_step_kernel<<<(total + threads - 1) / threads, threads, 0, stream>>>(sc, w, x, u, xn, model, order);
}

// ---- plant step with its scratch in static device memory (ihm2b_sim_step_static_dev) -------------------------------
// For callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): the
// [B] launch order + 2 counters of the partition above live in a file-scope array instead of a buffer the handle owns.
// Every GPU that loads the library holds the array (SIM_STATIC_MAX_B + 2 int32 = 256 KB) until the process exits; all
// calls on a device share it, so calls that overlap on the GPU (different streams) race -- serialise them.  Nothing in
// it has to survive a call: the counters are cleared and order[0 .. B) is rewritten by every call.
constexpr int SIM_STATIC_MAX_B = 65536;
static __device__ __align__(256) int32_t g_sim_static_scratch[SIM_STATIC_MAX_B + 2];

int sim_static_max_batch() { return SIM_STATIC_MAX_B; }

// Device address of the array on the current device.  In host code the array's name is only a stand-in for the symbol:
// the address a kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current.
static cudaError_t sim_static_scratch(int32_t** out) {
  static std::mutex mtx;
  static int32_t* addr[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mtx);
  if (!addr[dev]) {
    int32_t* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_sim_static_scratch);
    if (e != cudaSuccess) return e;
    addr[dev] = p;
  }
  *out = addr[dev];
  return cudaSuccess;
}

// Same launches as launch_sim_step; the partition runs for every batch size of a switching model (also B <= 8), so the
// scratch is what the call uses.  No allocation: cudaGetSymbolAddress, cudaMemsetAsync and kernel launches only.
cudaError_t launch_sim_step_static(const ihm2b_config& cfg, const KCfg& cf, const Work& w, const double* x, const double* u,
                                   double* xn, int32_t* model, cudaStream_t stream) {
  if (cf.B > SIM_STATIC_MAX_B) return cudaErrorInvalidValue;
  int32_t* scratch = nullptr;
  cudaError_t e = sim_static_scratch(&scratch);
  if (e != cudaSuccess) return e;
  SimCfg sc;
  sc.B = cf.B; sc.n_grid = cf.n_grid; sc.model = cfg.sim_model; sc.n_sub = cfg.sim_steps > 0 ? cfg.sim_steps : 1;
  sc.h = cfg.dt / sc.n_sub; sc.sw = cfg.sim_switch;
  for (int i = 0; i < 16; i++) sc.A[i] = cfg.sim_A[i];
  for (int i = 0; i < 4; i++) sc.b[i] = cfg.sim_b[i];
  const int threads = 128, total = cf.B * 4;
  const int32_t* order = nullptr;
  if (cfg.sim_model == IHM2B_SIM_KIN6_DYN6 || cfg.sim_model == IHM2B_SIM_KIN6C_DYN6C) {
    e = cudaMemsetAsync(scratch + cf.B, 0, 2 * sizeof(int32_t), stream);   // the two counters start at zero
    if (e != cudaSuccess) return e;
    sim_partition_kernel<<<(cf.B + 255) / 256, 256, 0, stream>>>(cf.B, cfg.sim_model, cfg.sim_switch, w.params, x, w.active,
                                                                 scratch + cf.B, scratch);
    order = scratch;
  }
  sim_step_kernel<<<(total + threads - 1) / threads, threads, 0, stream>>>(sc, w, x, u, xn, model, order);
  return cudaGetLastError();
}

// closed-loop bookkeeping after solve + plant step (python/main.py:503-517): an instance stops at
// its first solver failure (status not in {0, 2}) or NaN in the plant (state kept), and is done
// when s > lap_length + 1.
// Lap time as python/main.py:561-570 computes it from the state log: dt x (second - first index at which
// mod(s, lap_length) jumps by more than 1 between consecutive states); NaN until two jumps have been seen.
// cost_sum: OCP cost at the solution (stats[7]) summed over the control steps taken.
__global__ void closed_loop_post_kernel(int B, int step, double dt, const int32_t* __restrict__ track, const double* __restrict__ lap,
                                        const int32_t* __restrict__ status, const double* __restrict__ u0,
                                        const double* __restrict__ xn, const double* __restrict__ stats,
                                        double* __restrict__ xs, double* __restrict__ xlog,
                                        double* __restrict__ ulog, int32_t* __restrict__ done, int32_t* __restrict__ fail,
                                        int32_t* __restrict__ active_mask, int32_t* __restrict__ lap_state,
                                        double* __restrict__ lap_time, double* __restrict__ cost_sum) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  bool active = done[b] < 0 && fail[b] < 0;
  double uu0 = 0.0, uu1 = 0.0;
  if (active) {
    const int st = status[b];
    if (st != IHM2B_SUCCESS && st != IHM2B_MAXITER) { fail[b] = step; active = false; }
    else { uu0 = u0[b * 2]; uu1 = u0[b * 2 + 1]; if (cost_sum) cost_sum[b] += stats[(size_t)b * 8 + 7]; }
  }
  if (active) {
    bool isnan_ = false;
    for (int r = 0; r < 8; r++) isnan_ |= !(xn[b * 8 + r] == xn[b * 8 + r]);
    if (isnan_) fail[b] = step;
    else {
      const double L = lap[track[b]], sa = xs[b * 8], sb = xn[b * 8];
      const double m0 = sa - L * floor(sa / L), m1 = sb - L * floor(sb / L);   // np.mod
      if (fabs(m1 - m0) > 1.0) {
        const int jumps = lap_state[b * 2];
        if (jumps == 0) lap_state[b * 2 + 1] = step;
        else if (jumps == 1 && lap_time) lap_time[b] = dt * (step - lap_state[b * 2 + 1]);
        lap_state[b * 2] = jumps + 1;
      }
      for (int r = 0; r < 8; r++) xs[b * 8 + r] = xn[b * 8 + r];
      if (xn[b * 8] > lap[track[b]] + 1.0) done[b] = step;
    }
  }
  if (active_mask && (done[b] >= 0 || fail[b] >= 0)) active_mask[b] = 0;   // later steps skip this instance
  if (ulog) { ulog[((size_t)step * B + b) * 2] = uu0; ulog[((size_t)step * B + b) * 2 + 1] = uu1; }
  if (xlog) for (int r = 0; r < 8; r++) xlog[((size_t)(step + 1) * B + b) * 8 + r] = xs[b * 8 + r];
}

void launch_closed_loop_post(int B, int step, double dt, const int32_t* track, const double* lap, const int32_t* status,
                             const double* u0, const double* xn, const double* stats, double* xs, double* xlog,
                             double* ulog, int32_t* done, int32_t* fail, int32_t* active_mask, int32_t* lap_state,
                             double* lap_time, double* cost_sum, cudaStream_t stream) {
  closed_loop_post_kernel<<<(B + 127) / 128, 128, 0, stream>>>(B, step, dt, track, lap, status, u0, xn, stats, xs, xlog, ulog,
                                                               done, fail, active_mask, lap_state, lap_time, cost_sum);
}

}  // namespace ihm2b
