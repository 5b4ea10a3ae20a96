// qp.cu -- batched structured QP: Mehrotra primal-dual interior point with a Riccati KKT
// solve, followed by the SQP-RTI full step.
//
// Replaces HPIPM's d_ocp_qp_ipm_solve (selected by the reference through
// qp_solver = PARTIAL_CONDENSING_HPIPM, new_python/controller.py:300; implementation not in
// the reference tree) and the feedback phase of acados' ocp_nlp_sqp_rti for the OCP of
// python/mpc.py:29-101 (LINEAR_LS cost, box bounds idxbx/idxbu, two rate rows C x + D u).
//
// Mapping: ONE WARP PER PROBLEM, the whole horizon resident in shared memory (one record
// of RS doubles per stage: [B A], c, multipliers/slacks, iterate, costate, Riccati gains,
// candidate point), 3 problems per SM.  Lane l = (a = l & 7, q = l >> 3): a is the row of
// every 8-row object (state x_a, row a of the cost-to-go matrix P, row a of [B A], inequality
// row a); the four quarters q split the columns of the two dense products of the Riccati
// step and otherwise compute redundantly, so that every reduction is a shuffle inside an
// 8-lane segment and every hand-over goes through a few hundred bytes of scratch.
//
// Formulation (variables z_k = (u_k, x_k), x_0 fixed to the measured state):
//   min sum 1/2 z'Hz + g'z   s.t.  x_{k+1} = A x_k + B u_k + c_k,   lb <= C z <= ub
// Residual ("delta") form of the Newton system, as in the CPU oracle: each iteration
// recomputes the stationarity / dynamics / slack residuals at the current point, factorises
// H + C' diag(lam/t) C once by a backward Riccati sweep, and solves predictor and corrector
// with that factor.  Equality multipliers are advanced through the costate recursion of the
// full Newton step, which needs no stored cost-to-go matrices.
#include "common.cuh"

namespace ihm2b {

namespace {

constexpr double TAU = 0.995, THR0 = 0.1;
constexpr unsigned FULL = 0xffffffffu;

// stage record (doubles)
constexpr int R_BA = 0;     // [8][10]  (B | A) row-major
constexpr int R_C = 80;     // [8]
constexpr int R_LT = 88;    // lamL[8] lamU[8] tL[8] tU[8]
constexpr int R_X = 120;    // [8]
constexpr int R_U = 128;    // [2]
constexpr int R_PI = 130;   // [8]
constexpr int R_K = 138;    // [8][2]
constexpr int R_KF = 154;   // kff[2] kfc[2]
constexpr int R_MI = 158;   // i00 i01 i11 pad
constexpr int R_D = 162;    // dlam_aff*dt_aff: L[8] U[8]
constexpr int R_XA = 178;   // [8]  step dx of the current solve (affine, then corrected)
constexpr int R_UA = 186;   // [2]  step du
constexpr int R_RB = 188;   // [8]  dynamics residual  A x + B u + c - x_next
constexpr int R_RG = 196;   // [8]  stationarity residual, state part
constexpr int RS = 204;
// scratch (doubles, after the N+1 records)
constexpr int S_W = 0;      // [8][10]
constexpr int S_M = 80;     // [8][10]  rows 2..9 of M
constexpr int S_P = 160;    // [8][8]
constexpr int S_V = 224;    // [8]
constexpr int S_RB = 232;   // [8]
constexpr int S_G = 240;    // gamma[8] psi[8] dl[8]
constexpr int S_MU = 264;   // M00 M01 M11 mvu0 mvu1 pad[3]
constexpr int S_TOT = 272;

__device__ __forceinline__ double gsum(double v) {
  v += __shfl_xor_sync(FULL, v, 1);
  v += __shfl_xor_sync(FULL, v, 2);
  v += __shfl_xor_sync(FULL, v, 4);
  return v;
}
// NaN-propagating max
__device__ __forceinline__ double nmax(double a, double b) { return (b > a || b != b) ? b : a; }
__device__ __forceinline__ double gmaxr(double v) {
  v = nmax(v, __shfl_xor_sync(FULL, v, 1));
  v = nmax(v, __shfl_xor_sync(FULL, v, 2));
  v = nmax(v, __shfl_xor_sync(FULL, v, 4));
  return v;
}

// Inequality row r of a stage of type ty (0: k = 0, 1: 0 < k < N, 2: k = N):
//   v = (iu >= 0 ? u[iu] : 0) + sx * x[ix]   in [lb, ub]
// python/mpc.py:77-99; stage 0 carries no state bounds (x_0 is fixed); the terminal set is
// idxbx_e = [1,3,4,5] with the n / v_x / T / delta limits (literal, mpc.py:81-83).
struct Row { int iu, ix; double sx, lb, ub; bool act; };

__device__ __forceinline__ Row row_def(const KCfg& cf, int ty, int r) {
  Row R;
  R.iu = -1; R.ix = -1; R.sx = 0.0; R.lb = -1.0; R.ub = 1.0; R.act = false;
  if (ty < 2) {
    if (r < 4) {
      R.act = true;
      R.iu = r & 1;
      if (r >= 2) { R.ix = 6 + (r & 1); R.sx = -1.0; }
      const double bd = (r == 0) ? cf.T_max : (r == 1) ? cf.delta_max : (r == 2) ? cf.g_T_max : cf.g_delta_max;
      R.lb = -bd; R.ub = bd;
    } else if (ty == 1) {
      R.act = true;
      R.sx = 1.0;
      R.ix = (r == 4) ? 1 : (r == 5) ? 3 : (r == 6) ? 6 : 7;
      const double bd = (r == 4) ? cf.n_max : (r == 5) ? cf.v_x_max : (r == 6) ? cf.T_max : cf.delta_max;
      R.lb = (r == 5) ? 0.0 : -bd; R.ub = bd;
    }
  } else if (r < 4) {
    R.act = true;
    R.sx = 1.0;
    R.ix = (r == 0) ? 1 : (r == 1) ? 3 : (r == 2) ? 4 : 5;
    const double bd = (r == 0) ? cf.n_max : (r == 1) ? cf.v_x_max : (r == 2) ? cf.T_max : cf.delta_max;
    R.lb = -bd; R.ub = bd;
  }
  return R;
}

// row value from a point held in memory (x[8], u[2])
__device__ __forceinline__ double row_value_mem(const Row& R, const double* x, const double* u) {
  double v = (R.iu >= 0) ? u[R.iu] : 0.0;
  if (R.ix >= 0) v = fma(R.sx, x[R.ix], v);
  return v;
}
// row value from a point held in registers (x_a per lane, u replicated)
__device__ __forceinline__ double row_value_reg(const Row& R, double xa, double u0, double u1) {
  const double xs = __shfl_sync(FULL, xa, R.ix < 0 ? 0 : R.ix, GROUP);
  double v = (R.iu == 0) ? u0 : (R.iu == 1) ? u1 : 0.0;
  if (R.ix >= 0) v = fma(R.sx, xs, v);
  return v;
}

// What state lane a receives from C' q (q_r in sq[0..7]).
//   ty < 2: rows (u_T, u_d, u_T - T, u_d - delta, n, v_x, T, delta);  ty = 2: rows (n, v_x, v_y, r)
__device__ __forceinline__ double ct_state(const double* sq, int ty, int a) {
  if (ty < 2) {
    switch (a) {
      case 1: return sq[4];
      case 3: return sq[5];
      case 6: return sq[6] - sq[2];
      case 7: return sq[7] - sq[3];
      default: return 0.0;
    }
  }
  switch (a) {
    case 1: return sq[0];
    case 3: return sq[1];
    case 4: return sq[2];
    case 5: return sq[3];
    default: return 0.0;
  }
}
// diagonal of C' diag(q) C seen by state lane a
__device__ __forceinline__ double ctc_diag(const double* sq, int ty, int a) {
  if (ty < 2) {
    switch (a) {
      case 1: return sq[4];
      case 3: return sq[5];
      case 6: return sq[6] + sq[2];
      case 7: return sq[7] + sq[3];
      default: return 0.0;
    }
  }
  return ct_state(sq, ty, a);
}

}  // namespace

// ---------------------------------------------------------------------------------------
// grid = B, block = 32, dynamic shared memory = ((N+1) * RS + S_TOT) doubles.
// x0 [B][8]; on exit the iterate w.x / w.u holds the QP solution (RTI full step).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) qp_rti_kernel(const KCfg cf, const Work w,
                                                     const double* __restrict__ x0,
                                                     double* __restrict__ u0_out,
                                                     int32_t* __restrict__ status_out) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x;
  const int a = lane & 7, q = lane >> 3;
  const int b = blockIdx.x;
  const int N = cf.N;
  double* const sc = smem + (N + 1) * RS;
  double* const sW = sc + S_W; double* const sM = sc + S_M; double* const sP = sc + S_P;
  double* const sv = sc + S_V; double* const srb = sc + S_RB; double* const sG = sc + S_G;
  double* const sMu = sc + S_MU;
#define REC(k) (smem + (k) * RS)

  // ------------------------------------------------------------------ load the problem
  {
    const double2* gBA = reinterpret_cast<const double2*>(w.BA + (size_t)b * N * 80);
    for (int e = lane; e < N * 40; e += 32) {
      const int k = e / 40, o = e - k * 40;
      reinterpret_cast<double2*>(REC(k) + R_BA)[o] = gBA[e];
    }
    const double* gc = w.c + (size_t)b * N * 8;
    for (int e = lane; e < N * 8; e += 32) REC(e >> 3)[R_C + (e & 7)] = gc[e];
    const double* gx = w.x + (size_t)b * (N + 1) * 8;
    for (int e = lane; e < (N + 1) * 8; e += 32) REC(e >> 3)[R_X + (e & 7)] = gx[e];
    const double* gu = w.u + (size_t)b * N * 2;
    for (int e = lane; e < N * 2; e += 32) REC(e >> 1)[R_U + (e & 1)] = gu[e];
  }
  const double xh = x0[(size_t)b * 8 + a];  // measured state, component a
  const double s0 = __shfl_sync(FULL, xh, 0, GROUP);
  __syncwarp();
  if (q == 0) REC(0)[R_X + a] = xh;         // x_0 of the iterate is the measured state
  if (lane < 2) REC(N)[R_U + lane] = 0.0;
  // cost (python/mpc.py:47-69, controller.py:320-385): constant near-diagonal GN Hessian
  const double hx = cf.cs * (cf.W[a] + (a == 6 ? cf.W[10] : a == 7 ? cf.W[11] : 0.0));
  const double hxN = cf.ct * cf.We[a];
  const double hu0 = cf.cs * (cf.W[8] + cf.W[10]), hu1 = cf.cs * (cf.W[9] + cf.W[11]);
  const double c10 = cf.cs * cf.W[10], c11 = cf.cs * cf.W[11];
  const double hxu0 = (a == 6) ? -c10 : 0.0;  // H[x_a][u_T]
  const double hxu1 = (a == 7) ? -c11 : 0.0;  // H[x_a][u_delta]
  const double gs_scale = -cf.cs * cf.W[0];   // g[s]_k = gs_scale * yref_k
  const double gsN = -cf.ct * cf.We[0] * (s0 + cf.s_target);
  __syncwarp();

  // ------------------------------------------------------------------ initialisation
  // slacks / multipliers cold (t = max(distance, thr0), lam = mu0 / t), pi = 0; 4 stages at a time
  for (int k = q; k <= N; k += 4) {
    const int ty = (k == 0) ? 0 : (k < N ? 1 : 2);
    double* r = REC(k);
    const Row R = row_def(cf, ty, a);
    const double v = row_value_mem(R, r + R_X, r + R_U);
    double lamL = 0.0, lamU = 0.0, tL = 1.0, tU = 1.0;
    if (R.act) {
      tL = fmax(v - R.lb, THR0); tU = fmax(R.ub - v, THR0);
      lamL = cf.mu0 / tL; lamU = cf.mu0 / tU;
    }
    r[R_LT + a] = lamL; r[R_LT + 8 + a] = lamU; r[R_LT + 16 + a] = tL; r[R_LT + 24 + a] = tU;
    r[R_PI + a] = 0.0;
  }
  __syncwarp();
  const double inv_n = 1.0 / (16.0 * N);  // one-sided inequalities: 2 * (4 + 8 (N-1) + 4)
  double mu = 0.0, res_g = 0.0, res_b = 0.0, res_d = 0.0, res_m = 0.0;
  double mag_g = 0.0, mag_b = 0.0, mag_d = 0.0;
  int it = 0, qp_status = 1;

  // ------------------------------------------------------------------ IPM iterations
  const int b0 = (q < 2) ? 3 * q : 2 * q + 2;              // first column of this quarter
  const int c0 = b0, c1 = b0 + 1, c2 = (q < 2) ? b0 + 2 : b0 + 1;  // (q >= 2: third column is a duplicate)
  while (true) {
    // ================= backward sweep 1: residuals, Riccati factorisation, predictor rhs
    double rg_m = 0.0, rb_m = 0.0, rd_m = 0.0, rm_m = 0.0, S0 = 0.0;
    double pa;
    {
      double* r = REC(N);
      const Row R = row_def(cf, 2, a);
      const double v = row_value_mem(R, r + R_X, r + R_U);
      double gam = 0.0, psi = 0.0, dl = 0.0;
      if (R.act) {
        const double lamL = r[R_LT + a], lamU = r[R_LT + 8 + a], tL = r[R_LT + 16 + a], tU = r[R_LT + 24 + a];
        const double itL = 1.0 / tL, itU = 1.0 / tU;
        const double rdL = v - R.lb - tL, rdU = R.ub - v - tU;
        gam = lamL * itL + lamU * itU;
        psi = lamL * rdL * itL - lamU * rdU * itU;
        dl = lamL - lamU;
        rd_m = nmax(fabs(rdL), fabs(rdU));
        rm_m = nmax(lamL * tL, lamU * tU);
        S0 = lamL * tL + lamU * tU;
        mag_d = fmax(mag_d, fmax(fabs(R.lb), fabs(R.ub)));
      }
      if (q == 0) { sG[a] = gam; sG[8 + a] = psi; sG[16 + a] = dl; }
      __syncwarp();
      const double gx = (a == 0) ? gsN : 0.0;
      mag_g = fmax(mag_g, fabs(gx));
      const double base = hxN * r[R_X + a] + gx - r[R_PI + a];
      const double rgx = base - ct_state(sG + 16, 2, a);
      rg_m = fabs(rgx);
      if (q == 0) r[R_RG + a] = rgx;
      pa = base + ct_state(sG + 8, 2, a);
      const double pd = hxN + ctc_diag(sG, 2, a);
      sP[a * 8 + 2 * q] = (2 * q == a) ? pd : 0.0;
      sP[a * 8 + 2 * q + 1] = (2 * q + 1 == a) ? pd : 0.0;
      __syncwarp();
    }
#pragma unroll 1
    for (int k = N - 1; k >= 0; k--) {
      const int ty = (k == 0) ? 0 : 1;
      double* r = REC(k);
      const double* rn = REC(k + 1);
      // ---- S1: residuals of this stage
      {
        const Row R = row_def(cf, ty, a);
        const double v = row_value_mem(R, r + R_X, r + R_U);
        double gam = 0.0, psi = 0.0, dl = 0.0;
        if (R.act) {
          const double lamL = r[R_LT + a], lamU = r[R_LT + 8 + a], tL = r[R_LT + 16 + a], tU = r[R_LT + 24 + a];
          const double itL = 1.0 / tL, itU = 1.0 / tU;
          const double rdL = v - R.lb - tL, rdU = R.ub - v - tU;
          gam = lamL * itL + lamU * itU;
          psi = lamL * rdL * itL - lamU * rdU * itU;
          dl = lamL - lamU;
          rd_m = nmax(rd_m, nmax(fabs(rdL), fabs(rdU)));
          rm_m = nmax(rm_m, nmax(lamL * tL, lamU * tU));
          S0 += lamL * tL + lamU * tU;
          mag_d = fmax(mag_d, fmax(fabs(R.lb), fabs(R.ub)));
        }
        if (q == 0) { sG[a] = gam; sG[8 + a] = psi; sG[16 + a] = dl; }
      }
      const double u0 = r[R_U], u1 = r[R_U + 1];
      double atp = 0.0, btp0 = 0.0, btp1 = 0.0;
      {
        const double2* row = reinterpret_cast<const double2*>(r + R_BA + a * 10);
        const double2* xx = reinterpret_cast<const double2*>(r + R_X);
        double acc = r[R_C + a];
        mag_b = fmax(mag_b, fabs(acc));
        const double2 bb = row[0];
        acc = fma(bb.x, u0, acc); acc = fma(bb.y, u1, acc);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const double2 aa = row[1 + j], xj = xx[j];
          acc = fma(aa.x, xj.x, acc); acc = fma(aa.y, xj.y, acc);
        }
        const double rb = acc - rn[R_X + a];
        rb_m = nmax(rb_m, fabs(rb));
        if (q == 0) { srb[a] = rb; r[R_RB + a] = rb; }
#pragma unroll
        for (int i = 0; i < 8; i++) {
          const double pn = rn[R_PI + i];
          const double2 b2 = *reinterpret_cast<const double2*>(r + R_BA + i * 10);
          atp = fma(r[R_BA + i * 10 + 2 + a], pn, atp);
          btp0 = fma(b2.x, pn, btp0); btp1 = fma(b2.y, pn, btp1);
        }
      }
      const double gx = (a == 0) ? gs_scale * (s0 + cf.s_target * k / N) : 0.0;
      mag_g = fmax(mag_g, fabs(gx));
      const double base_x = hx * r[R_X + a] + hxu0 * u0 + hxu1 * u1 + gx + atp - r[R_PI + a];
      const double base_u0 = hu0 * u0 - c10 * r[R_X + 6] + btp0;
      const double base_u1 = hu1 * u1 - c11 * r[R_X + 7] + btp1;
      __syncwarp();  // (1) sG, srb, sP visible
      {
        const double rgx = base_x - ct_state(sG + 16, ty, a);
        if (k > 0) rg_m = nmax(rg_m, fabs(rgx));
        if (q == 0) r[R_RG + a] = rgx;
      }
      rg_m = nmax(rg_m, nmax(fabs(base_u0 - (sG[16] + sG[18])), fabs(base_u1 - (sG[17] + sG[19]))));
      const double gx_t = base_x + ct_state(sG + 8, ty, a);
      const double gu0_t = base_u0 + sG[8] + sG[10], gu1_t = base_u1 + sG[9] + sG[11];
      const double gam0 = sG[0], gam1 = sG[1], gam2 = sG[2], gam3 = sG[3];
      const double dgam = hx + ctc_diag(sG, ty, a);
      // ---- S2: W = P [B A] (columns of this quarter), v = P rb + p
      {
        double Pr[8];
        const double2* pr = reinterpret_cast<const double2*>(sP + a * 8);
#pragma unroll
        for (int j = 0; j < 4; j++) { const double2 t = pr[j]; Pr[2 * j] = t.x; Pr[2 * j + 1] = t.y; }
        double W0 = 0.0, W1 = 0.0, W2 = 0.0, va = pa;
#pragma unroll
        for (int j = 0; j < 8; j++) {
          const double* bj = r + R_BA + j * 10;
          W0 = fma(Pr[j], bj[c0], W0); W1 = fma(Pr[j], bj[c1], W1); W2 = fma(Pr[j], bj[c2], W2);
          va = fma(Pr[j], srb[j], va);
        }
        sW[a * 10 + c0] = W0; sW[a * 10 + c1] = W1;
        if (q < 2) sW[a * 10 + c2] = W2;
        if (q == 0) sv[a] = va;
      }
      __syncwarp();  // (2)
      // ---- S3: M = H~ + [B A]' W (rows 2.., columns of this quarter), uu block split by quarter
      double mvx = gx_t;
      {
        double M0 = 0.0, M1 = 0.0, M2 = 0.0, e0 = 0.0, e1 = 0.0;
        const int ub = (q == 2) ? 1 : 0;            // B column used by this quarter's uu entry
        const int uw = (q == 0) ? 0 : 1;            // W column
#pragma unroll
        for (int i = 0; i < 8; i++) {
          const double* bi = r + R_BA + i * 10;
          const double* wi = sW + i * 10;
          const double col = bi[2 + a], vi = sv[i];
          M0 = fma(col, wi[c0], M0); M1 = fma(col, wi[c1], M1); M2 = fma(col, wi[c2], M2);
          mvx = fma(col, vi, mvx);
          if (q < 3) e0 = fma(bi[ub], wi[uw], e0);
          else { e0 = fma(bi[0], vi, e0); e1 = fma(bi[1], vi, e1); }
        }
        // add H + C' Gamma C
        if (c0 == 2 + a) M0 += dgam;
        if (c1 == 2 + a) M1 += dgam;
        if (c2 == 2 + a) M2 += dgam;
        if (q == 0) { M0 += hxu0 - ((a == 6) ? gam2 : 0.0); M1 += hxu1 - ((a == 7) ? gam3 : 0.0); }
        sM[a * 10 + c0] = M0; sM[a * 10 + c1] = M1;
        if (q < 2) sM[a * 10 + c2] = M2;
        if (a == 0) {
          if (q < 3) sMu[q] = e0;
          else { sMu[3] = e0; sMu[4] = e1; }
        }
      }
      __syncwarp();  // (3)
      // ---- S4: eliminate u, update P and p
      {
        const double M00 = sMu[0] + hu0 + gam0 + gam2, M01 = sMu[1], M11 = sMu[2] + hu1 + gam1 + gam3;
        const double mvu0 = sMu[3] + gu0_t, mvu1 = sMu[4] + gu1_t;
        const double idet = 1.0 / (M00 * M11 - M01 * M01);
        const double i00 = M11 * idet, i01 = -M01 * idet, i11 = M00 * idet;
        const double Mx0 = sM[a * 10], Mx1 = sM[a * 10 + 1];
        const double K0 = -(i00 * Mx0 + i01 * Mx1), K1 = -(i01 * Mx0 + i11 * Mx1);
        pa = mvx + K0 * mvu0 + K1 * mvu1;
        if (q == 0) *reinterpret_cast<double2*>(r + R_K + 2 * a) = make_double2(K0, K1);
        if (lane == 0) {
          r[R_KF] = -(i00 * mvu0 + i01 * mvu1); r[R_KF + 1] = -(i01 * mvu0 + i11 * mvu1);
          r[R_MI] = i00; r[R_MI + 1] = i01; r[R_MI + 2] = i11;
        }
#pragma unroll
        for (int t = 0; t < 2; t++) {
          const int j = 2 * q + t;
          const double Mj0 = sM[j * 10], Mj1 = sM[j * 10 + 1];
          const double K0j = -(i00 * Mj0 + i01 * Mj1), K1j = -(i01 * Mj0 + i11 * Mj1);
          sP[a * 8 + j] = fma(Mx0, K0j, fma(Mx1, K1j, sM[a * 10 + 2 + j]));
        }
      }
    }
    __syncwarp();
    res_g = gmaxr(rg_m); res_b = gmaxr(rb_m); res_d = gmaxr(rd_m); res_m = gmaxr(rm_m);
    mag_g = gmaxr(mag_g); mag_b = gmaxr(mag_b); mag_d = gmaxr(mag_d);
    S0 = gsum(S0);
    mu = S0 * inv_n;
    if (w.dbg && b == 0 && lane == 0 && it < 64) {
      double* d = w.dbg + it * 8;
      d[0] = res_g; d[1] = res_b; d[2] = res_d; d[3] = res_m; d[4] = mu;
    }
    if (!(res_g == res_g) || !(res_b == res_b) || !(mu == mu)) { qp_status = 3; break; }
    if (res_g <= cf.tol * (1.0 + mag_g) && res_b <= cf.tol * (1.0 + mag_b) && res_d <= cf.tol * (1.0 + mag_d) &&
        res_m <= cf.tol_comp) { qp_status = 0; break; }
    if (it >= cf.qp_max_iter) break;

    // ================= forward sweeps (pred = 1: affine step; pred = 0: corrected step)
    double alpha = 1.0, mu_t = 0.0;
    bool bad = false;
    for (int pred = 1; pred >= 0; pred--) {
      double ratio = 0.0, S1 = 0.0, S2 = 0.0;
      double dx = 0.0;  // component a of the state step (x_0 is fixed)
#pragma unroll 1
      for (int k = 0; k <= N; k++) {
        const int ty = (k == 0) ? 0 : (k < N ? 1 : 2);
        double* r = REC(k);
        double u0n = 0.0, u1n = 0.0;
        if (k < N) {
          const double2 Kc = *reinterpret_cast<const double2*>(r + R_K + 2 * a);
          u0n = gsum(Kc.x * dx) + r[R_KF + (pred ? 0 : 2)];
          u1n = gsum(Kc.y * dx) + r[R_KF + (pred ? 1 : 3)];
        }
        {
          const Row R = row_def(cf, ty, a);
          const double cdz = row_value_reg(R, dx, u0n, u1n);
          if (R.act) {
            const double lamL = r[R_LT + a], lamU = r[R_LT + 8 + a], tL = r[R_LT + 16 + a], tU = r[R_LT + 24 + a];
            const double itL = 1.0 / tL, itU = 1.0 / tU;
            const double v = row_value_mem(R, r + R_X, r + R_U);
            const double dtL = cdz + (v - R.lb - tL), dtU = -cdz + (R.ub - v - tU);
            double dlL, dlU;
            if (pred) {
              dlL = -lamL - lamL * dtL * itL; dlU = -lamU - lamU * dtU * itU;
              S1 += lamL * dtL + tL * dlL + lamU * dtU + tU * dlU;
              S2 += dlL * dtL + dlU * dtU;
              if (q == 0) { r[R_D + a] = dlL * dtL; r[R_D + 8 + a] = dlU * dtU; }
            } else {
              dlL = -lamL - (r[R_D + a] - mu_t + lamL * dtL) * itL;
              dlU = -lamU - (r[R_D + 8 + a] - mu_t + lamU * dtU) * itU;
            }
            ratio = nmax(ratio, nmax(nmax(-dlL / lamL, -dlU / lamU), nmax(-dtL * itL, -dtU * itU)));
          }
        }
        if (q == 0) r[R_XA + a] = dx;
        if (k < N) {
          if (lane == 0) { r[R_UA] = u0n; r[R_UA + 1] = u1n; }
          const double2* row = reinterpret_cast<const double2*>(r + R_BA + a * 10);
          const double2 bb = row[0];
          double acc = r[R_RB + a];
          acc = fma(bb.x, u0n, acc); acc = fma(bb.y, u1n, acc);
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const double2 aa = row[1 + j];
            acc = fma(aa.x, __shfl_sync(FULL, dx, 2 * j, GROUP), acc);
            acc = fma(aa.y, __shfl_sync(FULL, dx, 2 * j + 1, GROUP), acc);
          }
          dx = acc;
        }
      }
      __syncwarp();
      ratio = gmaxr(ratio);
      if (!(ratio == ratio)) { bad = true; break; }
      if (!pred) {
        alpha = TAU * ((ratio > 1.0) ? 1.0 / ratio : 1.0);
        break;
      }
      const double al_aff = (ratio > 1.0) ? 1.0 / ratio : 1.0;
      S1 = gsum(S1); S2 = gsum(S2);
      const double mu_aff = (S0 + al_aff * (S1 + al_aff * S2)) * inv_n;
      const double sr = mu_aff / mu;
      // centring target, floored so the iterates stay on the central path at ~0.1 tol_comp
      mu_t = fmax(sr * sr * sr * mu, 0.1 * cf.tol_comp);

      // ================= backward sweep 2: corrector right-hand side through the same factor
      double dp;
      {
        const double* r = REC(N);
        const Row R = row_def(cf, 2, a);
        const double chi = R.act ? (r[R_D + a] - mu_t) / r[R_LT + 16 + a] - (r[R_D + 8 + a] - mu_t) / r[R_LT + 24 + a] : 0.0;
        if (q == 0) sG[a] = chi;
        __syncwarp();
        dp = ct_state(sG, 2, a);
        __syncwarp();
      }
#pragma unroll 1
      for (int k = N - 1; k >= 0; k--) {
        const int ty = (k == 0) ? 0 : 1;
        double* r = REC(k);
        const Row R = row_def(cf, ty, a);
        const double chi = R.act ? (r[R_D + a] - mu_t) / r[R_LT + 16 + a] - (r[R_D + 8 + a] - mu_t) / r[R_LT + 24 + a] : 0.0;
        if (q == 0) sG[a] = chi;
        __syncwarp();
        double dmx = ct_state(sG, ty, a);
        double dmu0 = sG[0] + sG[2], dmu1 = sG[1] + sG[3];
        const double2 bb = *reinterpret_cast<const double2*>(r + R_BA + a * 10);
        dmu0 += gsum(bb.x * dp);
        dmu1 += gsum(bb.y * dp);
#pragma unroll
        for (int i = 0; i < 8; i++) dmx = fma(r[R_BA + i * 10 + 2 + a], __shfl_sync(FULL, dp, i, GROUP), dmx);
        const double2 Kc = *reinterpret_cast<const double2*>(r + R_K + 2 * a);
        const double i00 = r[R_MI], i01 = r[R_MI + 1], i11 = r[R_MI + 2];
        if (lane == 0) {
          r[R_KF + 2] = r[R_KF] - (i00 * dmu0 + i01 * dmu1);
          r[R_KF + 3] = r[R_KF + 1] - (i01 * dmu0 + i11 * dmu1);
        }
        dp = dmx + Kc.x * dmu0 + Kc.y * dmu1;
        __syncwarp();
      }
    }
    if (w.dbg && b == 0 && lane == 0 && it < 64) { w.dbg[it * 8 + 5] = alpha; w.dbg[it * 8 + 6] = mu_t; }
    if (bad) { qp_status = 3; break; }
    if (alpha < 1e-12) { qp_status = 2; break; }

    // ================= backward sweep 3: step; costate recursion of the Newton step -> d pi
    {
      double dpi = 0.0;  // d pi_{k+1}[a]
#pragma unroll 1
      for (int k = N; k >= 0; k--) {
        const int ty = (k == 0) ? 0 : (k < N ? 1 : 2);
        double* r = REC(k);
        const Row R = row_def(cf, ty, a);
        const double cdz = row_value_mem(R, r + R_XA, r + R_UA);
        double lamL = 0.0, lamU = 0.0, tL = 1.0, tU = 1.0, dlp = 0.0;
        if (R.act) {
          lamL = r[R_LT + a]; lamU = r[R_LT + 8 + a]; tL = r[R_LT + 16 + a]; tU = r[R_LT + 24 + a];
          const double v = row_value_mem(R, r + R_X, r + R_U);
          const double dtL = cdz + (v - R.lb - tL), dtU = -cdz + (R.ub - v - tU);
          const double dlL = -lamL - (r[R_D + a] - mu_t + lamL * dtL) / tL;
          const double dlU = -lamU - (r[R_D + 8 + a] - mu_t + lamU * dtU) / tU;
          dlp = dlL - dlU;
          lamL = fma(alpha, dlL, lamL); lamU = fma(alpha, dlU, lamU);
          tL = fma(alpha, dtL, tL); tU = fma(alpha, dtU, tU);
        }
        const double xc = r[R_XA + a], xo = r[R_X + a], po = r[R_PI + a], rgx = r[R_RG + a];
        const double ua0 = (k < N) ? r[R_UA] : 0.0, ua1 = (k < N) ? r[R_UA + 1] : 0.0;
        const double uo = (lane < 2) ? r[R_U + lane] : 0.0;
        double atp = 0.0;
        if (k < N) {
#pragma unroll
          for (int i = 0; i < 8; i++) atp = fma(r[R_BA + i * 10 + 2 + a], __shfl_sync(FULL, dpi, i, GROUP), atp);
        }
        __syncwarp();
        if (q == 0) {
          sG[a] = dlp;
          if (R.act) { r[R_LT + a] = lamL; r[R_LT + 8 + a] = lamU; r[R_LT + 16 + a] = tL; r[R_LT + 24 + a] = tU; }
          r[R_X + a] = fma(alpha, xc, xo);
        }
        if (lane < 2 && k < N) r[R_U + lane] = fma(alpha, (lane == 0) ? ua0 : ua1, uo);
        __syncwarp();
        dpi = rgx + ((k < N) ? hx * xc + hxu0 * ua0 + hxu1 * ua1 : hxN * xc) + atp - ct_state(sG, ty, a);
        if (q == 0) r[R_PI + a] = fma(alpha, dpi, po);
        __syncwarp();
      }
    }
    it++;
  }

  // ------------------------------------------------------------------ outputs (RTI full step)
  __syncwarp();
  {
    double* gx = w.x + (size_t)b * (N + 1) * 8;
    for (int e = lane; e < (N + 1) * 8; e += 32) gx[e] = REC(e >> 3)[R_X + (e & 7)];
    double* gu = w.u + (size_t)b * N * 2;
    for (int e = lane; e < N * 2; e += 32) gu[e] = REC(e >> 1)[R_U + (e & 1)];
  }
  if (lane == 0) {
    int st;
    if (qp_status == 3) st = IHM2B_FAILURE;
    else if (qp_status == 0) st = IHM2B_SUCCESS;
    else {
      // accept an IPM that stalled above the tight tolerance but is converged to 1e-6
      const bool loose = res_g <= 1e-6 * (1.0 + mag_g) && res_b <= 1e-6 * (1.0 + mag_b) &&
                         res_d <= 1e-6 * (1.0 + mag_d) && res_m <= 1e-6;
      st = loose ? IHM2B_SUCCESS : IHM2B_QP_FAILURE;
    }
    w.status[b] = st;
    if (status_out) status_out[b] = st;
    double* sp = w.stats + (size_t)b * 8;
    sp[0] = it; sp[1] = 1.0; sp[2] = mu; sp[3] = res_g; sp[4] = res_b; sp[5] = res_d; sp[6] = res_m; sp[7] = 0.0;
  }
  if (u0_out && lane < 2) u0_out[(size_t)b * 2 + lane] = REC(0)[R_U + lane];
#undef REC
}

int qp_smem_bytes(int N) { return ((N + 1) * RS + S_TOT) * (int)sizeof(double); }

int launch_qp_rti(const KCfg& cf, const Work& w, const double* x0, double* u0, int32_t* status,
                  cudaStream_t stream) {
  const int bytes = qp_smem_bytes(cf.N);
  static int configured = 0;
  if (configured < bytes) {
    cudaError_t e = cudaFuncSetAttribute(qp_rti_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return (int)e;
    configured = bytes;
  }
  qp_rti_kernel<<<cf.B, 32, bytes, stream>>>(cf, w, x0, u0, status);
  return 0;
}

}  // namespace ihm2b
