// qp.cu -- batched structured QP: Mehrotra primal-dual interior point with a Riccati KKT
// solve, followed by the SQP-RTI full step.
//
// Replaces HPIPM's d_ocp_qp_ipm_solve (selected by the reference through
// qp_solver = PARTIAL_CONDENSING_HPIPM, new_python/controller.py:300; implementation not in
// the reference tree) and the feedback phase of acados' ocp_nlp_sqp_rti for the OCP of
// python/mpc.py:29-101 (LINEAR_LS cost, box bounds idxbx/idxbu, two rate rows C x + D u).
//
// Mapping: 8 lanes own one problem, lane a = row a of every 8-row object (state x_a, row a
// of the cost-to-go matrix P, row a of [B A], inequality row a).  A warp carries 4 problems,
// groups only ever synchronise among their own 8 lanes.  The stage data streams from global
// memory (contiguous 640 B per stage and problem); what several lanes need goes through a
// small per-group shared-memory scratch.
//
// Formulation (absolute variables z_k = (u_k, x_k), x_0 fixed to the measured state):
//   min sum 1/2 z'Hz + g'z   s.t.  x_{k+1} = A x_k + B u_k + c_k,   lb <= C z <= ub
// Per IPM iteration: one Riccati factorisation of H + C' diag(lam/t) C, predictor solve,
// corrector solve with the same factor, one step length for primal and dual variables.
// The Newton system is solved for the new point z+ directly ("absolute form"), so the linear
// residuals contract exactly by (1 - alpha) per iteration; they are tracked through that
// factor and are not recomputed.
#include "common.cuh"

namespace ihm2b {

namespace {

constexpr int GS = 212;  // doubles of shared scratch per group (208 used; 212*2 = 8 mod 32 banks)
constexpr double TAU = 0.995, THR0 = 0.1;

__device__ __forceinline__ double gsum(unsigned gm, double v) {
  v += __shfl_xor_sync(gm, v, 1);
  v += __shfl_xor_sync(gm, v, 2);
  v += __shfl_xor_sync(gm, v, 4);
  return v;
}
__device__ __forceinline__ double gmaxr(unsigned gm, double v) {
  v = fmax(v, __shfl_xor_sync(gm, v, 1));
  v = fmax(v, __shfl_xor_sync(gm, v, 2));
  v = fmax(v, __shfl_xor_sync(gm, v, 4));
  return v;
}
// NaN-propagating max for residual tracking (fmax drops NaNs)
__device__ __forceinline__ double nmax(double a, double b) { return (b > a || b != b) ? b : a; }

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

__device__ __forceinline__ double row_value(unsigned gm, const Row& R, double xa, double u0, double u1) {
  const double xs = __shfl_sync(gm, xa, R.ix < 0 ? 0 : R.ix, GROUP);
  double v = (R.iu == 0) ? u0 : (R.iu == 1) ? u1 : 0.0;
  if (R.ix >= 0) v = fma(R.sx, xs, v);
  return v;
}

// Scatter of per-row quantities q_r (in shared sq[0..7]) through C': what state lane a and
// the two controls receive from C' q.
//   ty < 2: rows (u_T, u_d, u_T - T, u_d - delta, n, v_x, T, delta)
//   ty = 2: rows (n, v_x, v_y, r)
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

struct Ptrs {
  const double* BA; const double* c;
  double* x; double* u; double* LT; double* K; double* kff; double* Minv; double* kfc; double* D;
  double* xa; double* ua;
};

}  // namespace

// ---------------------------------------------------------------------------------------
// blockDim = 128 (16 problems per CTA), dynamic shared memory = 16 * GS doubles.
// x0 [B][8]; on exit the iterate w.x / w.u holds the QP solution (RTI full step).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) qp_rti_kernel(const KCfg cf, const Work w,
                                                      const double* __restrict__ x0,
                                                      double* __restrict__ u0_out,
                                                      int32_t* __restrict__ status_out) {
  extern __shared__ double smem[];
  const int tid = threadIdx.x;
  const int a = tid & 7;
  const int gl = tid >> 3;
  const int b = blockIdx.x * (blockDim.x >> 3) + gl;
  if (b >= cf.B) return;  // whole group leaves together; groups never sync across each other
  const unsigned gm = 0xFFu << (tid & 24);
  const int N = cf.N;

  double* sBA = smem + gl * GS;  // [8][10]
  double* sW = sBA + 80;         // [8][10]
  double* sv = sBA + 160;        // [8]
  double* sc = sBA + 168;        // [8]
  double* sK = sBA + 176;        // [8][2]
  double* sG = sBA + 192;        // gamma[8], omega[8]

  Ptrs P;
  P.BA = w.BA + (size_t)b * N * 80;
  P.c = w.c + (size_t)b * N * 8;
  P.x = w.x + (size_t)b * (N + 1) * 8;
  P.u = w.u + (size_t)b * N * 2;
  P.LT = w.LT + (size_t)b * (N + 1) * 32;
  P.K = w.K + (size_t)b * N * 16;
  P.kff = w.kff + (size_t)b * N * 2;
  P.Minv = w.Minv + (size_t)b * N * 4;
  P.kfc = w.kfc + (size_t)b * N * 2;
  P.D = w.D + (size_t)b * (N + 1) * 16;
  P.xa = w.xa + (size_t)b * (N + 1) * 8;
  P.ua = w.ua + (size_t)b * N * 2;

  const double xh = x0[(size_t)b * 8 + a];                       // measured state, lane a
  const double s0 = __shfl_sync(gm, xh, 0, GROUP);
  // cost (python/mpc.py:47-69, controller.py:320-385): constant near-diagonal GN Hessian
  const double hx = cf.cs * (cf.W[a] + (a == 6 ? cf.W[10] : a == 7 ? cf.W[11] : 0.0));
  const double hxN = cf.ct * cf.We[a];
  const double hu0 = cf.cs * (cf.W[8] + cf.W[10]), hu1 = cf.cs * (cf.W[9] + cf.W[11]);
  const double hxu0 = (a == 6) ? -cf.cs * cf.W[10] : 0.0;  // H[x_a][u_T]
  const double hxu1 = (a == 7) ? -cf.cs * cf.W[11] : 0.0;  // H[x_a][u_delta]
  const double gs_scale = -cf.cs * cf.W[0];               // g[s]_k = gs_scale * yref_k
  const double gsN = -cf.ct * cf.We[0] * (s0 + cf.s_target);

  // ------------------------------------------------------------------ initialisation
  double res_g0 = 0.0, res_b0 = 0.0, res_d0 = 0.0, mag_g = 0.0, mag_b = 0.0, mag_d = 0.0;
  double S0 = 0.0;
  int nineq = 0;
  {
    double xk = xh;
    if (true) P.x[a] = xh;  // x_0 of the iterate is the measured state
    for (int k = 0; k <= N; k++) {
      const int ty = (k == 0) ? 0 : (k < N ? 1 : 2);
      double u0 = 0.0, u1 = 0.0;
      if (k < N) { const double2 uu = *reinterpret_cast<const double2*>(P.u + 2 * k); u0 = uu.x; u1 = uu.y; }
      const Row R = row_def(cf, ty, a);
      const double v = row_value(gm, R, xk, u0, u1);
      double lamL = 0.0, lamU = 0.0, tL = 1.0, tU = 1.0;
      if (R.act) {
        tL = fmax(v - R.lb, THR0); tU = fmax(R.ub - v, THR0);
        lamL = cf.mu0 / tL; lamU = cf.mu0 / tU;
        res_d0 = nmax(res_d0, nmax(fabs(v - R.lb - tL), fabs(R.ub - v - tU)));
        mag_d = fmax(mag_d, fmax(fabs(R.lb), fabs(R.ub)));
        S0 += lamL * tL + lamU * tU;
        nineq += 2;
      }
      double* lt = P.LT + k * 32;
      lt[a] = lamL; lt[8 + a] = lamU; lt[16 + a] = tL; lt[24 + a] = tU;
      // stationarity residual with pi = 0:  H z + g - C'(lamL - lamU)
      sG[a] = lamL - lamU;
      __syncwarp(gm);
      const double gx = (a == 0) ? ((k < N) ? gs_scale * (s0 + cf.s_target * k / N) : gsN) : 0.0;
      mag_g = fmax(mag_g, fabs(gx));
      if (k > 0) {
        const double rgx = ((k < N) ? hx * xk + hxu0 * u0 + hxu1 * u1 : hxN * xk) + gx - ct_state(sG, ty, a);
        res_g0 = nmax(res_g0, fabs(rgx));
      }
      if (k < N) {
        const double x6 = __shfl_sync(gm, xk, 6, GROUP), x7 = __shfl_sync(gm, xk, 7, GROUP);
        const double rgu0 = hu0 * u0 - cf.cs * cf.W[10] * x6 - (sG[0] + sG[2]);
        const double rgu1 = hu1 * u1 - cf.cs * cf.W[11] * x7 - (sG[1] + sG[3]);
        res_g0 = nmax(res_g0, nmax(fabs(rgu0), fabs(rgu1)));
        // dynamics residual  A x + B u + c - x_next
        const double2* row = reinterpret_cast<const double2*>(P.BA + (size_t)k * 80 + a * 10);
        const double2 bb = row[0];
        double acc = P.c[k * 8 + a];
        mag_b = fmax(mag_b, fabs(acc));
        acc = fma(bb.x, u0, acc); acc = fma(bb.y, u1, acc);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const double2 aa = row[1 + j];
          acc = fma(aa.x, __shfl_sync(gm, xk, 2 * j, GROUP), acc);
          acc = fma(aa.y, __shfl_sync(gm, xk, 2 * j + 1, GROUP), acc);
        }
        const double xn = P.x[(k + 1) * 8 + a];
        res_b0 = nmax(res_b0, fabs(acc - xn));
        xk = xn;
      }
      __syncwarp(gm);
    }
    res_g0 = gmaxr(gm, res_g0); res_b0 = gmaxr(gm, res_b0); res_d0 = gmaxr(gm, res_d0);
    mag_g = gmaxr(gm, mag_g); mag_b = gmaxr(gm, mag_b); mag_d = gmaxr(gm, mag_d);
    S0 = gsum(gm, S0);
    nineq = (int)(gsum(gm, (double)nineq) + 0.5);
  }
  const double inv_n = 1.0 / (double)nineq;
  double mu = S0 * inv_n;
  double res_m = cf.mu0;       // every product starts at mu0
  double rho = 1.0;            // contraction factor of the linear residuals
  int it = 0, qp_status = 1;
  if (!(res_g0 == res_g0) || !(res_b0 == res_b0) || !(mu == mu)) { qp_status = 3; }

  // ------------------------------------------------------------------ IPM iterations
  while (qp_status == 1) {
    const bool conv = rho * res_g0 <= cf.tol * (1.0 + mag_g) && rho * res_b0 <= cf.tol * (1.0 + mag_b) &&
                      rho * res_d0 <= cf.tol * (1.0 + mag_d) && res_m <= cf.tol_comp;
    if (conv) { qp_status = 0; break; }
    if (it >= cf.qp_max_iter) break;

    // ---------------- backward pass 1: Riccati factorisation + predictor gradient
    double Pr[8], pa;
    {
      const Row R = row_def(cf, 2, a);
      const double* lt = P.LT + N * 32;
      const double lamL = lt[a], lamU = lt[8 + a], itL = 1.0 / lt[16 + a], itU = 1.0 / lt[24 + a];
      sG[a] = R.act ? lamL * itL + lamU * itU : 0.0;
      sG[8 + a] = R.act ? lamL * (1.0 + R.lb * itL) + lamU * (R.ub * itU - 1.0) : 0.0;
      __syncwarp(gm);
      const double dg = ctc_diag(sG, 2, a), dom = ct_state(sG + 8, 2, a);
#pragma unroll
      for (int j = 0; j < 8; j++) Pr[j] = (j == a) ? hxN + dg : 0.0;
      pa = ((a == 0) ? gsN : 0.0) - dom;
      __syncwarp(gm);
    }
    for (int k = N - 1; k >= 0; k--) {
      const int ty = (k == 0) ? 0 : 1;
      {  // stage data -> shared
        const double2* row = reinterpret_cast<const double2*>(P.BA + (size_t)k * 80 + a * 10);
        double2* dst = reinterpret_cast<double2*>(sBA + a * 10);
#pragma unroll
        for (int j = 0; j < 5; j++) dst[j] = row[j];
        sc[a] = P.c[k * 8 + a];
        const Row R = row_def(cf, ty, a);
        const double* lt = P.LT + k * 32;
        const double lamL = lt[a], lamU = lt[8 + a], itL = 1.0 / lt[16 + a], itU = 1.0 / lt[24 + a];
        sG[a] = R.act ? lamL * itL + lamU * itU : 0.0;
        sG[8 + a] = R.act ? lamL * (1.0 + R.lb * itL) + lamU * (R.ub * itU - 1.0) : 0.0;
      }
      __syncwarp(gm);
      {  // W = P [B A],  v = P c + p
        double Wr[10];
#pragma unroll
        for (int q = 0; q < 10; q++) Wr[q] = 0.0;
        double va = pa;
#pragma unroll
        for (int j = 0; j < 8; j++) {
          const double2* rj = reinterpret_cast<const double2*>(sBA + j * 10);
#pragma unroll
          for (int q = 0; q < 5; q++) {
            const double2 e = rj[q];
            Wr[2 * q] = fma(Pr[j], e.x, Wr[2 * q]);
            Wr[2 * q + 1] = fma(Pr[j], e.y, Wr[2 * q + 1]);
          }
          va = fma(Pr[j], sc[j], va);
        }
        double2* dst = reinterpret_cast<double2*>(sW + a * 10);
#pragma unroll
        for (int q = 0; q < 5; q++) dst[q] = make_double2(Wr[2 * q], Wr[2 * q + 1]);
        sv[a] = va;
      }
      __syncwarp(gm);
      double Mr[10], mvx, M00 = 0.0, M01 = 0.0, M11 = 0.0, mvu0 = 0.0, mvu1 = 0.0;
      {
#pragma unroll
        for (int q = 0; q < 10; q++) Mr[q] = 0.0;
        mvx = 0.0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
          const double col = sBA[i * 10 + 2 + a];
          const double2 bi = *reinterpret_cast<const double2*>(sBA + i * 10);
          const double2* wi = reinterpret_cast<const double2*>(sW + i * 10);
          const double vi = sv[i];
#pragma unroll
          for (int q = 0; q < 5; q++) {
            const double2 e = wi[q];
            Mr[2 * q] = fma(col, e.x, Mr[2 * q]);
            Mr[2 * q + 1] = fma(col, e.y, Mr[2 * q + 1]);
            if (q == 0) {
              M00 = fma(bi.x, e.x, M00); M01 = fma(bi.x, e.y, M01); M11 = fma(bi.y, e.y, M11);
            }
          }
          mvx = fma(col, vi, mvx);
          mvu0 = fma(bi.x, vi, mvu0); mvu1 = fma(bi.y, vi, mvu1);
        }
      }
      {  // add H + C' Gamma C and the predictor gradient g - C' omega
        const double g0 = sG[0], g1 = sG[1], g2 = sG[2], g3 = sG[3];
        M00 += hu0 + g0 + g2; M11 += hu1 + g1 + g3;
        mvu0 -= sG[8] + sG[10]; mvu1 -= sG[9] + sG[11];
        Mr[0] += hxu0 - ((a == 6) ? g2 : 0.0);
        Mr[1] += hxu1 - ((a == 7) ? g3 : 0.0);
        const double dg = hx + ctc_diag(sG, ty, a);
#pragma unroll
        for (int q = 0; q < 8; q++) Mr[2 + q] += (q == a) ? dg : 0.0;
        const double gx = (a == 0) ? gs_scale * (s0 + cf.s_target * k / N) : 0.0;
        mvx += gx - ct_state(sG + 8, ty, a);
      }
      const double idet = 1.0 / (M00 * M11 - M01 * M01);
      const double i00 = M11 * idet, i01 = -M01 * idet, i11 = M00 * idet;
      const double K0 = -(i00 * Mr[0] + i01 * Mr[1]), K1 = -(i01 * Mr[0] + i11 * Mr[1]);
      const double kf0 = -(i00 * mvu0 + i01 * mvu1), kf1 = -(i01 * mvu0 + i11 * mvu1);
      pa = mvx + K0 * mvu0 + K1 * mvu1;
      *reinterpret_cast<double2*>(sK + 2 * a) = make_double2(K0, K1);
      *reinterpret_cast<double2*>(P.K + k * 16 + 2 * a) = make_double2(K0, K1);
      if (a == 0) {
        *reinterpret_cast<double2*>(P.kff + 2 * k) = make_double2(kf0, kf1);
        *reinterpret_cast<double2*>(P.Minv + 4 * k) = make_double2(i00, i01);
        *reinterpret_cast<double2*>(P.Minv + 4 * k + 2) = make_double2(i11, 0.0);
      }
      __syncwarp(gm);
#pragma unroll
      for (int q = 0; q < 8; q++) {
        const double2 kq = *reinterpret_cast<const double2*>(sK + 2 * q);
        Pr[q] = fma(Mr[0], kq.x, fma(Mr[1], kq.y, Mr[2 + q]));
      }
    }

    // ---------------- forward passes (pred = 1: affine step; pred = 0: corrected step)
    double alpha = 1.0, mu_t = 0.0;
    for (int pred = 1; pred >= 0; pred--) {
      double ratio = 0.0, S1 = 0.0, S2 = 0.0;
      double xk = xh;
      for (int k = 0; k <= N; k++) {
        const int ty = (k == 0) ? 0 : (k < N ? 1 : 2);
        double u0 = 0.0, u1 = 0.0;
        if (k < N) {
          const double2 Kc = *reinterpret_cast<const double2*>(P.K + k * 16 + 2 * a);
          const double2 kf = *reinterpret_cast<const double2*>((pred ? P.kff : P.kfc) + 2 * k);
          u0 = gsum(gm, Kc.x * xk) + kf.x;
          u1 = gsum(gm, Kc.y * xk) + kf.y;
        }
        {  // inequality rows of this stage at the new point
          const Row R = row_def(cf, ty, a);
          const double v = row_value(gm, R, xk, u0, u1);
          if (R.act) {
            const double* lt = P.LT + k * 32;
            const double lamL = lt[a], lamU = lt[8 + a], tL = lt[16 + a], tU = lt[24 + a];
            const double itL = 1.0 / tL, itU = 1.0 / tU;
            const double dtL = v - R.lb - tL, dtU = R.ub - v - tU;
            double dlL, dlU;
            if (pred) {
              dlL = -lamL - lamL * dtL * itL; dlU = -lamU - lamU * dtU * itU;
              S1 += lamL * dtL + tL * dlL + lamU * dtU + tU * dlU;
              S2 += dlL * dtL + dlU * dtU;
              P.D[k * 16 + a] = dlL * dtL; P.D[k * 16 + 8 + a] = dlU * dtU;
            } else {
              const double DL = P.D[k * 16 + a], DU = P.D[k * 16 + 8 + a];
              dlL = -lamL - (DL - mu_t + lamL * dtL) * itL;
              dlU = -lamU - (DU - mu_t + lamU * dtU) * itU;
            }
            ratio = nmax(ratio, nmax(nmax(-dlL / lamL, -dlU / lamU), nmax(-dtL * itL, -dtU * itU)));
          }
        }
        P.xa[k * 8 + a] = xk;
        if (k < N) {
          if (a == 0) *reinterpret_cast<double2*>(P.ua + 2 * k) = make_double2(u0, u1);
          const double2* row = reinterpret_cast<const double2*>(P.BA + (size_t)k * 80 + a * 10);
          const double2 bb = row[0];
          double acc = P.c[k * 8 + a];
          acc = fma(bb.x, u0, acc); acc = fma(bb.y, u1, acc);
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const double2 aa = row[1 + j];
            acc = fma(aa.x, __shfl_sync(gm, xk, 2 * j, GROUP), acc);
            acc = fma(aa.y, __shfl_sync(gm, xk, 2 * j + 1, GROUP), acc);
          }
          xk = acc;
        }
      }
      ratio = gmaxr(gm, ratio);
      if (pred) {
        const double al_aff = (ratio > 1.0) ? 1.0 / ratio : 1.0;
        S1 = gsum(gm, S1); S2 = gsum(gm, S2);
        const double mu_aff = (S0 + al_aff * (S1 + al_aff * S2)) * inv_n;
        const double sr = mu_aff / mu;
        mu_t = fmax(sr * sr * sr * mu, 0.1 * cf.tol_comp);
        if (!(ratio == ratio)) { qp_status = 3; break; }

        // ---------------- backward pass 2: corrector gradient through the same factor
        double dp;
        {
          const Row R = row_def(cf, 2, a);
          const double* lt = P.LT + N * 32;
          const double chi = R.act ? (P.D[N * 16 + a] - mu_t) / lt[16 + a] - (P.D[N * 16 + 8 + a] - mu_t) / lt[24 + a] : 0.0;
          sG[a] = chi;
          __syncwarp(gm);
          dp = ct_state(sG, 2, a);
          __syncwarp(gm);
        }
        for (int k = N - 1; k >= 0; k--) {
          const int ty = (k == 0) ? 0 : 1;
          const Row R = row_def(cf, ty, a);
          const double* lt = P.LT + k * 32;
          const double chi = R.act ? (P.D[k * 16 + a] - mu_t) / lt[16 + a] - (P.D[k * 16 + 8 + a] - mu_t) / lt[24 + a] : 0.0;
          sG[a] = chi;
          __syncwarp(gm);
          double dmx = ct_state(sG, ty, a);
          double dmu0 = sG[0] + sG[2], dmu1 = sG[1] + sG[3];
          const double2 bb = *reinterpret_cast<const double2*>(P.BA + (size_t)k * 80 + a * 10);
          dmu0 += gsum(gm, bb.x * dp);
          dmu1 += gsum(gm, bb.y * dp);
#pragma unroll
          for (int i = 0; i < 8; i++) dmx = fma(P.BA[(size_t)k * 80 + i * 10 + 2 + a], __shfl_sync(gm, dp, i, GROUP), dmx);
          const double2 Kc = *reinterpret_cast<const double2*>(P.K + k * 16 + 2 * a);
          const double2 m0 = *reinterpret_cast<const double2*>(P.Minv + 4 * k);
          const double i11 = P.Minv[4 * k + 2];
          const double2 kf = *reinterpret_cast<const double2*>(P.kff + 2 * k);
          if (a == 0)
            *reinterpret_cast<double2*>(P.kfc + 2 * k) =
                make_double2(kf.x - (m0.x * dmu0 + m0.y * dmu1), kf.y - (m0.y * dmu0 + i11 * dmu1));
          dp = dmx + Kc.x * dmu0 + Kc.y * dmu1;
          __syncwarp(gm);
        }
      } else {
        const double amax = (ratio > 0.0) ? 1.0 / ratio : 2.0;
        alpha = fmin(1.0, TAU * amax);
        if (!(ratio == ratio)) qp_status = 3;
      }
    }
    if (qp_status == 3) break;
    if (alpha < 1e-12) { qp_status = 2; break; }

    // ---------------- update sweep
    {
      double Sn = 0.0, rm = 0.0;
      for (int k = 0; k <= N; k++) {
        const int ty = (k == 0) ? 0 : (k < N ? 1 : 2);
        const double xc = P.xa[k * 8 + a];
        double u0 = 0.0, u1 = 0.0;
        if (k < N) { const double2 uu = *reinterpret_cast<const double2*>(P.ua + 2 * k); u0 = uu.x; u1 = uu.y; }
        const Row R = row_def(cf, ty, a);
        const double v = row_value(gm, R, xc, u0, u1);
        if (R.act) {
          double* lt = P.LT + k * 32;
          double lamL = lt[a], lamU = lt[8 + a], tL = lt[16 + a], tU = lt[24 + a];
          const double dtL = v - R.lb - tL, dtU = R.ub - v - tU;
          const double dlL = -lamL - (P.D[k * 16 + a] - mu_t + lamL * dtL) / tL;
          const double dlU = -lamU - (P.D[k * 16 + 8 + a] - mu_t + lamU * dtU) / tU;
          lamL = fma(alpha, dlL, lamL); lamU = fma(alpha, dlU, lamU);
          tL = fma(alpha, dtL, tL); tU = fma(alpha, dtU, tU);
          lt[a] = lamL; lt[8 + a] = lamU; lt[16 + a] = tL; lt[24 + a] = tU;
          Sn += lamL * tL + lamU * tU;
          rm = nmax(rm, nmax(lamL * tL, lamU * tU));
        }
        const double xo = P.x[k * 8 + a];
        P.x[k * 8 + a] = fma(alpha, xc - xo, xo);
        if (k < N && a < 2) {
          const double uo = P.u[2 * k + a];
          P.u[2 * k + a] = fma(alpha, ((a == 0) ? u0 : u1) - uo, uo);
        }
      }
      S0 = gsum(gm, Sn);
      res_m = gmaxr(gm, rm);
      mu = S0 * inv_n;
      rho *= (1.0 - alpha);
      it++;
      if (!(mu == mu)) { qp_status = 3; break; }
    }
  }

  // ------------------------------------------------------------------ outputs
  if (a == 0) {
    int st;
    if (qp_status == 3) st = IHM2B_FAILURE;
    else if (qp_status == 0) st = IHM2B_SUCCESS;
    else {
      const bool loose = rho * res_g0 <= 1e-6 * (1.0 + mag_g) && rho * res_b0 <= 1e-6 * (1.0 + mag_b) &&
                         rho * res_d0 <= 1e-6 * (1.0 + mag_d) && res_m <= 1e-6;
      st = loose ? IHM2B_SUCCESS : IHM2B_QP_FAILURE;
    }
    w.status[b] = st;
    if (status_out) status_out[b] = st;
    double* sp = w.stats + (size_t)b * 8;
    sp[0] = it; sp[1] = 1.0; sp[2] = mu; sp[3] = rho * res_g0; sp[4] = rho * res_b0; sp[5] = rho * res_d0;
    sp[6] = res_m; sp[7] = 0.0;
  }
  if (u0_out && a < 2) u0_out[(size_t)b * 2 + a] = P.u[a];
}

void launch_qp_rti(const KCfg& cf, const Work& w, const double* x0, double* u0, int32_t* status,
                   cudaStream_t stream) {
  const int per_cta = 128 / GROUP;
  const int grid = (cf.B + per_cta - 1) / per_cta;
  qp_rti_kernel<<<grid, 128, per_cta * GS * sizeof(double), stream>>>(cf, w, x0, u0, status);
}

}  // namespace ihm2b
