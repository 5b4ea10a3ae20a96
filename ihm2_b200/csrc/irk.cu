// irk.cu -- 4-stage Gauss-Legendre collocation step of the controller model with forward
// sensitivities (acados IRK, sim_method_num_stages = 4, num_steps = 1, configured at
// new_python/controller.py:305-307 of the reference; acados' sim_irk is not in its tree).
//
// The stage equations K_i = f(x + h sum_j a_ij K_j, u) are solved to round-off, and the
// sensitivities follow from the implicit-function theorem on the same system.  fkin6 is a
// cascade, which the solver exploits instead of factorising the dense 32 x 32 Newton matrix:
//   (T, delta)  linear first-order lags      -> closed form (4 x 4 solve done once per
//                                               instance on the host, api.cu)
//   (v_x, v_y)  depend on themselves, T, delta -> Newton on 8 unknowns
//   r           explicit in the others
//   (s, n, psi) depend on all of the above     -> Newton on 12 unknowns
// The block-triangular Newton matrix gives the same converged point as the dense iteration.
//
// Mapping: a block owns NI = 32 shooting intervals and keeps their LU factors, stage slopes and
// Jacobians in SHARED memory (372 doubles per interval, interval-interleaved so every access is
// conflict-free).  Phase 1: four lanes per interval (one collocation stage each) run the two Newton solves in
// registers, exchanging slopes and pivot rows by shuffles.  Phase 2: one
// thread per (interval, 2-3 columns of d/d(x,u)) -- 128 threads, the column is warp-uniform -- runs
// the block forward substitution of its column through the shared factors and writes its
// column of [B A].  Phase 3: four threads per interval form c and phi (two rows each).  (A thread-per-interval
// version with the factors in a 5.4 KB local-memory frame was bound by local-memory misses:
// 5.3 GB of DRAM traffic per launch against 126 MB algorithmic.)
#include "model.cuh"

namespace ihm2b {

namespace {

constexpr int NS = 4;
constexpr int MAX_NEWTON = 12;

constexpr int NI = 32;     // intervals per block; shared arrays are [entry][NI]

// in-place LU without pivoting (the matrices are I - h A (x) J, close to the identity);
// element (r, c) at M[(r * n + c) * NI]; L below the diagonal (unit), U above, reciprocal pivots on it
template <int n, typename Real>
__device__ __forceinline__ void lu_factor(Real* M) {
  // left-looking, one row at a time in registers: row r is reduced against the finished rows c < r, so the
  // shared-memory loads never wait on a store of the same sweep (the right-looking form is a chain of
  // load - fma - store round trips through shared memory, ~10x slower here)
#pragma unroll
  for (int r = 0; r < n; r++) {
    Real row[n];
#pragma unroll
    for (int j = 0; j < n; j++) row[j] = M[(r * n + j) * NI];
#pragma unroll
    for (int c = 0; c < r; c++) {
      const Real l = row[c] * M[(c * n + c) * NI];   // reciprocal pivot of row c
      row[c] = l;
#pragma unroll
      for (int j = c + 1; j < n; j++) row[j] = fma(-l, M[(c * n + j) * NI], row[j]);
    }
    row[r] = Real(1.0) / row[r];  // store the reciprocal pivot
#pragma unroll
    for (int j = 0; j < n; j++) M[(r * n + j) * NI] = row[j];
  }
}
template <int n, typename Real>
__device__ __forceinline__ void lu_solve(const Real* M, Real* b) {
#pragma unroll
  for (int r = 1; r < n; r++) {
    Real v = b[r];
#pragma unroll
    for (int c = 0; c < r; c++) v = fma(-M[(r * n + c) * NI], b[c], v);
    b[r] = v;
  }
#pragma unroll
  for (int r = n - 1; r >= 0; r--) {
    Real v = b[r];
#pragma unroll
    for (int c = r + 1; c < n; c++) v = fma(-M[(r * n + c) * NI], b[c], v);
    b[r] = v * M[(r * n + r) * NI];
  }
}

// shared-memory record of one interval (offsets in entries)
constexpr int OFF_MV = 0;            // LU of the (v_x, v_y) Newton matrix [8][8]
constexpr int OFF_MP = 64;           // LU of the (s, n, psi) Newton matrix [12][12]
constexpr int OFF_K = 208;           // converged stage slopes K[4][8]
constexpr int OFF_J = 240;           // Jacobians at the 4 collocation points, 33 kept entries each
constexpr int OFF_XN = 372;          // x_next [8]
constexpr int PER_INT = 380;
// kept entries of J (JMASK): rows 0..2 x columns 0..5, rows 3..5 x columns {3,4,6,7,9}
__device__ __forceinline__ constexpr int jslot(int row, int col) {
  return row < 3 ? (col < 6 ? row * 6 + col : -1)
                 : (col == 3 ? 18 + (row - 3) * 5 : col == 4 ? 19 + (row - 3) * 5 : col == 6 ? 20 + (row - 3) * 5
                    : col == 7 ? 21 + (row - 3) * 5 : col == 9 ? 22 + (row - 3) * 5 : -1);
}

// quantities of a collocation point that depend only on (T_i, delta_i, d delta_i / dt)
template <typename Real>
struct VPoint {
  Real cb, sb, bdot, cdb, Fm;
};

}  // namespace


// ---- phase 1, four lanes per interval (lane t of the group owns collocation stage t): the stage evaluations
// (tanh / sincos / curvature lookup), the rows of the Newton matrices and the LU factorisations are spread over the
// four lanes and exchanged by 4-lane shuffles; nothing touches shared memory until the converged factors, slopes and
// Jacobians are written for phase 2.  All 32 lanes of the warp (8 intervals) run the same number of Newton iterations;
// an interval that has converged stops updating and re-evaluates at the point its factors belong to, so its result
// does not depend on its neighbours in the warp.
template <typename Real> __device__ __forceinline__ constexpr Real newton_tol() { return sizeof(Real) == 8 ? Real(1e-14) : Real(2e-6); }
constexpr unsigned FULLW = 0xffffffffu;
template <typename Real>
__device__ __forceinline__ Real g4(Real v, int src) { return __shfl_sync(FULLW, v, src, 4); }
template <typename Real>
__device__ __forceinline__ Real g4max(Real v) {
  v = fmax(v, __shfl_xor_sync(FULLW, v, 1)); v = fmax(v, __shfl_xor_sync(FULLW, v, 2));
  return v;
}
template <typename Real>
__device__ __forceinline__ Real g4sum(Real v) {
  v += __shfl_xor_sync(FULLW, v, 1); v += __shfl_xor_sync(FULLW, v, 2);
  return v;
}

// LU (no pivoting, unit L, reciprocal pivots on the diagonal: the layout lu_solve reads) of the n x n matrix whose rows
// t * RPL .. t * RPL + RPL - 1 lane t holds, and the solution of M x = b; x returned in b (own rows).
template <int RPL, typename Real>
__device__ __forceinline__ void lu4_factor_solve(Real (&M)[RPL][4 * RPL], Real (&b)[RPL], const int t) {
  constexpr int n = 4 * RPL;
#pragma unroll
  for (int p = 0; p < n; p++) {
    const int owner = p / RPL, pr = p % RPL;
    const Real rp = g4(Real(1.0) / M[pr][p], owner);
    Real prow[n];
#pragma unroll
    for (int c = p + 1; c < n; c++) prow[c] = g4(M[pr][c], owner);
    const Real pb = g4(b[pr], owner);
#pragma unroll
    for (int r = 0; r < RPL; r++) {
      const bool below = (t * RPL + r > p);
      const Real l = M[r][p] * rp;
      if (below) {
        M[r][p] = l;
#pragma unroll
        for (int c = p + 1; c < n; c++) M[r][c] = fma(-l, prow[c], M[r][c]);
        b[r] = fma(-l, pb, b[r]);
      }
    }
    if (t == owner) M[pr][p] = rp;
  }
  Real xs[n];
#pragma unroll
  for (int p = n - 1; p >= 0; p--) {
    const int owner = p / RPL, pr = p % RPL;
    Real v = b[pr];
#pragma unroll
    for (int c = p + 1; c < n; c++) v = fma(-M[pr][c], xs[c], v);
    xs[p] = g4(v * M[pr][p], owner);
  }
#pragma unroll
  for (int r = 0; r < RPL; r++) {
    Real v = Real(0.0);
#pragma unroll
    for (int o = 0; o < 4; o++) v = (t == o) ? xs[o * RPL + r] : v;
    b[r] = v;
  }
}

template <typename Real>
__device__ bool irk_newton4(const int t, const Real* __restrict__ x, Real u0, Real u1, const Real* __restrict__ p,
                            const TrackView& tr, Real h, const Real* __restrict__ A4, const Real* __restrict__ b4,
                            const Real* __restrict__ wT, const Real* __restrict__ wD, const double* __restrict__ lag_d,
                            Real* __restrict__ sm) {
  const Real m = p[IHM2B_P_M], lR = p[IHM2B_P_LR], lF = p[IHM2B_P_LF];
  const Real C = lR / (lR + lF), im = Real(1.0) / m;
  const Real itT = Real(1.0) / p[IHM2B_P_TT], itD = Real(1.0) / p[IHM2B_P_TD];
  Real ha[NS];
#pragma unroll
  for (int j = 0; j < NS; j++) ha[j] = h * A4[t * NS + j];
  // stage value of a component from the slopes of the four lanes
  auto stage_val = [&](Real x0, Real Kown) {
    Real v = x0;
#pragma unroll
    for (int j = 0; j < NS; j++) v = fma(ha[j], g4(Kown, j), v);
    return v;
  };
  // ---- actuators: K = w (u - x) / t  (closed form, every lane knows all four)
  // (u - x) / t of the two lags arrives in fp64 (lag_d): in fp32 the difference of two numbers of size 400 times 1 / t_T = 1000
  // would cost three digits of the torque rows
  const Real c6 = (Real)lag_d[0], c7 = (Real)lag_d[1];
  Real K[NX];
  K[6] = wT[t] * c6; K[7] = wD[t] * c7;
  Real Ti = x[6], Di = x[7];
#pragma unroll
  for (int j = 0; j < NS; j++) { Ti = fma(ha[j], wT[j] * c6, Ti); Di = fma(ha[j], wD[j] * c7, Di); }
  // ---- (v_x, v_y): Newton on 8 unknowns, rows 2t, 2t + 1 on this lane
  VPoint<Real> vp;
  {
    Real sd, cd;
    sincos(Di, &sd, &cd);
    const Real td = sd / cd, Ctd = C * td, den = Real(1.0) + Ctd * Ctd;
    vp.cb = rsqrt(den); vp.sb = Ctd * vp.cb;
    vp.bdot = C * (Real(1.0) + td * td) / den * K[7];
    vp.cdb = cd * vp.cb + sd * vp.sb;
    vp.Fm = p[IHM2B_P_CM0] * Ti;
  }
  auto veval = [&](Real vx, Real vy, Real& f3, Real& f4, Real& dvd) {
    const Real th = tanh(Real(10.0) * vx);
    const Real poly = p[IHM2B_P_CR0] + p[IHM2B_P_CR1] * vx + p[IHM2B_P_CR2] * vx * vx;
    const Real FRx = Real(0.5) * vp.Fm - poly * th;
    const Real vdot = (FRx * vp.cb + Real(0.5) * vp.Fm * vp.cdb) * im;
    f3 = vdot * vp.cb - vp.bdot * vy;
    f4 = vdot * vp.sb + vp.bdot * vx;
    dvd = (-(p[IHM2B_P_CR1] + Real(2.0) * p[IHM2B_P_CR2] * vx) * th - poly * Real(10.0) * (Real(1.0) - th * th)) * vp.cb * im;
  };
  { Real d; veval(x[3], x[4], K[3], K[4], d); }
  bool ok = true, conv = false;
  Real Mv[2][8];
  Real P3 = K[3], P4 = K[4];      // the point the factors belong to (frozen at convergence)
#pragma unroll 1
  for (int it = 0;; it++) {
    if (!conv) { P3 = K[3]; P4 = K[4]; }
    const Real vx = stage_val(x[3], P3), vy = stage_val(x[4], P4);
    Real f3, f4, dvd;
    veval(vx, vy, f3, f4, dvd);
    Real R[2] = {P3 - f3, P4 - f4};
    const Real J33 = dvd * vp.cb, J34 = -vp.bdot, J43 = dvd * vp.sb + vp.bdot;
#pragma unroll
    for (int j = 0; j < NS; j++) {
      const Real e = (t == j) ? Real(1.0) : Real(0.0);
      Mv[0][2 * j] = e - ha[j] * J33; Mv[0][2 * j + 1] = -ha[j] * J34;
      Mv[1][2 * j] = -ha[j] * J43;    Mv[1][2 * j + 1] = e;
    }
    lu4_factor_solve<2, Real>(Mv, R, t);
    const Real k3 = P3 - R[0], k4 = P4 - R[1];
    const Real dmax = g4max(fmax(fabs(R[0]), fabs(R[1]))), kmax = g4max(fmax(fabs(k3), fabs(k4)));
    if (!conv) {
      K[3] = k3; K[4] = k4;
      if (!(dmax == dmax)) { ok = false; conv = true; }
      else if (dmax <= newton_tol<Real>() * (Real(1.0) + kmax) || it >= MAX_NEWTON) conv = true;
    }
    if (__all_sync(FULLW, conv)) break;
  }
  // ---- r: explicit, r' = l_R v_y' - beta'   (python/models.py:304)
  K[5] = lR * K[4] - vp.bdot;
  const Real vxi = stage_val(x[3], K[3]), vyi = stage_val(x[4], K[4]), ri = stage_val(x[5], K[5]);
  // ---- (s, n, psi): Newton on 12 unknowns, rows 3t .. 3t + 2 on this lane
  {
    Real kap, dkap, sp, cp;
    kappa_lookup(tr, x[0], kap, dkap);
    sincos(x[2], &sp, &cp);
    const Real sdot = (x[3] * cp - x[4] * sp) / (Real(1.0) + kap * x[1]);
    K[0] = sdot; K[1] = x[3] * sp + x[4] * cp; K[2] = x[5] - kap * sdot;
  }
  Real Mp[3][12];
  Real Q0 = K[0], Q1 = K[1], Q2 = K[2];
  conv = !ok;
#pragma unroll 1
  for (int it = 0;; it++) {
    if (!conv) { Q0 = K[0]; Q1 = K[1]; Q2 = K[2]; }
    const Real s_ = stage_val(x[0], Q0), n_ = stage_val(x[1], Q1), psi = stage_val(x[2], Q2);
    Real kap, dkap, sp, cp;
    kappa_lookup(tr, s_, kap, dkap);
    sincos(psi, &sp, &cp);
    const Real iD = Real(1.0) / (Real(1.0) + kap * n_);
    const Real num = vxi * cp - vyi * sp;
    const Real sdot = num * iD;
    Real R[3] = {Q0 - sdot, Q1 - (vxi * sp + vyi * cp), Q2 - (ri - kap * sdot)};
    Real J[3][3];
    J[0][0] = -sdot * dkap * n_ * iD; J[0][1] = -sdot * kap * iD; J[0][2] = -(vxi * sp + vyi * cp) * iD;
    J[1][0] = Real(0.0); J[1][1] = Real(0.0); J[1][2] = num;
    J[2][0] = -dkap * sdot - kap * J[0][0]; J[2][1] = -kap * J[0][1]; J[2][2] = -kap * J[0][2];
#pragma unroll
    for (int j = 0; j < NS; j++)
#pragma unroll
      for (int a = 0; a < 3; a++)
#pragma unroll
        for (int c = 0; c < 3; c++) Mp[a][3 * j + c] = ((t == j && a == c) ? Real(1.0) : Real(0.0)) - ha[j] * J[a][c];
    lu4_factor_solve<3, Real>(Mp, R, t);
    const Real k0 = Q0 - R[0], k1 = Q1 - R[1], k2 = Q2 - R[2];
    const Real dmax = g4max(fmax(fabs(R[0]), fmax(fabs(R[1]), fabs(R[2]))));
    const Real kmax = g4max(fmax(fabs(k0), fmax(fabs(k1), fabs(k2))));
    if (!conv) {
      K[0] = k0; K[1] = k1; K[2] = k2;
      if (!(dmax == dmax)) { ok = false; conv = true; }
      else if (dmax <= newton_tol<Real>() * (Real(1.0) + kmax) || it >= MAX_NEWTON) conv = true;
    }
    if (__all_sync(FULLW, conv)) break;
  }
  // ---- hand over to phase 2 / 3: x_next, slopes, factors, Jacobian at this lane's collocation point
#pragma unroll
  for (int a = 0; a < NX; a++) {
    const Real v = g4sum(h * b4[t] * K[a]);           // increment x_next - x: added to the fp64 state in phase 3
    if (t == 0) sm[(OFF_XN + a) * NI] = v;
    sm[(OFF_K + t * 8 + a) * NI] = K[a];
  }
#pragma unroll
  for (int r = 0; r < 2; r++)
#pragma unroll
    for (int c = 0; c < 8; c++) sm[(OFF_MV + (2 * t + r) * 8 + c) * NI] = Mv[r][c];
#pragma unroll
  for (int r = 0; r < 3; r++)
#pragma unroll
    for (int c = 0; c < 12; c++) sm[(OFF_MP + (3 * t + r) * 12 + c) * NI] = Mp[r][c];
  {
    Real xs[NX], f[NX], J[NX][NZ];
#pragma unroll
    for (int a = 0; a < NX; a++) xs[a] = stage_val(x[a], K[a]);
    fkin6<true, Real>(xs, u0, u1, p, tr, f, J);
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
      for (int c = 0; c < NZ; c++)
        if (jslot(r, c) >= 0) sm[(OFF_J + t * 33 + jslot(r, c)) * NI] = ((JMASK[r] >> c) & 1u) ? J[r][c] : Real(0.0);
  }
  return g4max(ok ? Real(0.0) : Real(1.0)) == Real(0.0);
}

// ---- phase 2: NC columns of d x_next / d (x, u) at once (columns col0, col0 + 4, ...):  (I - h A (x) J) dK = J_x E + J_u
// by block forward substitution through the shared factors.  Every factor / Jacobian entry read from shared memory serves
// all NC right-hand sides.  `col0` is uniform over the warp.
template <int n, int NC, typename Real>
__device__ __forceinline__ void lu_solve_multi(const Real* M, Real (&b)[NC][n]) {
#pragma unroll
  for (int r = 1; r < n; r++)
#pragma unroll
    for (int c = 0; c < r; c++) {
      const Real l = M[(r * n + c) * NI];
#pragma unroll
      for (int q = 0; q < NC; q++) b[q][r] = fma(-l, b[q][c], b[q][r]);
    }
#pragma unroll
  for (int r = n - 1; r >= 0; r--) {
#pragma unroll
    for (int c = r + 1; c < n; c++) {
      const Real m = M[(r * n + c) * NI];
#pragma unroll
      for (int q = 0; q < NC; q++) b[q][r] = fma(-m, b[q][c], b[q][r]);
    }
    const Real ip = M[(r * n + r) * NI];
#pragma unroll
    for (int q = 0; q < NC; q++) b[q][r] *= ip;
  }
}

template <int NC, typename Real>
__device__ void irk_sens_columns(const int col0, const Real* __restrict__ sm, Real h, const Real* __restrict__ A4,
                                 const Real* __restrict__ b4, const Real* __restrict__ wT,
                                 const Real* __restrict__ wD, Real itT, Real itD, Real (&Scol)[NC][NX]) {
  const Real* const Mv = sm + OFF_MV * NI;
  const Real* const Mp = sm + OFF_MP * NI;
#define JC(i, r, c) sm[(OFF_J + (i) * 33 + jslot(r, c)) * NI]
  Real hA[NS][NS], hb[NS];
#pragma unroll
  for (int i = 0; i < NS; i++) {
    hb[i] = h * b4[i];
#pragma unroll
    for (int j = 0; j < NS; j++) hA[i][j] = h * A4[i * NS + j];
  }
  // stage sums h A dK and output rows e_a + h b' dK of one quantity, for all columns
  auto stage_sum = [&](const Real (&dk)[NC][NS], Real (&dx)[NC][NS]) {
#pragma unroll
    for (int q = 0; q < NC; q++)
#pragma unroll
      for (int i = 0; i < NS; i++) {
        Real v = Real(0.0);
#pragma unroll
        for (int j = 0; j < NS; j++) v = fma(hA[i][j], dk[q][j], v);
        dx[q][i] = v;
      }
  };
  auto out_row = [&](int a, const Real (&dk)[NC][NS]) {
#pragma unroll
    for (int q = 0; q < NC; q++) {
      Real v = (a == col0 + 4 * q) ? Real(1.0) : Real(0.0);
#pragma unroll
      for (int i = 0; i < NS; i++) v = fma(hb[i], dk[q][i], v);
      Scol[q][a] = v;
    }
  };
  // direct term of row `row` of the Jacobian at stage i for column col (structural zeros are not stored)
  auto direct = [&](int i, int row, int col) {
    Real v = Real(0.0);
    if (row < 3) {
      if (col == 0) v = JC(i, row, 0);
      else if (col == 1) v = JC(i, row, 1);
      else if (col == 2) v = JC(i, row, 2);
      else if (col == 3) v = JC(i, row, 3);
      else if (col == 4) v = JC(i, row, 4);
      else if (col == 5) v = JC(i, row, 5);
    } else {
      if (col == 3) v = JC(i, row, 3);
      else if (col == 4) v = JC(i, row, 4);
      else if (col == 6) v = JC(i, row, 6);
      else if (col == 7) v = JC(i, row, 7);
      else if (col == 9) v = JC(i, row, 9);
    }
    return v;
  };
  Real dK[NC][NS], dX6[NC][NS], dX7[NC][NS], dX3[NC][NS], dX4[NC][NS], dX5[NC][NS];
  // actuators (closed form)
#pragma unroll
  for (int q = 0; q < NC; q++)
#pragma unroll
    for (int i = 0; i < NS; i++) { const int col = col0 + 4 * q; dK[q][i] = (col == 6) ? -wT[i] * itT : (col == 8) ? wT[i] * itT : Real(0.0); }
  stage_sum(dK, dX6); out_row(6, dK);
#pragma unroll
  for (int q = 0; q < NC; q++)
#pragma unroll
    for (int i = 0; i < NS; i++) { const int col = col0 + 4 * q; dK[q][i] = (col == 7) ? -wD[i] * itD : (col == 9) ? wD[i] * itD : Real(0.0); }
  stage_sum(dK, dX7); out_row(7, dK);
  // (v_x, v_y): columns that do not reach these rows have a zero right-hand side and come out zero
  Real dK3[NC][NS], dK4[NC][NS];
  {
    Real Rv[NC][8];
#pragma unroll
    for (int i = 0; i < NS; i++)
#pragma unroll
      for (int a = 0; a < 2; a++) {
        const Real j6 = JC(i, 3 + a, 6), j7 = JC(i, 3 + a, 7);
#pragma unroll
        for (int q = 0; q < NC; q++) Rv[q][2 * i + a] = fma(j6, dX6[q][i], fma(j7, dX7[q][i], direct(i, 3 + a, col0 + 4 * q)));
      }
    lu_solve_multi<8, NC, Real>(Mv, Rv);
#pragma unroll
    for (int q = 0; q < NC; q++)
#pragma unroll
      for (int i = 0; i < NS; i++) { dK3[q][i] = Rv[q][2 * i]; dK4[q][i] = Rv[q][2 * i + 1]; }
  }
  stage_sum(dK3, dX3); stage_sum(dK4, dX4);
  out_row(3, dK3); out_row(4, dK4);
  // r
#pragma unroll
  for (int i = 0; i < NS; i++) {
    const Real j3 = JC(i, 5, 3), j6 = JC(i, 5, 6), j7 = JC(i, 5, 7);
#pragma unroll
    for (int q = 0; q < NC; q++) dK[q][i] = fma(j3, dX3[q][i], fma(j6, dX6[q][i], fma(j7, dX7[q][i], direct(i, 5, col0 + 4 * q))));
  }
  stage_sum(dK, dX5); out_row(5, dK);
  // (s, n, psi)
  Real Rp[NC][12];
#pragma unroll
  for (int i = 0; i < NS; i++)
#pragma unroll
    for (int a = 0; a < 3; a++) {
      const Real j3 = JC(i, a, 3), j4 = JC(i, a, 4), j5 = JC(i, a, 5);
#pragma unroll
      for (int q = 0; q < NC; q++) Rp[q][3 * i + a] = fma(j3, dX3[q][i], fma(j4, dX4[q][i], fma(j5, dX5[q][i], direct(i, a, col0 + 4 * q))));
    }
  lu_solve_multi<12, NC, Real>(Mp, Rp);
#pragma unroll
  for (int a = 0; a < 3; a++) {
    Real dk[NC][NS];
#pragma unroll
    for (int q = 0; q < NC; q++)
#pragma unroll
      for (int i = 0; i < NS; i++) dk[q][i] = Rp[q][3 * i + a];
    out_row(a, dk);
  }
#undef JC
}

#ifndef LIN_MIN_BLOCKS
#define LIN_MIN_BLOCKS 2
#endif
constexpr int NCG = 4;     // column groups: a thread of phase 2 handles the columns cg, cg + 4, (cg + 8)
// grid: ceil(B*N / NI) blocks of NI * NCG threads; dynamic shared memory PER_INT * NI scalars.
// Real = double: 2 blocks per SM by shared memory; 8 warps per SM leave phase 1 its registers.
// Real = float (ihm2b_config::linearize_fp32, the mixed-precision path): the collocation system, its Newton solves and
// the sensitivities in fp32; the iterate is read and [B A], c, phi are written in fp64, the state increment of the step
// is added to the fp64 state and c = phi - A x - B u is formed in fp64, so the QP sees dynamics that are consistent to
// round-off with an integrator that is accurate to ~1e-6 relative.
template <typename Real>
__global__ void __launch_bounds__(NI * NCG, LIN_MIN_BLOCKS) linearize_irk_kernel(const KCfg cf, const Work w) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Real* const smem = reinterpret_cast<Real*>(smem_raw);
  __shared__ int s_ok[NI];
  const int tid = threadIdx.x;
  const int j = tid & (NI - 1), cg = tid / NI;           // interval of the block, column group (warp-uniform)
  const int gid = blockIdx.x * NI + j;
  const bool in_range = gid < cf.B * cf.N;
  const int b = in_range ? gid / cf.N : 0, k = in_range ? gid - b * cf.N : 0;
  const bool live = in_range && !(w.active && !w.active[b]);   // SQP mode: instance already converged or failed
  Real* const sm = smem + j;
  Real wa[8], tabA[16], tabb[4];
  for (int i = 0; i < 8; i++) wa[i] = (Real)__ldg(w.actw + (size_t)b * 8 + i);
  for (int i = 0; i < 16; i++) tabA[i] = (Real)cf.irk_A[i];
  for (int i = 0; i < 4; i++) tabb[i] = (Real)cf.irk_b[i];
  const Real itT = (Real)(1.0 / __ldg(w.params + (size_t)b * NP + IHM2B_P_TT)), itD = (Real)(1.0 / __ldg(w.params + (size_t)b * NP + IHM2B_P_TD));
  const Real hh = (Real)cf.dt;
  const size_t sk = (size_t)b * cf.N + k;
  {
    // phase 1: lanes 4q .. 4q + 3 of warp wp work on interval 8 wp + q of the block
    const int t = tid & 3, j1 = (tid >> 5) * 8 + ((tid & 31) >> 2);
    const int gid1 = blockIdx.x * NI + j1;
    const bool in1 = gid1 < cf.B * cf.N;
    const int b1 = in1 ? gid1 / cf.N : 0, k1 = in1 ? gid1 - b1 * cf.N : 0;   // out of range: a valid dummy, results unused
    const bool live1 = in1 && !(w.active && !w.active[b1]);
    if (__any_sync(0xffffffffu, live1)) {     // a warp whose 8 intervals are all masked out (SQP / line search / closed loop) skips
      const int tk = w.track[b1];
      TrackView tr{cf.n_grid, w.s_ref + (size_t)tk * cf.n_grid, w.kappa_ref + (size_t)tk * cf.n_grid};
      Real p[NP], x[NX], wa1[8];
      for (int i = 0; i < NP; i++) p[i] = (Real)__ldg(w.params + (size_t)b1 * NP + i);
      for (int i = 0; i < 8; i++) wa1[i] = (Real)__ldg(w.actw + (size_t)b1 * 8 + i);
      for (int i = 0; i < NX; i++) x[i] = (Real)w.x[((size_t)b1 * (cf.N + 1) + k1) * NX + i];
      const Real u0 = (Real)w.u[((size_t)b1 * cf.N + k1) * NU], u1 = (Real)w.u[((size_t)b1 * cf.N + k1) * NU + 1];
      const double lag_d[2] = {(w.u[((size_t)b1 * cf.N + k1) * NU] - w.x[((size_t)b1 * (cf.N + 1) + k1) * NX + 6]) / __ldg(w.params + (size_t)b1 * NP + IHM2B_P_TT),
                               (w.u[((size_t)b1 * cf.N + k1) * NU + 1] - w.x[((size_t)b1 * (cf.N + 1) + k1) * NX + 7]) / __ldg(w.params + (size_t)b1 * NP + IHM2B_P_TD)};
      const bool ok = irk_newton4<Real>(t, x, u0, u1, p, tr, hh, tabA, tabb, wa1, wa1 + 4, lag_d, smem + j1);
      if (t == 0) s_ok[j1] = ok;
    }
  }
  __syncthreads();
  double* ba = w.BA + sk * NX * NZ;
  if (live) {
    // warps 0, 1 hold three columns (cg, cg + 4, cg + 8), warps 2, 3 two
    if (cg < 2) {
      Real Scol[3][NX];
      irk_sens_columns<3, Real>(cg, sm, hh, tabA, tabb, wa, wa + 4, itT, itD, Scol);
#pragma unroll
      for (int q = 0; q < 3; q++) {
        const int col = cg + 4 * q, oc = (col < NX) ? 2 + col : col - NX;     // [B A] column of d/d(x,u) column `col`
#pragma unroll
        for (int a = 0; a < NX; a++) ba[a * NZ + oc] = (double)Scol[q][a];
      }
    } else {
      Real Scol[2][NX];
      irk_sens_columns<2, Real>(cg, sm, hh, tabA, tabb, wa, wa + 4, itT, itD, Scol);
#pragma unroll
      for (int q = 0; q < 2; q++) {
        const int col = cg + 4 * q, oc = 2 + col;
#pragma unroll
        for (int a = 0; a < NX; a++) ba[a * NZ + oc] = (double)Scol[q][a];
      }
    }
  }
  __syncthreads();
  if (live) {   // phase 3: the thread of column group cg forms rows 2 cg, 2 cg + 1 of c and phi (all four warps at work)
    const bool ok = s_ok[j] != 0;
    double x[NX];
    for (int i = 0; i < NX; i++) x[i] = w.x[((size_t)b * (cf.N + 1) + k) * NX + i];
    const double u0 = w.u[sk * NU], u1 = w.u[sk * NU + 1];
#pragma unroll
    for (int ii = 0; ii < 2; ii++) {
      const int i = 2 * cg + ii;
      const double xn = w.x[((size_t)b * (cf.N + 1) + k) * NX + i] + (double)sm[(OFF_XN + i) * NI];     // fp64 state + the step's increment
      double ci = ok ? xn - ba[i * NZ] * u0 - ba[i * NZ + 1] * u1 : nan("");
      for (int c = 0; c < NX; c++) ci = fma(-ba[i * NZ + 2 + c], x[c], ci);
      w.c[sk * NX + i] = ci;
      w.phi[sk * NX + i] = ok ? xn : nan("");
    }
  }
}

namespace {
template <typename Real>
void launch_irk_instance(int slot, const KCfg& cf, const Work& w, cudaStream_t stream) {
  const int total = cf.B * cf.N;
  const int bytes = PER_INT * NI * (int)sizeof(Real);
  // the dynamic shared-memory limit is a per-device attribute of each kernel instance: remembered per (instance, device)
  static bool configured[2][64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return;
  if (dev < 0 || dev >= 64 || !configured[slot][dev]) {
    if (cudaFuncSetAttribute(linearize_irk_kernel<Real>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) return;   // the launch below then reports the error
    if (dev >= 0 && dev < 64) configured[slot][dev] = true;
  }
  linearize_irk_kernel<Real><<<(total + NI - 1) / NI, NI * NCG, bytes, stream>>>(cf, w);
}
}  // namespace

void launch_linearize_irk(const KCfg& cf, const Work& w, cudaStream_t stream) {
  if (cf.lin_fp32) launch_irk_instance<float>(1, cf, w, stream);
  else launch_irk_instance<double>(0, cf, w, stream);
}

}  // namespace ihm2b
