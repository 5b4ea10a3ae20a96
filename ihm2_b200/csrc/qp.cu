// qp.cu -- batched structured QP: Mehrotra primal-dual interior point with a Riccati KKT
// solve, followed by the SQP-RTI full step.
//
// Replaces HPIPM's d_ocp_qp_ipm_solve (selected by the reference through
// qp_solver = PARTIAL_CONDENSING_HPIPM, new_python/controller.py:300; implementation not in
// the reference tree) and the feedback phase of acados' ocp_nlp_sqp_rti for the OCP of
// python/mpc.py:29-101 (LINEAR_LS cost, box bounds idxbx/idxbu, two rate rows C x + D u).
//
// Mapping: ONE WARP PER PROBLEM, 11 problems per SM (19.2 KB of shared memory, 168 registers).
//  * serial sweeps (Riccati factorisation, forward / backward substitutions, costate recursion) run on
//    the FP64 tensor cores: every 8x8 (x 8 | 10 | 16) product of a stage is one or a few
//    mma.sync.m8n8k4.f64 (DMMA), lane l = (g = l >> 2, t = l & 3) owning the fragments.  The
//    cost-to-go matrix P never leaves its accumulator registers; matrix-vector products repeat the
//    vector in the 8 columns of the right operand.  A dependent DFMA costs 23 cycles on B200, a
//    DMMA 26 (scripts/micro/dmma_lat.cu), so a length-8 contraction drops from ~190 to ~50 cycles
//    and the per-lane operand traffic from 36 shared-memory loads per stage to 4-6.
//    What these sweeps hand over between stages and sweeps (gains, residuals, steps: 58 doubles per
//    stage) stays in shared memory for the whole solve; the [B A] operand fragments are read
//    straight from L2 into registers PF stages ahead of their use.
//  * everything that is independent across stages (inequality rows, residuals, step lengths,
//    updates) runs stage-parallel with lane l = (a = l & 7, q = l >> 3): row / state a of stages
//    q, q+4, q+8, ... directly on global memory (256 contiguous bytes per warp instruction,
//    L2-resident while the problem is in flight), so the divisions and the row logic stay off the
//    serial chains.  Values a pass needs from the other rows of its stage come by 8-lane shuffles.
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

constexpr double TAU = 0.995;
constexpr unsigned FULL = 0xffffffffu;
constexpr int PER_STAGE = 58;   // doubles of shared memory per stage
constexpr int PER_STAGE_NL = PER_STAGE + 20;   // + barrier weights of the dense h rows [4], linearised h rows [16]
#ifndef PF_
#define PF_ 2
#endif
#ifndef QP_MIN_BLOCKS
#define QP_MIN_BLOCKS 12   // problems (warps) per SM the register allocation is held to (168 registers for 9..12: three warps on a
                           // sub-partition; residency is 11 by shared memory -- 12 is the value ptxas allocates without spills at)
#endif
#ifndef QP_MIN_BLOCKS_NL
#define QP_MIN_BLOCKS_NL 8
#endif
#ifndef QP_EXACT_MU
#define QP_EXACT_MU 1e-3
#endif
#ifndef PAR_UNROLL
#define PAR_UNROLL 1
#endif
#define PAR_STR_(x) #x
#define PAR_STR(x) PAR_STR_(x)
#define PAR_LOOP _Pragma(PAR_STR(unroll PAR_UNROLL))
constexpr int PF = PF_;         // [B A] prefetch distance (stages) of the serial sweeps
#ifndef QP_WPB
#define QP_WPB 4                // problems (warps) per block: 3 blocks of 4 independent warps fit an SM (12 problems; one-warp blocks
#endif                          // stop at 11, each block paying the 1 KB of shared memory the system reserves), one warp per scheduler and block

__device__ __forceinline__ double gsum(double v) {   // over the 8 lanes of a segment
  v += __shfl_xor_sync(FULL, v, 1);
  v += __shfl_xor_sync(FULL, v, 2);
  v += __shfl_xor_sync(FULL, v, 4);
  return v;
}
__device__ __forceinline__ double wsum(double v) {   // over the warp
  v = gsum(v);
  v += __shfl_xor_sync(FULL, v, 8);
  v += __shfl_xor_sync(FULL, v, 16);
  return v;
}
// D = A (8x4, row) * B (4x8, col) + D on the FP64 tensor cores.  Lane l holds a = A[l>>2][l&3],
// b = B[l&3][l>>2], (c0, c1) = D[l>>2][2(l&3)], D[l>>2][2(l&3)+1].
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// Predicated shared-memory stores (one lane of the warp writes; no divergent branch on the serial chains)
__device__ __forceinline__ void sts_if(bool p, double* ptr, double v) {
  asm volatile("{.reg .pred p; setp.ne.b32 p, %0, 0; @p st.shared.f64 [%1], %2;}"
               :: "r"((int)p), "r"((unsigned)__cvta_generic_to_shared(ptr)), "d"(v) : "memory");
}
__device__ __forceinline__ void sts2_if(bool p, double* ptr, double v0, double v1) {
  asm volatile("{.reg .pred p; setp.ne.b32 p, %0, 0; @p st.shared.v2.f64 [%1], {%2, %3};}"
               :: "r"((int)p), "r"((unsigned)__cvta_generic_to_shared(ptr)), "d"(v0), "d"(v1) : "memory");
}
// Serial sweep over the stages with the per-lane [B A] operand fragments fetched from global memory (L2)
// PF stages ahead into a rotating register file: DIR = -1 walks k = N-1 .. 0, DIR = +1 walks k = 0 .. N-1.
// NT > 0 with NT % PF == 0 drops the remainder code at compile time.
template <int NT, int PF, int DIR, class Frag, class Load, class Body>
__device__ __forceinline__ void sweep(const int N, Load load, Body body) {
  Frag f[PF];
  const int k0 = (DIR > 0) ? 0 : N - 1;
#pragma unroll
  for (int d = 0; d < PF; d++) f[d] = load(max(0, min(N - 1, k0 + DIR * d)));
  int k = k0, left = N;
#pragma unroll 1
  for (; left >= PF; left -= PF, k += DIR * PF) {
#pragma unroll
    for (int d = 0; d < PF; d++) {
      const Frag cur = f[d];
      f[d] = load(max(0, min(N - 1, k + DIR * (d + PF))));
      body(k + DIR * d, cur);
    }
  }
  if (NT <= 0 || NT % PF != 0) {
#pragma unroll
    for (int d = 0; d < PF; d++)
      if (d < left) body(k + DIR * d, f[d]);
  }
}
struct FragB { double a0, a1, b0, b1, e0a, e0b, e1a, e1b, xa, xb, pa, pb, u0, u1; };
struct ResIn { double xa, pia, p1, p2, l1, l2, u0, u1, c, xn, xj, q0, q1, s0, s1; };
struct RowIn { double lamL, lamU, tL, tU, d1, d2, uv, xv, xa, pia, atp, b0, b1; };
struct Frag4 { double a, b, c, d; };
struct Frag3 { double a, b, c; };
struct Frag2 { double a, b; };
// NaN-propagating max
__device__ __forceinline__ double nmax(double a, double b) { return (b > a || b != b) ? b : a; }
__device__ __forceinline__ double wmax(double v) {
#pragma unroll
  for (int m = 1; m < 32; m <<= 1) v = nmax(v, __shfl_xor_sync(FULL, v, m));
  return v;
}
// The step-length ratio tests use fmax (three instructions against six, NaNs ignored): a NaN in a step reaches the
// slacks with the update, and the residual norms / the barrier parameter of the next iteration -- which do propagate
// it -- end the solve with status 3.
__device__ __forceinline__ double wmaxf(double v) {
#pragma unroll
  for (int m = 1; m < 32; m <<= 1) v = fmax(v, __shfl_xor_sync(FULL, v, m));
  return v;
}

// Inequality row r = a of a stage:  v = cu * u[iu] + cx * x[ix]  in [lb, ub]   (branch-free form)
// python/mpc.py:77-99.  Path stages: (u_T, u_d, u_T - T, u_d - delta, n, v_x, T, delta), rows 4..7
// absent at k = 0 (x_0 is fixed).  Terminal: idxbx_e = [1,3,4,5] with the n / v_x / T / delta
// limits (literal, mpc.py:81-83).
struct RowDef { double cu, cx, lb, ub; int iu, ix; };

__device__ __forceinline__ RowDef row_path(const KCfg& cf, int r) {
  RowDef R;
  R.iu = r & 1;
  R.cu = (r < 4) ? 1.0 : 0.0;
  R.ix = (r == 2 || r == 6) ? 6 : (r == 3 || r == 7) ? 7 : (r == 4) ? 1 : (r == 5) ? 3 : 0;
  R.cx = (r < 2) ? 0.0 : (r < 4) ? -1.0 : 1.0;
  const double bd = (r == 0 || r == 6) ? cf.T_max : (r == 1 || r == 7) ? cf.delta_max : (r == 2) ? cf.g_T_max
                    : (r == 3) ? cf.g_delta_max : (r == 4) ? cf.n_max : cf.v_x_max;
  R.lb = (r == 5) ? 0.0 : -bd; R.ub = bd;
  return R;
}
__device__ __forceinline__ RowDef row_term(const KCfg& cf, int r) {
  RowDef R;
  R.iu = 0; R.cu = 0.0; R.cx = 1.0;
#ifdef IHM2B_EXPERIMENT_TERMINAL_6_7   // experiment only: the terminal box on (n, v_x, T, delta) instead of the literal set
  R.ix = (r == 0) ? 1 : (r == 1) ? 3 : (r == 2) ? 6 : 7;
#else
  R.ix = (r == 0) ? 1 : (r == 1) ? 3 : (r == 2) ? 4 : 5;
#endif
  const double bd = (r == 0) ? cf.n_max : (r == 1) ? cf.v_x_max : (r == 2) ? cf.T_max : cf.delta_max;
  R.lb = -bd; R.ub = bd;
  return R;
}

// Value of a row at (u, x).  Every pass evaluates rows through this one expression: the slack residuals
// v - lb - t are differences of nearly equal numbers near convergence, and the passes must agree on v
// to the last bit (a row value carried along as v + alpha C dz instead stalls the IPM).
__device__ __forceinline__ double row_value(const RowDef& R, bool term, double u, double x) {
  return fma(R.cx, x, term ? 0.0 : R.cu * u);
}

// C' q seen by state a:  c1 * q[i1] + c2 * q[i2]   (diagonal of C' diag(q) C: c1 q[i1] + |c2| q[i2])
struct CtDef { int i1, i2; double c1, c2; };
__device__ __forceinline__ CtDef ct_path(int a) {
  CtDef d;
  d.i1 = (a == 1) ? 4 : (a == 3) ? 5 : (a == 6) ? 6 : 7; d.c1 = (a == 1 || a == 3 || a == 6 || a == 7) ? 1.0 : 0.0;
  d.i2 = (a == 6) ? 2 : 3; d.c2 = (a == 6 || a == 7) ? -1.0 : 0.0;
  return d;
}
__device__ __forceinline__ CtDef ct_term(int a) {
  CtDef d;
#ifdef IHM2B_EXPERIMENT_TERMINAL_6_7
  d.i1 = (a == 1) ? 0 : (a == 3) ? 1 : (a == 6) ? 2 : 3; d.c1 = (a == 1 || a == 3 || a == 6 || a == 7) ? 1.0 : 0.0;
#else
  d.i1 = (a == 1) ? 0 : (a == 3) ? 1 : (a == 4) ? 2 : 3; d.c1 = (a == 1 || a == 3 || a == 4 || a == 5) ? 1.0 : 0.0;
#endif
  d.i2 = 0; d.c2 = 0.0;
  return d;
}


// ---------------------------------------------------------------------------------------
// Soft rows (NL): a soft two-sided row  lb - sL <= v <= ub + sU,  sL, sU >= 0  with the penalty
// z (sL + sU) + Z/2 (sL^2 + sU^2) (acados idxsbx / idxsh with zl = zu, Zl = Zu:
// old/generate_acaods_interface.py:383-398, 413, 417, 438, 449).  The slack pair and the multipliers of
// s >= 0 are eliminated row by row from the Newton system (HPIPM does the same with its Zs_inv vectors), which only
// changes the barrier weight and the first-order term the row hands to the Riccati recursion:
//   D = Z + lam/t + lamS/s,   gamma_eff = (lam/t)(Z + lamS/s)/D,   q_eff = q - (lam/t)(r_s + q + q_S)/D.
// ---------------------------------------------------------------------------------------
struct RowSt { double lamL, lamU, tL, tU, sL, sU, lsL, lsU; };   // state of one row (s, ls: soft rows only)
struct RowA { double gam, psi, dl; };
struct Norms { double rd, rm, rs, S0; };

// residual-side quantities of one row at value v: barrier weight gam, first-order term psi (= q_eff - lam, lower minus
// upper), multiplier difference dl; accumulates the residual norms and the complementarity sum
__device__ __forceinline__ RowA soft_row_A(const bool act, const bool soft, const double v, const double lb, const double ub,
                                           const double isc, const double Z, const double zz, const RowSt& r, Norms& nm) {
  RowA o{0.0, 0.0, 0.0};
  if (act) {
    const double itL = 1.0 / r.tL, itU = 1.0 / r.tU;
    const double rdL = soft ? v + r.sL - lb - r.tL : v - lb - r.tL;
    const double rdU = soft ? ub + r.sU - v - r.tU : ub - v - r.tU;
    double gL = r.lamL * itL, gU = r.lamU * itU;
    double pL = r.lamL * rdL * itL, pU = r.lamU * rdU * itU;
    nm.rd = nmax(nm.rd, nmax(fabs(rdL), fabs(rdU)) * isc);
    nm.rm = nmax(nm.rm, nmax(r.lamL * r.tL, r.lamU * r.tU));
    nm.S0 += r.lamL * r.tL + r.lamU * r.tU;
    if (soft) {
      const double gSL = r.lsL / r.sL, gSU = r.lsU / r.sU;
      const double rsL = Z * r.sL + zz - r.lamL - r.lsL, rsU = Z * r.sU + zz - r.lamU - r.lsU;
      nm.rs = nmax(nm.rs, nmax(fabs(rsL), fabs(rsU)));
      nm.rm = nmax(nm.rm, nmax(r.lsL * r.sL, r.lsU * r.sU));
      nm.S0 += r.lsL * r.sL + r.lsU * r.sU;
      const double iDL = 1.0 / (Z + gL + gSL), iDU = 1.0 / (Z + gU + gSU);
      pL -= gL * (rsL + (r.lamL + pL) + r.lsL) * iDL;
      pU -= gU * (rsU + (r.lamU + pU) + r.lsU) * iDU;
      gL *= (Z + gSL) * iDL; gU *= (Z + gSU) * iDU;
    }
    o.gam = gL + gU; o.psi = pL - pU; o.dl = r.lamL - r.lamU;
  }
  return o;
}

// Newton step of one row for a given primal step cdz = C dz of the row.  cL, cU / cSL, cSU: second-order corrections of
// the complementarity products (0 in the predictor, dlam_aff dt_aff - mu_t in the corrector).
struct RowStep { double dlL, dlU, dtL, dtU, dsL, dsU, dlsL, dlsU; };
__device__ __forceinline__ RowStep soft_row_step(const bool soft, const double v, const double lb, const double ub, const double cdz,
                                                 const double Z, const double zz, const RowSt& r, const double cL, const double cU,
                                                 const double cSL, const double cSU) {
  RowStep o;
  const double itL = 1.0 / r.tL, itU = 1.0 / r.tU;
  const double rdL = soft ? v + r.sL - lb - r.tL : v - lb - r.tL;
  const double rdU = soft ? ub + r.sU - v - r.tU : ub - v - r.tU;
  o.dsL = 0.0; o.dsU = 0.0; o.dlsL = 0.0; o.dlsU = 0.0;
  if (soft) {
    const double isL = 1.0 / r.sL, isU = 1.0 / r.sU;
    const double gL = r.lamL * itL, gU = r.lamU * itU;
    const double qL = r.lamL + (cL + r.lamL * rdL) * itL, qU = r.lamU + (cU + r.lamU * rdU) * itU;
    const double wL = Z * r.sL + zz - r.lamL - r.lsL + qL + r.lsL + cSL * isL;
    const double wU = Z * r.sU + zz - r.lamU - r.lsU + qU + r.lsU + cSU * isU;
    const double iDL = 1.0 / (Z + gL + r.lsL * isL), iDU = 1.0 / (Z + gU + r.lsU * isU);
    o.dsL = -(wL + gL * cdz) * iDL;
    o.dsU = -(wU - gU * cdz) * iDU;
    o.dlsL = -r.lsL - (cSL + r.lsL * o.dsL) * isL;
    o.dlsU = -r.lsU - (cSU + r.lsU * o.dsU) * isU;
  }
  o.dtL = cdz + o.dsL + rdL; o.dtU = -cdz + o.dsU + rdU;
  o.dlL = -r.lamL - (cL + r.lamL * o.dtL) * itL;
  o.dlU = -r.lamU - (cU + r.lamU * o.dtU) * itU;
  return o;
}

// Row a of h (a < 5) as seen by the kernels that work on global memory (SQP residuals, merit, cost): bounds, the
// position of the row's coefficients in a stage's 16 doubles of Work::HC, and what state a receives from the h rows.
struct HLane {
  int i0, i1, i2, co, eo, r1, r2, so1;
  bool dense, rate;
  double ratec, lb, ub, hc2r;
  __device__ HLane(const KCfg& cf, int a, double itT, double itD) {
    ub = (a < 2) ? 0.0 : (a == 2) ? cf.T_dot_max : (a == 3) ? cf.delta_dot_max : cf.a_lat_max;
    lb = (a < 2) ? -1e3 : -ub;
    i0 = (a < 2) ? 0 : (a == 2) ? 6 : (a == 3) ? 7 : 3; i1 = (a < 2) ? 1 : 4; i2 = (a < 2) ? 2 : 6;
    dense = (a < 2) || (a == 4); rate = (a == 2 || a == 3);
    co = (a == 0) ? 0 : (a == 1) ? 4 : 8;
    eo = (a == 0) ? 3 : (a == 1) ? 7 : (a == 2) ? 13 : (a == 3) ? 14 : (a == 4) ? 12 : 15;
    ratec = (a == 2) ? itT : (a == 3) ? itD : 0.0;
    r1 = (a < 3) ? 0 : 4; r2 = (a < 3) ? 1 : (a == 6) ? 2 : 3;
    so1 = (a < 3) ? a : (a == 3) ? 8 : (a == 4) ? 9 : (a == 6) ? 10 : (a == 7) ? 11 : 15;
    hc2r = (a == 6) ? -itT : (a == 7) ? -itD : 0.0;
  }
  // value of the row at (x, u) of its stage (the same expression as the QP kernel's hrow_val)
  __device__ double value(const double* __restrict__ h, const double* __restrict__ x, double u0, double u1, int a) const {
    const double c0 = dense ? h[co] : -ratec, c1 = dense ? h[co + 1] : ratec;
    const double c2 = dense ? h[co + 2] : 0.0, c3 = (a == 4) ? h[co + 3] : 0.0;
    const double X1 = (a == 2) ? u0 : (a == 3) ? u1 : x[i1];
    return fma(c0, x[i0], fma(c1, X1, fma(c2, x[i2], fma(c3, x[7], h[eo]))));
  }
};
// value of a soft row's slack penalty with the slack eliminated: the cheapest feasible slack of a row at value v is
// its violation, so the NLP in (x, u, s) equals the NLP in (x, u) with cost + sum z viol + Z/2 viol^2
__device__ __forceinline__ double soft_penalty(double zs, double Zs, double viol) { return viol * (zs + 0.5 * Zs * viol); }

// corrector right-hand side of one row as chi = A - mu_t B (A from the second-order products d, B from the centring)
__device__ __forceinline__ void soft_row_chi(const bool soft, const double Z, const RowSt& r, const double dL, const double dU,
                                             const double dSL, const double dSU, double& A, double& B) {
  const double itL = 1.0 / r.tL, itU = 1.0 / r.tU;
  double aL = dL * itL, bL = itL, aU = dU * itU, bU = itU;
  if (soft) {
    const double isL = 1.0 / r.sL, isU = 1.0 / r.sU;
    const double gL = r.lamL * itL, gU = r.lamU * itU;
    const double fL = gL / (Z + gL + r.lsL * isL), fU = gU / (Z + gU + r.lsU * isU);
    aL = aL * (1.0 - fL) - fL * dSL * isL; bL = bL * (1.0 - fL) - fL * isL;
    aU = aU * (1.0 - fU) - fU * dSU * isU; bU = bU * (1.0 - fU) - fU * isU;
  }
  A = aL - aU; B = bL - bU;
}

}  // namespace

// ---------------------------------------------------------------------------------------
// grid = ceil(B / WPB), block = 32 WPB (independent warps, one problem each), dynamic shared memory = (N+1) * PER_STAGE
// WPB = QP_WPB for the N = 40 instances, 1 for the instances with the horizon as a run-time value (a long horizon would
// not fit four problems into the shared memory of a block: N up to 480 hard-bounded / 360 with soft rows stays possible).
// doubles per problem.
// x0 [B][8]; on exit the iterate w.x / w.u holds the QP solution (RTI full step).
// ---------------------------------------------------------------------------------------
// NL: soft nonlinear constraints h(x, u) and soft bounds (16 row slots per stage instead of 8, see Work::HC)
template <int NT, bool NL, int WPB>  // NT > 0: horizon known at compile time (all shared-memory offsets fold into immediates)
__global__ void __launch_bounds__(32 * WPB, (NL ? QP_MIN_BLOCKS_NL : QP_MIN_BLOCKS) / WPB) qp_rti_kernel(const KCfg cf, const Work w,
                                                     const double* __restrict__ x0,
                                                     double* __restrict__ u0_out,
                                                     int32_t* __restrict__ status_out) {
  extern __shared__ double smem_block[];
  const int lane = threadIdx.x & 31;
  const int a = lane & 7, q = lane >> 3;
  const int slot = blockIdx.x * WPB + (threadIdx.x >> 5);
  if (slot >= cf.B) return;
  const int b = w.order ? w.order[slot] : slot;
  if (w.active && !w.active[b]) return;      // SQP mode: instance already converged or failed
  const int N = (NT > 0) ? NT : cf.N, N1 = N + 1;
  double* const smem = smem_block + (threadIdx.x >> 5) * (N1 * (NL ? PER_STAGE_NL : PER_STAGE));
  // Shared memory (structure of arrays, stage-major): what the serial sweeps touch.
  double* const pGAM = smem;             // [N1][8]   gamma -> chi -> C'(dlamL - dlamU) per state
  double* const pGX = pGAM + N1 * 8;     // [N1][8]   Newton gradient (state part), later dx
  double* const pGU = pGX + N1 * 8;      // [N1][2]   Newton gradient (control part), later du
  double* const pRG = pGU + N1 * 2;      // [N1][8]   stationarity residual (state part), later d pi
  double* const pRB = pRG + N1 * 8;      // [N1][8]   dynamics residual
  double* const pK = pRB + N1 * 8;       // [N1][8][2]
  double* const pKF = pK + N1 * 16;      // [N1][4]   kff[2], kfc[2]
  double* const pMI = pKF + N1 * 4;      // [N1][4]   i00 i01 i11 pad
  double* const pGH = pMI + N1 * 4;      // NL: [N1][4]  barrier weights of the rows h0, h1, h4 (and 0)
  double* const pHC = pGH + (NL ? N1 * 4 : 0);   // NL: [N1][16] linearised h rows (Work::HC)
  constexpr int RS = NL ? 16 : 8;        // row slots per stage
  // Global memory (L2-resident while the problem is in flight): what only the stage-parallel
  // phases touch, accessed as 256 contiguous bytes per warp instruction; and [B A], whose operand
  // fragments the serial sweeps load into registers PF stages ahead of their use.
  const double* __restrict__ const gBA = w.BA + (size_t)b * N * 80;   // [N][8][10]  (B | A) row-major
  const double* __restrict__ const pC = w.c + (size_t)b * N * 8;      // [N][8]
  double* __restrict__ const pX = w.x + (size_t)b * N1 * 8;           // [N1][8]   iterate, updated in place
  double* __restrict__ const pU = w.u + (size_t)b * N * 2;            // [N][2]
  double* __restrict__ const pPI = w.xa + (size_t)b * N1 * 8;         // [N1][8]   costate
  double* __restrict__ const pATP = w.atp + (size_t)b * N * 8;        // [N][8]    A_k' pi_{k+1}, carried (phase B)
  double* __restrict__ const pLL = w.LT + (size_t)b * N1 * 4 * RS;    // [N1][RS]  lamL, lamU, tL, tU
  double* __restrict__ const pLU = pLL + N1 * RS;
  double* __restrict__ const pTL = pLU + N1 * RS;
  double* __restrict__ const pTU = pTL + N1 * RS;
  double* __restrict__ const pD1 = w.D + (size_t)b * N1 * 2 * RS;     // [N1][RS]  psi -> dlam_aff*dt_aff (L) -> dlamL
  double* __restrict__ const pD2 = pD1 + N1 * RS;                     // [N1][RS]  dl  -> dlam_aff*dt_aff (U) -> dlamU
  double* const pDX = pGX; double* const pDU = pGU;

  // ------------------------------------------------------------------ load the problem
  double mag_b = 0.0;
  {
    for (int e = lane; e < N * 8; e += 32) mag_b = fmax(mag_b, fabs(pC[e]));
    mag_b = wmax(mag_b);
  }
  const double xh = x0[(size_t)b * 8 + a];  // measured state, component a
  const double s0 = __shfl_sync(FULL, xh, 0, GROUP);
  __syncwarp();
  if (q == 0) pX[a] = xh;                   // x_0 of the iterate is the measured state
  // cost (python/mpc.py:47-69, controller.py:320-385): constant near-diagonal GN Hessian
  const double hx = cf.cs * (cf.W[a] + (a == 6 ? cf.W[10] : a == 7 ? cf.W[11] : 0.0));
  const double hxN = cf.ct * cf.We[a];
  const double hu0 = cf.cs * (cf.W[8] + cf.W[10]), hu1 = cf.cs * (cf.W[9] + cf.W[11]);
  const double c10 = cf.cs * cf.W[10], c11 = cf.cs * cf.W[11];
  const double hxu0 = (a == 6) ? -c10 : 0.0;  // H[x_a][u_T]
  const double hxu1 = (a == 7) ? -c11 : 0.0;  // H[x_a][u_delta]
  const double gs_scale = -cf.cs * cf.W[0];   // g[s]_k = gs_scale * yref_k
  const double gsN = -cf.ct * cf.We[0] * (s0 + cf.s_target);
  const double ds_ref = cf.s_target / N;
  const double mag_g = fmax(fmax(fabs(gs_scale * s0), fabs(gs_scale * (s0 + ds_ref * (N - 1)))), fabs(gsN));
  const RowDef Rp = row_path(cf, a), Rt = row_term(cf, a);
  // slack residuals are measured relative to the size of the row's bounds (the rows span 0.02 ... 1e6)
  const double iscP = 1.0 / (1.0 + fmin(fabs(Rp.lb), fabs(Rp.ub))), iscT = 1.0 / (1.0 + fmin(fabs(Rt.lb), fabs(Rt.ub)));
  const CtDef ctP = ct_path(a), ctT = ct_term(a);
  const int rounds = (N + 4) >> 2;           // stage-parallel loops: stage k = q + 4 i
  // ---- NL: row slot 8 + a of a stage is row a of h, all soft: a = 0, 1 the track boundaries, a = 4 a_lat (path stages).
  // Rows 2, 3 of h (T_dot, delta_dot) are the linear rows 2, 3 above (u - x within +-t x_dot_max): kept once, in that
  // form; their slots stay empty.  Of the static rows T and delta (terminal: the rows on v_y and r) are soft
  // (idxsbx = [2, 3] of idxbx / idxbx_e)
  double itT = 0.0, itD = 0.0;
  if (NL) { itT = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TT]; itD = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TD]; }
  const double Zp = cf.cs * cf.slack_Z, zp = cf.cs * cf.slack_z, Zt = cf.ct * cf.slack_Z, zt = cf.ct * cf.slack_z;
  const double hub = (a < 2) ? 0.0 : (a == 2) ? cf.T_dot_max : (a == 3) ? cf.delta_dot_max : cf.a_lat_max;
  const double hlb = (a < 2) ? -1e3 : -hub;
  const double isc1 = 1.0 / (1.0 + fmin(fabs(hlb), fabs(hub)));
  const int hi0 = (a < 2) ? 0 : (a == 2) ? 6 : (a == 3) ? 7 : 3, hi1 = (a < 2) ? 1 : 4, hi2 = (a < 2) ? 2 : 6;
  const bool hdense = (a < 2) || (a == 4), hrate = (a == 2 || a == 3);
  const int hco = (a == 0) ? 0 : (a == 1) ? 4 : 8;                 // the row's coefficients in pHC (dense rows)
  const int heo = (a == 0) ? 3 : (a == 1) ? 7 : (a == 2) ? 13 : (a == 3) ? 14 : (a == 4) ? 12 : 15;
  const double ratec = (a == 2) ? itT : (a == 3) ? itD : 0.0;      // h2, h3: -ratec on T / delta, +ratec on the control
  // what state a receives from the h rows: coefficient pHC[so1] of row hr1, and pHC[4 + a] of h1 (a < 3) or the
  // constant -1/t of the rate row hr2 (a = 6, 7)
  const int hr1 = (a < 3) ? 0 : 4, hr2 = (a < 3) ? 1 : (a == 6) ? 2 : 3;
  const int so1 = (a < 3) ? a : (a == 3) ? 8 : (a == 4) ? 9 : (a == 6) ? 10 : 11;
  const double hc2r = (a == 6) ? -itT : (a == 7) ? -itD : 0.0;
  struct HRow { double c0, c1, c2, c3, e; };
  auto hrow_coef = [&](int k) {
    const double* h = pHC + k * 16;
    HRow r;
    r.c0 = hdense ? h[hco] : -ratec; r.c1 = hdense ? h[hco + 1] : ratec;
    r.c2 = hdense ? h[hco + 2] : 0.0; r.c3 = (a == 4) ? h[hco + 3] : 0.0;
    r.e = h[heo];
    return r;
  };
  // value of h row a at the stage whose state component a this lane holds in xa (u0v, u1v: the stage's controls)
  auto hrow_val = [&](const HRow& r, double xa_, double u0v, double u1v) {
    const double X0 = __shfl_sync(FULL, xa_, hi0, GROUP), X1s = __shfl_sync(FULL, xa_, hi1, GROUP);
    const double X2 = __shfl_sync(FULL, xa_, hi2, GROUP), X3 = __shfl_sync(FULL, xa_, 7, GROUP);
    const double X1 = (a == 2) ? u0v : (a == 3) ? u1v : X1s;
    return fma(r.c0, X0, fma(r.c1, X1, fma(r.c2, X2, fma(r.c3, X3, r.e))));
  };
  auto hrow_cdz = [&](const HRow& r, int k) {
    const double X1 = hrate ? pGU[k * 2 + (a - 2)] : pGX[k * 8 + hi1];
    return fma(r.c0, pGX[k * 8 + hi0], fma(r.c1, X1, fma(r.c2, pGX[k * 8 + hi2], r.c3 * pGX[k * 8 + 7])));
  };
  // SS arrays (NL)
  // a lane has at most one soft row per stage (path: lanes 0, 1, 4 their h row, lanes 6, 7 the T / delta bound;
  // terminal: lanes 0, 1 their h row, lanes 2, 3 the v_y / r bound): the slack arrays carry 8 slots per stage, slot = lane
  double* __restrict__ const pSL = NL ? w.SS + (size_t)b * N1 * 8 * 8 : nullptr;   // [N1][8] each
  double* __restrict__ const pSU = pSL + N1 * 8;
  double* __restrict__ const pLSL = pSU + N1 * 8;
  double* __restrict__ const pLSU = pLSL + N1 * 8;
  double* __restrict__ const pDS1 = pLSU + N1 * 8;   // dlamS_aff * ds_aff (L) -> dlamSL
  double* __restrict__ const pDS2 = pDS1 + N1 * 8;
  double* __restrict__ const pE1 = pDS2 + N1 * 8;    // dsL, dsU of the corrected step
  double* __restrict__ const pE2 = pE1 + N1 * 8;
  double* __restrict__ const pCA = NL ? w.CAB + (size_t)b * N1 * 2 * 16 : nullptr;  // [N1][16]: [a] state a, [8 + a] control a
  double* __restrict__ const pCB = pCA + N1 * 16;
  if (NL) {
    const double* __restrict__ gHC = w.HC + (size_t)b * N1 * 16;
    for (int e = lane; e < N1 * 16; e += 32) pHC[e] = gHC[e];
    for (int e = lane; e < N1 * 4; e += 32) pGH[e] = 0.0;
  }
  __syncwarp();

  // ------------------------------------------------------------------ initialisation
  // Row pass "A" of one (stage, row): barrier Hessian weight gamma, first-order terms psi / dl for the
  // residuals, residual norms; leaves what the Riccati sweep adds to M ready to load:
  // pGAM[k][j] = diagonal of H~_xx, pMI[k] = (M_uu diagonal addends, M_xu addends of rows T and delta).
  // Called by all 32 lanes (segment shuffles inside).
  double rd_m = 0.0, rm_m = 0.0, S0 = 0.0;
  auto rows_A = [&](int k, bool on, bool term, bool act, const RowDef& R, double v, double lamL, double lamU,
                    double tL, double tU) {
    double gam = 0.0, psi = 0.0, dl = 0.0;
    if (act) {
      const double itL = 1.0 / tL, itU = 1.0 / tU;
      const double rdL = v - R.lb - tL, rdU = R.ub - v - tU;
      gam = lamL * itL + lamU * itU;
      psi = lamL * rdL * itL - lamU * rdU * itU;
      dl = lamL - lamU;
      rd_m = nmax(rd_m, nmax(fabs(rdL), fabs(rdU)) * (term ? iscT : iscP));
      rm_m = nmax(rm_m, nmax(lamL * tL, lamU * tU));
      S0 += lamL * tL + lamU * tU;
    }
    const CtDef& ct = term ? ctT : ctP;
    const double g1 = __shfl_sync(FULL, gam, ct.i1, GROUP), g2 = __shfl_sync(FULL, gam, ct.i2, GROUP);
    const double gp2 = __shfl_sync(FULL, gam, (a + 2) & 7, GROUP);
    if (on) {
      pGAM[k * 8 + a] = (term ? hxN : hx) + ct.c1 * g1 + fabs(ct.c2) * g2;
      pD1[k * 8 + a] = psi; pD2[k * 8 + a] = dl;
      if (a < 4 && !term) pMI[k * 4 + a] = (a < 2) ? ((a == 0 ? hu0 : hu1) + gam + gp2) : (-(a == 2 ? c10 : c11) - gam);
    }
  };
  // NL form of the row pass: both row slots of the lane, soft rows with their slacks eliminated; the rows' first-order
  // terms are handed over per state / control (pD1 / pD2 [k][a] state a, [k][8 + a] control a), the barrier weights
  // of the dense rows h0, h1, h4 go to pGH for the Riccati sweep (their C' gam C is not diagonal)
  double rs_m = 0.0;
  auto rows_A_nl = [&](int k, bool on, bool term, bool act0, double v0, const RowSt& r0, bool act1, double v1, const RowSt& r1) {
    const RowDef& R = term ? Rt : Rp;
    const double Z = term ? Zt : Zp, zz = term ? zt : zp;
    const bool soft0 = term ? (a == 2 || a == 3) : (a >= 6);
    Norms nm{rd_m, rm_m, rs_m, S0};
    // a lane has at most one soft row: its h row if it has one (then its static row is hard), else its static row
    const double iscR = term ? iscT : iscP;
    const RowSt& rS = act1 ? r1 : r0;
    const RowA AS = soft_row_A(act1 || act0, act1 || soft0, act1 ? v1 : v0, act1 ? hlb : R.lb, act1 ? hub : R.ub,
                               act1 ? isc1 : iscR, Z, zz, rS, nm);
    const RowA AH = soft_row_A(act1 && act0, false, v0, R.lb, R.ub, iscR, Z, zz, r0, nm);
    const RowA A0 = act1 ? AH : AS, A1 = act1 ? AS : RowA{0.0, 0.0, 0.0};
    rd_m = nm.rd; rm_m = nm.rm; rs_m = nm.rs; S0 = nm.S0;
    const CtDef& ct = term ? ctT : ctP;
    const double* h = pHC + k * 16;
    const double hc1 = (a != 5) ? h[so1] : 0.0, hc2 = (a < 3) ? h[4 + a] : hc2r;
    const int ap2 = (a + 2) & 7;
    const double g1 = __shfl_sync(FULL, A0.gam, ct.i1, GROUP), g2 = __shfl_sync(FULL, A0.gam, ct.i2, GROUP), gp2 = __shfl_sync(FULL, A0.gam, ap2, GROUP);
    const double p1 = __shfl_sync(FULL, A0.psi, ct.i1, GROUP), p2 = __shfl_sync(FULL, A0.psi, ct.i2, GROUP), pp2 = __shfl_sync(FULL, A0.psi, ap2, GROUP);
    const double l1 = __shfl_sync(FULL, A0.dl, ct.i1, GROUP), l2 = __shfl_sync(FULL, A0.dl, ct.i2, GROUP), lp2 = __shfl_sync(FULL, A0.dl, ap2, GROUP);
    const double hg2 = __shfl_sync(FULL, A1.gam, hr2, GROUP), ug = __shfl_sync(FULL, A1.gam, ap2, GROUP);
    const double hp1 = __shfl_sync(FULL, A1.psi, hr1, GROUP), hp2 = __shfl_sync(FULL, A1.psi, hr2, GROUP), up = __shfl_sync(FULL, A1.psi, ap2, GROUP);
    const double hl1 = __shfl_sync(FULL, A1.dl, hr1, GROUP), hl2 = __shfl_sync(FULL, A1.dl, hr2, GROUP), ul = __shfl_sync(FULL, A1.dl, ap2, GROUP);
    if (on) {
      pGAM[k * 8 + a] = (term ? hxN : hx) + ct.c1 * g1 + fabs(ct.c2) * g2 + ((a >= 6) ? hc2 * hc2 * hg2 : 0.0);
      pD1[k * 16 + a] = ct.c1 * p1 + ct.c2 * p2 + hc1 * hp1 + hc2 * hp2;
      pD2[k * 16 + a] = ct.c1 * l1 + ct.c2 * l2 + hc1 * hl1 + hc2 * hl2;
      if (!term) {
        if (a < 2) {
          const double rc = (a == 0) ? itT : itD;
          pD1[k * 16 + 8 + a] = A0.psi + pp2 + rc * up;
          pD2[k * 16 + 8 + a] = A0.dl + lp2 + rc * ul;
          pMI[k * 4 + a] = (a == 0 ? hu0 : hu1) + A0.gam + gp2 + rc * rc * ug;
        } else if (a < 4) {
          pMI[k * 4 + a] = -(a == 2 ? c10 : c11) - A0.gam - ratec * ratec * A1.gam;
        }
      }
      if (a < 2) pGH[k * 4 + a] = A1.gam;
      if (a == 4) pGH[k * 4 + 2] = A1.gam;
    }
  };
  // slacks / multipliers cold (t = max(distance, thr0), lam = mu0 / t), pi = 0
  if constexpr (NL) {
    PAR_LOOP for (int i = 0; i < rounds; i++) {
      const int kk = q + 4 * i;
      const bool on = (kk <= N);
      const int k = on ? kk : N;
      const bool term = (k == N);
      const RowDef& R = term ? Rt : Rp;
      const bool act0 = on && (term ? (a < 4) : (a < 4 || k > 0)), act1 = on && (a < 2 || (a == 4 && !term));
      const bool soft0 = term ? (a == 2 || a == 3) : (a >= 6);
      const double xa_ = pX[k * 8 + a], u0v = pU[(term ? 0 : k) * 2], u1v = pU[(term ? 0 : k) * 2 + 1];
      const double v0 = row_value(R, term, R.iu ? u1v : u0v, __shfl_sync(FULL, xa_, R.ix, GROUP));
      const HRow hr = hrow_coef(k);
      const double v1 = hrow_val(hr, xa_, u0v, u1v);
      RowSt r0{0.0, 0.0, 1.0, 1.0, 1.0, 1.0, 0.0, 0.0}, r1 = r0;
      // the slacks start where they absorb the row's violation at the initial point (t = thr0 there): the soft rows
      // are primal feasible from the first iterate on, whatever the warm start violates
      if (act0) {
        if (soft0) {
          r0.sL = fmax(cf.thr0, R.lb - v0 + cf.thr0); r0.sU = fmax(cf.thr0, v0 - R.ub + cf.thr0);
          r0.lsL = cf.mu0 / r0.sL; r0.lsU = cf.mu0 / r0.sU;
        }
        r0.tL = fmax(soft0 ? v0 + r0.sL - R.lb : v0 - R.lb, cf.thr0); r0.tU = fmax(soft0 ? R.ub + r0.sU - v0 : R.ub - v0, cf.thr0);
        r0.lamL = cf.mu0 / r0.tL; r0.lamU = cf.mu0 / r0.tU;
      }
      if (act1) {
        r1.sL = fmax(cf.thr0, hlb - v1 + cf.thr0); r1.sU = fmax(cf.thr0, v1 - hub + cf.thr0);
        r1.lsL = cf.mu0 / r1.sL; r1.lsU = cf.mu0 / r1.sU;
        r1.tL = fmax(v1 + r1.sL - hlb, cf.thr0); r1.tU = fmax(hub + r1.sU - v1, cf.thr0);
        r1.lamL = cf.mu0 / r1.tL; r1.lamU = cf.mu0 / r1.tU;
      }
      if (on) {
        const int e0 = k * 16 + a, e1 = e0 + 8;
        pLL[e0] = r0.lamL; pLU[e0] = r0.lamU; pTL[e0] = r0.tL; pTU[e0] = r0.tU;
        pLL[e1] = r1.lamL; pLU[e1] = r1.lamU; pTL[e1] = r1.tL; pTU[e1] = r1.tU;
        {
          const RowSt& rs = act1 ? r1 : r0;      // the lane's soft row, if it has one (defaults otherwise)
          const int es = k * 8 + a;
          pSL[es] = rs.sL; pSU[es] = rs.sU; pLSL[es] = rs.lsL; pLSU[es] = rs.lsU;
        }
        pPI[k * 8 + a] = 0.0;
      }
      rows_A_nl(k, on, term, act0, v0, r0, act1, v1, r1);
    }
  } else
  PAR_LOOP for (int i = 0; i < rounds; i++) {
    const int kk = q + 4 * i;
    const bool on = (kk <= N);
    const int k = on ? kk : N;
    const bool term = (k == N);
    const RowDef& R = term ? Rt : Rp;
    const bool act = on && (term ? (a < 4) : (a < 4 || k > 0));
    const double v = row_value(R, term, pU[(term ? 0 : k) * 2 + R.iu], pX[k * 8 + R.ix]);
    double lamL = 0.0, lamU = 0.0, tL = 1.0, tU = 1.0;
    if (act) {
      tL = fmax(v - R.lb, cf.thr0); tU = fmax(R.ub - v, cf.thr0);
      lamL = cf.mu0 / tL; lamU = cf.mu0 / tU;
    }
    if (on) {
      pLL[k * 8 + a] = lamL; pLU[k * 8 + a] = lamU; pTL[k * 8 + a] = tL; pTU[k * 8 + a] = tU;
      pPI[k * 8 + a] = 0.0;
    }
    rows_A(k, on, term, act, R, v, lamL, lamU, tL, tU);
  }
  __syncwarp();
  // one-sided inequalities: 2 * (4 + 8 (N-1) + 4); NL: 20 at stage 0, 32 on the path, 20 at the end (soft rows count twice)
  const double inv_n = NL ? 1.0 / (32.0 * N + 8.0) : 1.0 / (16.0 * N);
  double mu = 0.0, res_g = 0.0, res_b = 0.0, res_d = 0.0, res_m = 0.0;
  int it = 0, qp_status = 1, at_floor = 0, short_steps = 0;
  const double mu_floor = 0.1 * cf.tol_comp;
  double prev_res_g = __longlong_as_double(0x7ff0000000000000LL);
  double mu_hist = 0.0;   // lane i < 8 keeps the barrier parameter of the last iteration with it % 8 == i
  double om_alpha = 0.0;  // 1 - (primal step length of the last iteration)

#ifdef IHM2B_PHASE_TIMERS
  long long tph[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tlast = clock64();
#define TICK(i) { const long long t_ = clock64(); tph[i] += t_ - tlast; tlast = t_; }
#else
#define TICK(i)
#endif
  while (true) {
    __syncwarp();
    TICK(0)
    // ================= B: stationarity and dynamics residuals of all stages
    S0 = wsum(S0);
    mu = S0 * inv_n;
    // the carried products below are exact up to rounding that accumulates over the iterations; the first iteration and
    // the second half of a solve (barrier parameter below QP_EXACT_MU: thin interiors make the iterates sensitive to the last
    // bits there -- with the carried products a borderline problem can end on the other side of a failure than the oracle --
    // and the residuals are
    // compared with the tolerances) form them anew
    const bool exact = (it == 0) || !(mu > QP_EXACT_MU);
    // B1, first iteration and end game only: the three [B A] products of every stage, A x + B u, A' pi+, B' pi+, as DMMA
    // matrix-vector products (no dependence between stages: the tensor pipe runs back to back); results parked in
    // the rb / rg / kfc slots.  Later iterations carry them along instead: the Newton step satisfies its own
    // recursions exactly, dx+ = A dx + B du + rb (forward sweep) and dpi = rhs + A' dpi+ (costate sweep), so
    //   rb <- (1 - alpha) rb,   A' pi+ <- A' pi+ + alpha_d A' dpi+   (the product the costate sweep forms anyway),
    // and B' pi+ (16 products per stage) is redone next to the costate update in phase J.  Nothing of the
    // solve error of the factorisation enters: the control rows of the stationarity residual, the only ones the
    // recursions do not pin, are still evaluated from the new point.
    if (exact) {
      const int g = lane >> 2, t = lane & 3;
      const int ra = (2 * t) * 10;
      const bool lo = (g < 2);
      auto load_B = [&](int k) {
        const double* p = gBA + k * 80;
        const double2 af = *reinterpret_cast<const double2*>(p + g * 10 + 2 + 2 * t);
        const double2 bd = *reinterpret_cast<const double2*>(p + g * 10);
        const double2 xf = *reinterpret_cast<const double2*>(pX + k * 8 + 2 * t);
        const double2 pf = *reinterpret_cast<const double2*>(pPI + (k + 1) * 8 + 2 * t);
        const double2 uf = *reinterpret_cast<const double2*>(pU + k * 2);
        return FragB{af.x, af.y, bd.x, bd.y, p[ra + 2 + g], p[ra + 12 + g], p[ra + (g & 1)], p[ra + 10 + (g & 1)],
                     xf.x, xf.y, pf.x, pf.y, uf.x, uf.y};
      };
      auto products = [&](const int k, const FragB f) {
        double R0 = fma(f.b0, f.u0, f.b1 * f.u1), R1 = R0, T0 = 0.0, T1 = 0.0, V0 = 0.0, V1 = 0.0;
        const double e1a = lo ? f.e1a : 0.0, e1b = lo ? f.e1b : 0.0;
        dmma(R0, R1, f.a0, f.xa); dmma(T0, T1, f.e0a, f.pa); dmma(V0, V1, e1a, f.pa);
        dmma(R0, R1, f.a1, f.xb); dmma(T0, T1, f.e0b, f.pb); dmma(V0, V1, e1b, f.pb);
        sts_if(t == 0, pRB + k * 8 + g, R0);                 // (A x + B u)[g]
        sts_if(t == 1, pRG + k * 8 + g, T0);                 // (A' pi+)[g]
        if (t == 1) pATP[k * 8 + g] = T0;
        sts_if(t == 2 && lo, pKF + k * 4 + 2 + g, V0);       // (B' pi+)[g]
      };
      sweep<NT, PF, +1, FragB>(N, load_B, products);
      __syncwarp();
    }
    // B2: the rest per state / row (stage-parallel); global-memory inputs fetched one round ahead
    double rg_m = 0.0, rb_m = 0.0;
    auto load_res = [&](int i) {
      const int kk = q + 4 * i;
      const int k = (kk <= N) ? kk : N;
      const bool term = (k == N);
      const CtDef& ct = term ? ctT : ctP;
      const int ku = term ? 0 : k, kn = term ? k : k + 1, j = a & 1;
      ResIn in;
      in.xa = pX[k * 8 + a]; in.pia = pPI[k * 8 + a];
      if (NL) { in.p1 = pD1[k * 16 + a]; in.l1 = pD2[k * 16 + a]; in.p2 = 0.0; in.l2 = 0.0; }   // handed over per state
      else { in.p1 = pD1[k * 8 + ct.i1]; in.p2 = pD1[k * 8 + ct.i2]; in.l1 = pD2[k * 8 + ct.i1]; in.l2 = pD2[k * 8 + ct.i2]; }
      in.u0 = pU[ku * 2]; in.u1 = pU[ku * 2 + 1];
      in.c = exact ? pC[(term ? 0 : k) * 8 + a] : 0.0; in.xn = exact ? pX[kn * 8 + a] : 0.0; in.xj = pX[k * 8 + 6 + j];
      if (NL) { in.q0 = pD2[k * 16 + 8 + j]; in.s0 = pD1[k * 16 + 8 + j]; in.q1 = 0.0; in.s1 = 0.0; }
      else { in.q0 = pD2[k * 8 + j]; in.q1 = pD2[k * 8 + 2 + j]; in.s0 = pD1[k * 8 + j]; in.s1 = pD1[k * 8 + 2 + j]; }
      return in;
    };
    ResIn rnx = load_res(0);
    PAR_LOOP for (int i = 0; i < rounds; i++) {
      const ResIn in = rnx;
      if (i + 1 < rounds) rnx = load_res(i + 1);
      const int k = q + 4 * i;
      if (k <= N) {
        const bool term = (k == N);
        const CtDef& ct = term ? ctT : ctP;
        const double xa_ = in.xa, pia = in.pia;
        const double psi_s = NL ? in.p1 : ct.c1 * in.p1 + ct.c2 * in.p2, dl_s = NL ? in.l1 : ct.c1 * in.l1 + ct.c2 * in.l2;
        double base_x;
        if (!term) {
          const double u0 = in.u0, u1 = in.u1;
          const double rb = exact ? pRB[k * 8 + a] + in.c - in.xn : om_alpha * pRB[k * 8 + a];
          rb_m = nmax(rb_m, fabs(rb));
          pRB[k * 8 + a] = rb;
          const int j = a & 1;
          const double atp = pRG[k * 8 + a], btp = pKF[k * 4 + 2 + j];
          const double gx = (a == 0) ? gs_scale * (s0 + cf.s_target * k / N) : 0.0;
          base_x = hx * xa_ + hxu0 * u0 + hxu1 * u1 + gx + atp - pia;
          const double base_u = (j == 0 ? hu0 * u0 - c10 * in.xj : hu1 * u1 - c11 * in.xj) + btp;
          const double rg_u = base_u - (in.q0 + in.q1);
          rg_m = nmax(rg_m, fabs(rg_u));
          if (a < 2) pGU[k * 2 + a] = base_u + in.s0 + in.s1;
        } else {
          base_x = hxN * xa_ + ((a == 0) ? gsN : 0.0) - pia;
        }
        const double rg_x = base_x - dl_s;
        if (k > 0) rg_m = nmax(rg_m, fabs(rg_x));
        pRG[k * 8 + a] = rg_x;
        pGX[k * 8 + a] = base_x + psi_s;
      }
    }
    res_g = wmax(NL ? nmax(rg_m, rs_m) : rg_m); res_b = wmax(rb_m); res_d = wmax(rd_m); res_m = wmax(rm_m);
    if (w.dbg && b == 0 && lane == 0 && it < 64) {
      double* d = w.dbg + it * 8;
      d[0] = res_g; d[1] = res_b; d[2] = res_d; d[3] = res_m; d[4] = mu;
    }
    if (!(res_g == res_g) || !(res_b == res_b) || !(mu == mu)) { qp_status = 3; break; }
    {
      const bool feas = res_b <= cf.tol * (1.0 + mag_b) && res_d <= cf.tol && res_m <= cf.tol_comp;
      if (feas && res_g <= cf.tol * (1.0 + mag_g)) { qp_status = 0; break; }
      // converged as far as the arithmetic allows: feasibility and complementarity at the tight tolerances, the
      // barrier parameter at its floor, stationarity within 1e-6 relative and no longer improving
      if (feas && res_g <= 1e-6 * (1.0 + mag_g) && mu <= 2.0 * mu_floor && res_g >= 0.1 * prev_res_g) { qp_status = 0; break; }
      prev_res_g = res_g;
      // numerical breakdown: four iterations with the barrier parameter at its floor and still no acceptable point
      if (mu <= 2.0 * mu_floor && ++at_floor >= 4) { qp_status = 1; break; }
      if (it >= cf.qp_max_iter) { qp_status = 1; break; }
      if (cf.infeas_exit) {
        // primal infeasible QP: eight iterations in a row with a step below 0.5 and no decrease of the barrier
        // parameter over them end the solve (status 4 after a few iterations instead of qp_max_iter)
        const double mu_old = __shfl_sync(FULL, mu_hist, it & 7);
        if (it >= 8 && short_steps >= 8 && mu >= mu_old) { qp_status = 4; break; }
        if (lane == (it & 7)) mu_hist = mu;
      }
    }
    __syncwarp();
    TICK(1)

    // ================= C: backward Riccati sweep (factorisation + predictor right-hand side)
    // FP64 tensor-core form.  Lane l = (g = l >> 2, t = l & 3) owns the m8n8k4 fragments; z is
    // ordered (x, u) here, so the extended stage matrix is E = [A | B | rb | 0] (8 x 16) and
    //   W' = E' P           (step 1, P symmetric: its accumulator tile is its own operand)
    //   M  = H~ + E' W      (step 2: tile 00 = M_xx, tile 01 = [M_xu | A' v], tile 11 = [M_uu | B' v])
    // The contraction index is permuted (slot t <-> rows 2t of the first, 2t+1 of the second
    // DMMA), which makes every accumulator pair (c0, c1) directly the operand pair of the next
    // product: no transposition between the two steps, P never leaves registers.
    {
      const int g = lane >> 2, t = lane & 3;
      const bool dgA = (g == 2 * t), dgB = (g == 2 * t + 1);       // this lane's accumulator sits on the diagonal
      const double pdg = pGAM[N * 8 + g];
      double P0 = dgA ? pdg : 0.0, P1 = dgB ? pdg : 0.0;             // P[g][2t], P[g][2t+1]
      double pp0 = pGX[N * 8 + 2 * t], pp1 = pGX[N * 8 + 2 * t + 1]; // p[2t], p[2t+1]
      const int ra = (2 * t) * 10;
      // per-lane sources of the accumulator initialisers (offsets in doubles from smem, stage stride added below)
      const bool e1on = (g <= 2);
      const int o01 = (t == 1) ? (int)(pGX - smem) + g : (int)(pMI - smem) + 2 + (g & 1);
      const int s01 = (t == 1) ? 8 : 4;
      const bool k01a = (t == 1) || (t == 0 && g == 6), k01b = (t == 0 && g == 7);
      const int o11 = (t == 1) ? (int)(pGU - smem) + (g & 1) : (int)(pMI - smem) + (g & 1);
      const int s11 = (t == 1) ? 2 : 4;
      const bool k11a = (t == 1 && g < 2) || (lane == 0), k11b = (lane == 4);
      // NL: C' gam C of the dense rows h0, h1, h4 as one more product into the xx tile: slot t = row (h0, h1, h4, -),
      // this lane holds the coefficient of state g in that row
      const bool hval = (t < 2) ? (g < 3) : (t == 2 && g >= 3 && g != 5);
      const int hoff = (t < 2) ? 4 * t + g : (g == 3) ? 8 : (g == 4) ? 9 : (g == 6) ? 10 : 11;
      if constexpr (NL) {   // terminal stage: h0, h1
        const double cH = hval ? pHC[N * 16 + hoff] : 0.0;
        dmma(P0, P1, pGH[N * 4 + t] * cH, cH);
      }
      auto load_E = [&](int k) {
        const double* p = gBA + k * 80 + ra;
        return Frag4{p[2 + g], p[12 + g], p[g & 1], p[10 + (g & 1)]};   // A[2t][g], A[2t+1][g], B[2t][g&1], B[2t+1][g&1]
      };
      auto riccati_stage = [&](const int k, const Frag4 f) {
        // operand fragments of E and accumulator initialisers (independent of P: ahead of the chain)
        const double E0a = f.a, E0b = f.b;
        double E1a = f.c, E1b = f.d;
        if (g == 2) { E1a = pRB[k * 8 + 2 * t]; E1b = pRB[k * 8 + 2 * t + 1]; }
        if (!e1on) { E1a = 0.0; E1b = 0.0; }
        const double dgv = pGAM[k * 8 + g], v01 = smem[o01 + k * s01], v11 = smem[o11 + k * s11];
        double T00a = dgA ? dgv : 0.0, T00b = dgB ? dgv : 0.0;
        double T01a = k01a ? v01 : 0.0, T01b = k01b ? v01 : 0.0;
        double T11a = k11a ? v11 : 0.0, T11b = k11b ? v11 : 0.0;
        if constexpr (NL) {
          const double cH = hval ? pHC[k * 16 + hoff] : 0.0;
          dmma(T00a, T00b, pGH[k * 4 + t] * cH, cH);
        }
        // ---- step 1: W' tiles (rows = columns of E, columns = rows of P)
        double W0a = 0.0, W0b = 0.0;
        double W1a = (g == 2) ? pp0 : 0.0, W1b = (g == 2) ? pp1 : 0.0;     // column 10: v = P rb + p
        dmma(W1a, W1b, E1a, P0); dmma(W0a, W0b, E0a, P0);
        dmma(W1a, W1b, E1b, P1); dmma(W0a, W0b, E0b, P1);
        // ---- step 2: M tiles, the uu / gradient tile first (the reciprocal chain hangs on it)
        dmma(T11a, T11b, E1a, W1a); dmma(T01a, T01b, E0a, W1a);
        dmma(T11a, T11b, E1b, W1b); dmma(T01a, T01b, E0b, W1b);
        dmma(T00a, T00b, E0a, W0a); dmma(T00a, T00b, E0b, W0b);
        // ---- eliminate u
        const double M00 = __shfl_sync(FULL, T11a, 0), M01 = __shfl_sync(FULL, T11b, 0), M11 = __shfl_sync(FULL, T11b, 4);
        const double mvu0 = __shfl_sync(FULL, T11a, 1), mvu1 = __shfl_sync(FULL, T11a, 5);
        const double xg0 = __shfl_sync(FULL, T01a, g << 2), xg1 = __shfl_sync(FULL, T01b, g << 2);        // M_xu row g
        const double xa0 = __shfl_sync(FULL, T01a, t << 3), xa1 = __shfl_sync(FULL, T01b, t << 3);        // row 2t
        const double xb0 = __shfl_sync(FULL, T01a, (t << 3) + 4), xb1 = __shfl_sync(FULL, T01b, (t << 3) + 4);  // row 2t+1
        const double mva = __shfl_sync(FULL, T01a, (t << 3) + 1), mvb = __shfl_sync(FULL, T01a, (t << 3) + 5);  // mv_x[2t], [2t+1]
        const double idet = 1.0 / (M00 * M11 - M01 * M01);
        const double i00 = M11 * idet, i01 = -M01 * idet, i11 = M00 * idet;
        const double Ka0 = -(i00 * xa0 + i01 * xa1), Ka1 = -(i01 * xa0 + i11 * xa1);   // K[:, 2t]
        const double Kb0 = -(i00 * xb0 + i01 * xb1), Kb1 = -(i01 * xb0 + i11 * xb1);   // K[:, 2t+1]
        P0 = fma(xg0, Ka0, fma(xg1, Ka1, T00a));
        P1 = fma(xg0, Kb0, fma(xg1, Kb1, T00b));
#ifndef QP_NO_SYMMETRISE   // rounding asymmetry of P drifts over the 40 stages once lam / t ~ 1e12 sits next to an O(1) Hessian: without this the
                          // interior point needs up to 50 % more iterations on thin interiors and fails where the oracle (which symmetrises) does not
        {   // P <- (P + P') / 2: the transposed entries P[2t][g], P[2t+1][g] sit in lanes (2t, g >> 1), (2t + 1, g >> 1)
          const int sa = ((2 * t) << 2) | (g >> 1), sb = ((2 * t + 1) << 2) | (g >> 1);
          const double a0 = __shfl_sync(FULL, P0, sa), a1 = __shfl_sync(FULL, P1, sa);
          const double b0 = __shfl_sync(FULL, P0, sb), b1 = __shfl_sync(FULL, P1, sb);
          P0 = 0.5 * (P0 + ((g & 1) ? a1 : a0));
          P1 = 0.5 * (P1 + ((g & 1) ? b1 : b0));
        }
#endif
        pp0 = fma(Ka0, mvu0, fma(Ka1, mvu1, mva));
        pp1 = fma(Kb0, mvu0, fma(Kb1, mvu1, mvb));
        sts2_if(g == 0, pK + k * 16 + 4 * t, Ka0, Ka1);
        sts2_if(g == 0, pK + k * 16 + 4 * t + 2, Kb0, Kb1);
        sts2_if(lane == 0, pKF + k * 4, -(i00 * mvu0 + i01 * mvu1), -(i01 * mvu0 + i11 * mvu1));
        sts2_if(lane == 0, pMI + k * 4, i00, i01);
        sts_if(lane == 0, pMI + k * 4 + 2, i11);
      };
      sweep<NT, PF, -1, Frag4>(N, load_E, riccati_stage);
      __syncwarp();
    }

    TICK(2)
    // ================= D..H: predictor (pred = 1), corrector (pred = 0)
    double alpha = 1.0, alpha_d = 1.0, mu_t = 0.0;
    for (int pred = 1; pred >= 0; pred--) {
      // ---- forward substitution: du = K dx + kf, dx+ = A dx + B du + rb
      // Matrix-vector products as DMMAs whose right operand repeats the vector in all 8 columns:
      // lane (g, t) holds (dx[2t], dx[2t+1]) going in and receives row g of the result.
      {
        const int g = lane >> 2, t = lane & 3;
        double dxa = 0.0, dxb = 0.0;  // x_0 is fixed
        const int kfo = pred ? 0 : 2;
        const bool lo = (g < 2), tlo = (t < 2);
        auto load_F = [&](int k) {
          const double* p = gBA + k * 80 + g * 10;
          const double2 af = *reinterpret_cast<const double2*>(p + 2 + 2 * t);
          return Frag3{af.x, af.y, p[t & 1]};                      // A[g][2t], A[g][2t+1], B[g][t&1]
        };
        auto forward_stage = [&](const int k, const Frag3 f) {
          sts2_if(g == 0, pDX + k * 8 + 2 * t, dxa, dxb);
          double Ka = pK[k * 16 + 4 * t + (g & 1)], Kb = pK[k * 16 + 4 * t + 2 + (g & 1)];  // K[g][2t], K[g][2t+1]
          double Bf = f.c;                                                                 // B[g][t]
          double kf = pKF[k * 4 + kfo + (g & 1)];
          const double rbg = pRB[k * 8 + g];
          if (!lo) { Ka = 0.0; Kb = 0.0; kf = 0.0; }
          if (!tlo) Bf = 0.0;
          double U0 = kf, U1 = kf, X0 = rbg, X1 = rbg;
          dmma(U0, U1, Ka, dxa); dmma(X0, X1, f.a, dxa);
          dmma(U0, U1, Kb, dxb); dmma(X0, X1, f.b, dxb);
          double duf = __shfl_sync(FULL, U0, (t & 1) << 2);      // du[t] for the k-slots 0, 1
          if (!tlo) duf = 0.0;
          dmma(X0, X1, Bf, duf);
          sts_if((lane & ~4) == 0, pDU + k * 2 + (lane >> 2), U0);
          dxa = __shfl_sync(FULL, X0, t << 3); dxb = __shfl_sync(FULL, X0, (t << 3) + 4);
        };
        sweep<NT, PF, +1, Frag3>(N, load_F, forward_stage);
        sts2_if(g == 0, pDX + N * 8 + 2 * t, dxa, dxb);
        if (lane < 2) pDU[N * 2 + lane] = 0.0;
      }
      __syncwarp();
      TICK(3)
      // ---- inequality rows at the step (stage-parallel): slack / multiplier steps, step length
      double ratio = 0.0, ratio_d = 0.0, S1 = 0.0, S1d = 0.0, S2 = 0.0;   // ratio: primal (slacks), ratio_d: dual (multipliers)
      // the global-memory inputs of a round (L2) are fetched one round ahead of their use
      auto load_rows = [&](int i) {
        const int kk = q + 4 * i;
        const int k = (kk <= N) ? kk : N;
        const bool term = (k == N);
        const RowDef& R = term ? Rt : Rp;
        const int e = k * 8 + a;
        RowIn in;
        in.lamL = pLL[e]; in.lamU = pLU[e]; in.tL = pTL[e]; in.tU = pTU[e];
        in.d1 = pred ? 0.0 : pD1[e]; in.d2 = pred ? 0.0 : pD2[e];
        in.uv = pU[(term ? 0 : k) * 2 + R.iu]; in.xv = pX[k * 8 + R.ix];
        in.xa = 0.0; in.pia = 0.0; in.atp = 0.0; in.b0 = 0.0; in.b1 = 0.0;
        return in;
      };
      if constexpr (NL) {
        PAR_LOOP for (int i = 0; i < rounds; i++) {
          const int kk = q + 4 * i;
          const bool on = (kk <= N);
          const int k = on ? kk : N;
          const bool term = (k == N);
          const RowDef& R = term ? Rt : Rp;
          const double Z = term ? Zt : Zp, zz = term ? zt : zp;
          const bool act0 = on && (term ? (a < 4) : (a < 4 || k > 0)), act1 = on && (a < 2 || (a == 4 && !term));
          const bool soft0 = term ? (a == 2 || a == 3) : (a >= 6);
          const int e0 = k * 16 + a, e1 = e0 + 8, es = k * 8 + a;
          const double xa_ = pX[k * 8 + a], u0v = pU[(term ? 0 : k) * 2], u1v = pU[(term ? 0 : k) * 2 + 1];
          const double v0 = row_value(R, term, R.iu ? u1v : u0v, __shfl_sync(FULL, xa_, R.ix, GROUP));
          const HRow hr = hrow_coef(k);
          const double v1 = hrow_val(hr, xa_, u0v, u1v);
          const double cdz0 = R.cu * pDU[k * 2 + R.iu] + R.cx * pDX[k * 8 + R.ix];
          const double cdz1 = hrow_cdz(hr, k);
          double ddl0 = 0.0, ddl1 = 0.0, A0 = 0.0, B0 = 0.0, A1 = 0.0, B1 = 0.0;
          // the lane's (possibly) soft row S -- its h row if it has one, else its static row -- and, for lanes with an
          // h row, the static row H (hard).  The slack algebra runs once per lane.
          {
            const bool sS = act1, actS = act1 || act0, softS = act1 || soft0;
            const int e = sS ? e1 : e0;
            double dS = 0.0, aS = 0.0, bS = 0.0;
            if (actS) {
              const double v = sS ? v1 : v0, lb = sS ? hlb : R.lb, ub = sS ? hub : R.ub, cdz = sS ? cdz1 : cdz0;
              RowSt r{pLL[e], pLU[e], pTL[e], pTU[e], 1.0, 1.0, 0.0, 0.0};
              if (softS) { r.sL = pSL[es]; r.sU = pSU[es]; r.lsL = pLSL[es]; r.lsU = pLSU[es]; }
              const double cL_ = pred ? 0.0 : pD1[e] - mu_t, cU_ = pred ? 0.0 : pD2[e] - mu_t;
              const double cSL = (softS && !pred) ? pDS1[es] - mu_t : 0.0, cSU = (softS && !pred) ? pDS2[es] - mu_t : 0.0;
              const RowStep st = soft_row_step(softS, v, lb, ub, cdz, Z, zz, r, cL_, cU_, cSL, cSU);
              // step-length ratios (the reciprocals of t and s are the ones soft_row_step forms)
              ratio_d = fmax(ratio_d, fmax(-st.dlL / r.lamL, -st.dlU / r.lamU));
              ratio = fmax(ratio, fmax(-st.dtL * (1.0 / r.tL), -st.dtU * (1.0 / r.tU)));
              if (softS) {
                ratio_d = fmax(ratio_d, fmax(-st.dlsL / r.lsL, -st.dlsU / r.lsU));
                ratio = fmax(ratio, fmax(-st.dsL * (1.0 / r.sL), -st.dsU * (1.0 / r.sU)));
              }
              if (pred) {
                S1 += r.lamL * st.dtL + r.lamU * st.dtU; S1d += r.tL * st.dlL + r.tU * st.dlU;
                const double dL = st.dlL * st.dtL, dU = st.dlU * st.dtU;
                double dSL = 0.0, dSU = 0.0;
                S2 += dL + dU;
                if (softS) {
                  dSL = st.dlsL * st.dsL; dSU = st.dlsU * st.dsU;
                  S1 += r.lsL * st.dsL + r.lsU * st.dsU; S1d += r.sL * st.dlsL + r.sU * st.dlsU; S2 += dSL + dSU;
                  pDS1[es] = dSL; pDS2[es] = dSU;
                }
                pD1[e] = dL; pD2[e] = dU;
                soft_row_chi(softS, Z, r, dL, dU, dSL, dSU, aS, bS);
              } else {
                pD1[e] = st.dlL; pD2[e] = st.dlU;
                dS = st.dlL - st.dlU;
                if (softS) { pDS1[es] = st.dlsL; pDS2[es] = st.dlsU; pE1[es] = st.dsL; pE2[es] = st.dsU; }
              }
            }
            double dH = 0.0, aH = 0.0, bH = 0.0;
            if (act1 && act0) {          // static row of a lane that also holds an h row: hard
              const RowSt r{pLL[e0], pLU[e0], pTL[e0], pTU[e0], 1.0, 1.0, 0.0, 0.0};
              const RowStep st = soft_row_step(false, v0, R.lb, R.ub, cdz0, Z, zz, r, pred ? 0.0 : pD1[e0] - mu_t,
                                               pred ? 0.0 : pD2[e0] - mu_t, 0.0, 0.0);
              ratio_d = fmax(ratio_d, fmax(-st.dlL / r.lamL, -st.dlU / r.lamU));
              ratio = fmax(ratio, fmax(-st.dtL * (1.0 / r.tL), -st.dtU * (1.0 / r.tU)));
              if (pred) {
                S1 += r.lamL * st.dtL + r.lamU * st.dtU; S1d += r.tL * st.dlL + r.tU * st.dlU;
                const double dL = st.dlL * st.dtL, dU = st.dlU * st.dtU;
                S2 += dL + dU;
                pD1[e0] = dL; pD2[e0] = dU;
                soft_row_chi(false, Z, r, dL, dU, 0.0, 0.0, aH, bH);
              } else {
                pD1[e0] = st.dlL; pD2[e0] = st.dlU;
                dH = st.dlL - st.dlU;
              }
            }
            ddl0 = sS ? dH : dS; ddl1 = sS ? dS : 0.0;
            A0 = sS ? aH : aS; B0 = sS ? bH : bS; A1 = sS ? aS : 0.0; B1 = sS ? bS : 0.0;
          }
          const CtDef& ct = term ? ctT : ctP;
          const double* h = pHC + k * 16;
          const double hc1 = (a != 5) ? h[so1] : 0.0, hc2 = (a < 3) ? h[4 + a] : hc2r;
          const int ap2 = (a + 2) & 7;
          if (pred) {
            // corrector right-hand side chi = A - mu_t B of every row, handed over per state / control
            const double a1 = __shfl_sync(FULL, A0, ct.i1, GROUP), a2 = __shfl_sync(FULL, A0, ct.i2, GROUP), a3 = __shfl_sync(FULL, A0, ap2, GROUP);
            const double b1 = __shfl_sync(FULL, B0, ct.i1, GROUP), b2 = __shfl_sync(FULL, B0, ct.i2, GROUP), b3 = __shfl_sync(FULL, B0, ap2, GROUP);
            const double ha1 = __shfl_sync(FULL, A1, hr1, GROUP), ha2 = __shfl_sync(FULL, A1, hr2, GROUP), ha3 = __shfl_sync(FULL, A1, ap2, GROUP);
            const double hb1 = __shfl_sync(FULL, B1, hr1, GROUP), hb2 = __shfl_sync(FULL, B1, hr2, GROUP), hb3 = __shfl_sync(FULL, B1, ap2, GROUP);
            if (on) {
              pCA[k * 16 + a] = ct.c1 * a1 + ct.c2 * a2 + hc1 * ha1 + hc2 * ha2;
              pCB[k * 16 + a] = ct.c1 * b1 + ct.c2 * b2 + hc1 * hb1 + hc2 * hb2;
              if (a < 2 && !term) {
                const double rc = (a == 0) ? itT : itD;
                pCA[k * 16 + 8 + a] = A0 + a3 + rc * ha3;
                pCB[k * 16 + 8 + a] = B0 + b3 + rc * hb3;
              }
            }
          } else {
            const double d1 = __shfl_sync(FULL, ddl0, ct.i1, GROUP), d2 = __shfl_sync(FULL, ddl0, ct.i2, GROUP);
            const double hd1 = __shfl_sync(FULL, ddl1, hr1, GROUP), hd2 = __shfl_sync(FULL, ddl1, hr2, GROUP);
            if (on) {
              const double hdz = term ? hxN * pDX[k * 8 + a] : hx * pDX[k * 8 + a] + hxu0 * pDU[k * 2] + hxu1 * pDU[k * 2 + 1];
              pGAM[k * 8 + a] = pRG[k * 8 + a] - (ct.c1 * d1 + ct.c2 * d2 + hc1 * hd1 + hc2 * hd2) + hdz;
            }
          }
        }
      } else {
      RowIn nxt = load_rows(0);
      PAR_LOOP for (int i = 0; i < rounds; i++) {
        const RowIn cur = nxt;
        if (i + 1 < rounds) nxt = load_rows(i + 1);
        const int kk = q + 4 * i;
        const bool on = (kk <= N);
        const int k = on ? kk : N;
        const bool term = (k == N);
        const RowDef& R = term ? Rt : Rp;
        const bool act = on && (term ? (a < 4) : (a < 4 || k > 0));
        double ddl = 0.0;   // dlamL - dlamU of this row (corrector pass)
        if (act) {
          const double cdz = R.cu * pDU[k * 2 + R.iu] + R.cx * pDX[k * 8 + R.ix];
          const double v = row_value(R, term, cur.uv, cur.xv);
          const double lamL = cur.lamL, lamU = cur.lamU, tL = cur.tL, tU = cur.tU;
          const double itL = 1.0 / tL, itU = 1.0 / tU;
          const double dtL = cdz + (v - R.lb - tL), dtU = -cdz + (R.ub - v - tU);
          if (pred) {
            const double dlL = -lamL - lamL * dtL * itL, dlU = -lamU - lamU * dtU * itU;
            S1 += lamL * dtL + lamU * dtU; S1d += tL * dlL + tU * dlU;
            S2 += dlL * dtL + dlU * dtU;
            pD1[k * 8 + a] = dlL * dtL; pD2[k * 8 + a] = dlU * dtU;
            ratio_d = fmax(ratio_d, fmax(1.0 + dtL * itL, 1.0 + dtU * itU));
            ratio = fmax(ratio, fmax(-dtL * itL, -dtU * itU));
          } else {
            const double dlL = -lamL - (cur.d1 - mu_t + lamL * dtL) * itL;
            const double dlU = -lamU - (cur.d2 - mu_t + lamU * dtU) * itU;
            pD1[k * 8 + a] = dlL; pD2[k * 8 + a] = dlU;
            ddl = dlL - dlU;
            ratio_d = fmax(ratio_d, fmax(-dlL * (1.0 / lamL), -dlU * (1.0 / lamU)));
            ratio = fmax(ratio, fmax(-dtL * itL, -dtU * itU));
          }
        }
        if (!pred) {
          // right-hand side of the costate recursion per state: rg - C'(dlamL - dlamU) + (H dz)_x,
          // left in the gamma slot (the rows of a stage sit in the 8 lanes of this segment)
          const CtDef& ct = term ? ctT : ctP;
          const double d1 = __shfl_sync(FULL, ddl, ct.i1, GROUP), d2 = __shfl_sync(FULL, ddl, ct.i2, GROUP);
          if (on) {
            const double hdz = term ? hxN * pDX[k * 8 + a] : hx * pDX[k * 8 + a] + hxu0 * pDU[k * 2] + hxu1 * pDU[k * 2 + 1];
            pGAM[k * 8 + a] = pRG[k * 8 + a] - (ct.c1 * d1 + ct.c2 * d2) + hdz;
          }
        }
      }
      }
      ratio = wmaxf(ratio); ratio_d = wmaxf(ratio_d);
      if (!cf.split_step) ratio = ratio_d = fmax(ratio, ratio_d);
      if (!pred) {
        // fraction to the boundary: 0.995 far from the solution, -> 1 with the barrier parameter (Mehrotra)
        const double tau_k = fmax(TAU, 1.0 - mu_t);
        alpha = tau_k * ((ratio > 1.0) ? 1.0 / ratio : 1.0);
        alpha_d = tau_k * ((ratio_d > 1.0) ? 1.0 / ratio_d : 1.0);
        break;
      }
      const double al_aff = (ratio > 1.0) ? 1.0 / ratio : 1.0, ad_aff = (ratio_d > 1.0) ? 1.0 / ratio_d : 1.0;
      S1 = wsum(S1); S1d = wsum(S1d); S2 = wsum(S2);
      const double mu_aff = (S0 + al_aff * S1 + ad_aff * S1d + al_aff * ad_aff * S2) * inv_n;
      const double sr = mu_aff / mu;
      // centring target, floored so the iterates stay on the central path at ~0.1 tol_comp
      mu_t = fmax(sr * sr * sr * mu, mu_floor);
      // ---- corrector term chi of every row (stage-parallel); what the backward sweep adds per state
      // (C' chi) goes to the gamma slot, the two control entries to the kfc slot
      auto load_chi = [&](int i) {
        const int kk = q + 4 * i;
        const int e = ((kk <= N) ? kk : N) * 8 + a;
        return Frag4{pD1[e], pD2[e], pTL[e], pTU[e]};
      };
      if constexpr (NL) {
        PAR_LOOP for (int i = 0; i < rounds; i++) {
          const int k = q + 4 * i;
          if (k <= N) {
            pGAM[k * 8 + a] = pCA[k * 16 + a] - mu_t * pCB[k * 16 + a];
            if (a < 2 && k < N) pKF[k * 4 + 2 + a] = pCA[k * 16 + 8 + a] - mu_t * pCB[k * 16 + 8 + a];
          }
        }
      } else {
      Frag4 cnx = load_chi(0);
      PAR_LOOP for (int i = 0; i < rounds; i++) {
        const Frag4 cc = cnx;
        if (i + 1 < rounds) cnx = load_chi(i + 1);
        const int kk = q + 4 * i;
        const bool on = (kk <= N);
        const int k = on ? kk : N;
        const bool term = (k == N);
        const bool act = on && (term ? (a < 4) : (a < 4 || k > 0));
        double chi = 0.0;
        if (act) chi = (cc.a - mu_t) * (1.0 / cc.c) - (cc.b - mu_t) * (1.0 / cc.d);   // reciprocals: half the instructions of a division
        const CtDef& ct = term ? ctT : ctP;
        const double c1v = __shfl_sync(FULL, chi, ct.i1, GROUP), c2v = __shfl_sync(FULL, chi, ct.i2, GROUP);
        const double cp2 = __shfl_sync(FULL, chi, (a + 2) & 7, GROUP);
        if (on) {
          pGAM[k * 8 + a] = ct.c1 * c1v + ct.c2 * c2v;
          if (a < 2 && !term) pKF[k * 4 + 2 + a] = chi + cp2;
        }
      }
      }
      __syncwarp();
      TICK(4)
      // ---- backward substitution of the corrector right-hand side C' chi through the same factor
      //   dmu = chi_u + B' dp+,  kfc = kff - M_uu^-1 dmu,  dp = (C' chi)_x + A' dp+ + K' dmu      (DMMA form)
      {
        const int g = lane >> 2, t = lane & 3;
        const bool lo = (g < 2), tlo = (t < 2);
        const int ra = (2 * t) * 10;
        double dpa = pGAM[N * 8 + 2 * t], dpb = pGAM[N * 8 + 2 * t + 1];
        auto load_E = [&](int k) {
          const double* p = gBA + k * 80 + ra;
          return Frag4{p[2 + g], p[12 + g], p[g & 1], p[10 + (g & 1)]};   // A[2t][g], A[2t+1][g], B[2t][g&1], B[2t+1][g&1]
        };
        auto backward_stage = [&](const int k, const Frag4 f) {
          const double E0a = f.a, E0b = f.b;
          double E1a = f.c, E1b = f.d;
          double Ui = pKF[k * 4 + 2 + (g & 1)];
          double Kf = pK[k * 16 + 2 * g + (t & 1)];                          // K[t][g]
          const double Xi = pGAM[k * 8 + g];
          const double kf0 = pKF[k * 4], kf1 = pKF[k * 4 + 1];
          const double i00 = pMI[k * 4], i01 = pMI[k * 4 + 1], i11 = pMI[k * 4 + 2];
          if (!lo) { E1a = 0.0; E1b = 0.0; Ui = 0.0; }
          if (!tlo) Kf = 0.0;
          double U0 = Ui, U1 = Ui, X0 = Xi, X1 = Xi;
          dmma(U0, U1, E1a, dpa); dmma(X0, X1, E0a, dpa);
          dmma(U0, U1, E1b, dpb); dmma(X0, X1, E0b, dpb);
          const double d0 = __shfl_sync(FULL, U0, 0), d1 = __shfl_sync(FULL, U0, 4);
          const double dmuf = tlo ? ((t == 0) ? d0 : d1) : 0.0;
          dmma(X0, X1, Kf, dmuf);
          sts2_if(lane == 0, pKF + k * 4 + 2, kf0 - (i00 * d0 + i01 * d1), kf1 - (i01 * d0 + i11 * d1));
          dpa = __shfl_sync(FULL, X0, t << 3); dpb = __shfl_sync(FULL, X0, (t << 3) + 4);
        };
        sweep<NT, PF, -1, Frag4>(N, load_E, backward_stage);
      }
      __syncwarp();
      TICK(5)
    }
    TICK(4)
    if (w.dbg && b == 0 && lane == 0 && it < 64) { w.dbg[it * 8 + 5] = alpha; w.dbg[it * 8 + 6] = mu_t; }
    if (fmin(alpha, alpha_d) < 1e-12) { qp_status = 2; break; }
    short_steps = (fmin(alpha, alpha_d) < 0.5) ? short_steps + 1 : 0;
    __syncwarp();

    // ================= I: costate recursion of the Newton step -> d pi (kept in the rg slot)
    //   dpi_k = [rg - C' dlam + H dz]_k + A_k' dpi_{k+1}                                   (DMMA form)
    {
      const int g = lane >> 2, t = lane & 3;
      const int ra = (2 * t) * 10;
      double da = pGAM[N * 8 + 2 * t], db = pGAM[N * 8 + 2 * t + 1];
      sts2_if(g == 0, pRG + N * 8 + 2 * t, da, db);
      auto load_A = [&](int k) {
        const double* p = gBA + k * 80 + ra;
        return Frag2{p[2 + g], p[12 + g]};                                  // A[2t][g], A[2t+1][g]
      };
      auto costate_stage = [&](const int k, const Frag2 f) {
        const double E0a = f.a, E0b = f.b;
        const double Xi = pGAM[k * 8 + g];
        double X0 = 0.0, X1 = 0.0;
        dmma(X0, X1, E0a, da); dmma(X0, X1, E0b, db);
        sts_if(t == 0, pK + k * 16 + g, X0);       // A' dpi+ (the gains are dead by now): carried into A' pi+ in phase J
        X0 += Xi;
        sts_if(t == 0, pRG + k * 8 + g, X0);       // applied to pi in the update phase
        da = __shfl_sync(FULL, X0, t << 3); db = __shfl_sync(FULL, X0, (t << 3) + 4);
      };
      sweep<NT, PF, -1, Frag2>(N, load_A, costate_stage);
      __syncwarp();
    }
    TICK(6)
    // ================= J + A: step of multipliers / slacks, and the row pass of the NEXT iteration on the
    // updated values while they are in registers, and the step of the primal point / costate (stage-parallel)
    rd_m = 0.0; rm_m = 0.0; S0 = 0.0;
    auto load_upd = [&](int i) {      // global-memory inputs of a round, fetched one round ahead
      const int kk = q + 4 * i;
      const int k = (kk <= N) ? kk : N;
      const bool term = (k == N);
      const RowDef& R = term ? Rt : Rp;
      const int e = k * 8 + a;
      RowIn in;
      in.lamL = pLL[e]; in.lamU = pLU[e]; in.tL = pTL[e]; in.tU = pTU[e]; in.d1 = pD1[e]; in.d2 = pD2[e];
      in.uv = pU[(term ? 0 : k) * 2 + R.iu]; in.xv = pX[k * 8 + R.ix];
      in.xa = pX[e]; in.pia = pPI[e];
      in.atp = pATP[(term ? 0 : k) * 8 + a];
      const double2 bb = *reinterpret_cast<const double2*>(gBA + (k > 0 ? k - 1 : 0) * 80 + a * 10);   // B_{k-1}[a][0..1]
      in.b0 = bb.x; in.b1 = bb.y;
      return in;
    };
    // carried products of phase B for the next iteration: (B_{k-1}' pi_k)[0..1] -> kfc slot of stage k - 1,
    // A_k' pi_{k+1} += alpha_d A_k' dpi_{k+1} -> global copy and the rg slot (where phase B2 picks it up)
    auto carry_products = [&](int k, bool on, bool term, double pin, double atp, double b0, double b1) {
      const double bt0 = gsum(b0 * pin), bt1 = gsum(b1 * pin);
      sts2_if(a == 0 && on && k >= 1, pKF + (k - 1) * 4 + 2, bt0, bt1);
      if (on && !term) {
        const double an = fma(alpha_d, pK[k * 16 + a], atp);
        pATP[k * 8 + a] = an; pRG[k * 8 + a] = an;
      }
    };
    if constexpr (NL) {
      rs_m = 0.0;
      PAR_LOOP for (int i = 0; i < rounds; i++) {
        const int kk = q + 4 * i;
        const bool on = (kk <= N);
        const int k = on ? kk : N;
        const bool term = (k == N);
        const RowDef& R = term ? Rt : Rp;
        const bool act0 = on && (term ? (a < 4) : (a < 4 || k > 0)), act1 = on && (a < 2 || (a == 4 && !term));
        const bool soft0 = term ? (a == 2 || a == 3) : (a >= 6);
        const int e0 = k * 16 + a, e1 = e0 + 8, es = k * 8 + a, ku = term ? 0 : k;
        const double xa_ = pX[k * 8 + a], pia = pPI[k * 8 + a], u0v = pU[ku * 2], u1v = pU[ku * 2 + 1];
        const HRow hr = hrow_coef(k);
        const double v0o = row_value(R, term, R.iu ? u1v : u0v, __shfl_sync(FULL, xa_, R.ix, GROUP));
        const double v1o = hrow_val(hr, xa_, u0v, u1v);
        const double cdz0 = R.cu * pDU[k * 2 + R.iu] + R.cx * pDX[k * 8 + R.ix];
        const double cdz1 = hrow_cdz(hr, k);
        // the primal point after the step, and the rows there
        const double xn_ = fma(alpha, pDX[k * 8 + a], xa_);
        const double u0n = term ? u0v : fma(alpha, pDU[k * 2], u0v), u1n = term ? u1v : fma(alpha, pDU[k * 2 + 1], u1v);
        const double v0 = row_value(R, term, R.iu ? u1n : u0n, __shfl_sync(FULL, xn_, R.ix, GROUP));
        const double v1 = hrow_val(hr, xn_, u0n, u1n);
        RowSt r0{0.0, 0.0, 1.0, 1.0, 1.0, 1.0, 0.0, 0.0}, r1 = r0;
        {
          const bool sS = act1, actS = act1 || act0, softS = act1 || soft0;
          const int e = sS ? e1 : e0;
          if (actS) {
            const double vo = sS ? v1o : v0o, lb = sS ? hlb : R.lb, ub = sS ? hub : R.ub, cdz = sS ? cdz1 : cdz0;
            RowSt r{pLL[e], pLU[e], pTL[e], pTU[e], 1.0, 1.0, 0.0, 0.0};
            double dsL = 0.0, dsU = 0.0;
            if (softS) { r.sL = pSL[es]; r.sU = pSU[es]; r.lsL = pLSL[es]; r.lsU = pLSU[es]; dsL = pE1[es]; dsU = pE2[es]; }
            const double rdL = softS ? vo + r.sL - lb - r.tL : vo - lb - r.tL;
            const double rdU = softS ? ub + r.sU - vo - r.tU : ub - vo - r.tU;
            r.tL = fma(alpha, cdz + dsL + rdL, r.tL); r.tU = fma(alpha, -cdz + dsU + rdU, r.tU);
            r.lamL = fma(alpha_d, pD1[e], r.lamL); r.lamU = fma(alpha_d, pD2[e], r.lamU);
            pLL[e] = r.lamL; pLU[e] = r.lamU; pTL[e] = r.tL; pTU[e] = r.tU;
            if (softS) {
              r.sL = fma(alpha, dsL, r.sL); r.sU = fma(alpha, dsU, r.sU);
              r.lsL = fma(alpha_d, pDS1[es], r.lsL); r.lsU = fma(alpha_d, pDS2[es], r.lsU);
              pSL[es] = r.sL; pSU[es] = r.sU; pLSL[es] = r.lsL; pLSU[es] = r.lsU;
            }
            if (sS) r1 = r; else r0 = r;
          }
          if (act1 && act0) {        // static row of a lane that also holds an h row: hard
            RowSt r{pLL[e0], pLU[e0], pTL[e0], pTU[e0], 1.0, 1.0, 0.0, 0.0};
            const double rdL = v0o - R.lb - r.tL, rdU = R.ub - v0o - r.tU;
            r.tL = fma(alpha, cdz0 + rdL, r.tL); r.tU = fma(alpha, -cdz0 + rdU, r.tU);
            r.lamL = fma(alpha_d, pD1[e0], r.lamL); r.lamU = fma(alpha_d, pD2[e0], r.lamU);
            pLL[e0] = r.lamL; pLU[e0] = r.lamU; pTL[e0] = r.tL; pTU[e0] = r.tU;
            r0 = r;
          }
        }
        rows_A_nl(k, on, term, act0, v0, r0, act1, v1, r1);
        const double pin = fma(alpha_d, pRG[k * 8 + a], pia);
        {
          const double2 bb = *reinterpret_cast<const double2*>(gBA + (k > 0 ? k - 1 : 0) * 80 + a * 10);
          carry_products(k, on, term, pin, pATP[(term ? 0 : k) * 8 + a], bb.x, bb.y);
        }
        if (on) {
          pX[k * 8 + a] = xn_;
          pPI[k * 8 + a] = pin;
          if (a < 2 && !term) pU[k * 2 + a] = a ? u1n : u0n;
        }
      }
    } else {
    RowIn unx = load_upd(0);
    PAR_LOOP for (int i = 0; i < rounds; i++) {
      const RowIn cur = unx;
      if (i + 1 < rounds) unx = load_upd(i + 1);
      const int kk = q + 4 * i;
      const bool on = (kk <= N);
      const int k = on ? kk : N;
      const bool term = (k == N);
      const RowDef& R = term ? Rt : Rp;
      const bool act = on && (term ? (a < 4) : (a < 4 || k > 0));
      double lamL = 0.0, lamU = 0.0, tL = 1.0, tU = 1.0, v = 0.0;
      if (act) {
        const double cdz = R.cu * pDU[k * 2 + R.iu] + R.cx * pDX[k * 8 + R.ix];
        v = row_value(R, term, cur.uv, cur.xv);
        tL = cur.tL; tU = cur.tU;
        const double dtL = cdz + (v - R.lb - tL), dtU = -cdz + (R.ub - v - tU);
        lamL = fma(alpha_d, cur.d1, cur.lamL);
        lamU = fma(alpha_d, cur.d2, cur.lamU);
        tL = fma(alpha, dtL, tL); tU = fma(alpha, dtU, tU);
        // the row at the updated point, from the values the primal update below will store
        v = row_value(R, term, fma(alpha, pDU[(term ? 0 : k) * 2 + R.iu], cur.uv), fma(alpha, pDX[k * 8 + R.ix], cur.xv));
        pLL[k * 8 + a] = lamL; pLU[k * 8 + a] = lamU; pTL[k * 8 + a] = tL; pTU[k * 8 + a] = tU;
      }
      rows_A(k, on, term, act, R, v, lamL, lamU, tL, tU);
      // step of the primal point and the costate: state a / control a of this stage.  The shuffles inside
      // rows_A separate these stores from the row loads of the stage above.
      const double pin = fma(alpha_d, pRG[k * 8 + a], cur.pia);
      carry_products(k, on, term, pin, cur.atp, cur.b0, cur.b1);
      if (on) {
        pX[k * 8 + a] = fma(alpha, pDX[k * 8 + a], cur.xa);
        pPI[k * 8 + a] = pin;
        if (a < 2 && !term) pU[k * 2 + a] = fma(alpha, pDU[k * 2 + a], pU[k * 2 + a]);
      }
    }
    }
    __syncwarp();
    TICK(7)
    om_alpha = 1.0 - alpha;
    it++;
  }
#ifdef IHM2B_PHASE_TIMERS
  if (w.dbg && b == 0 && lane == 0) for (int i = 0; i < 8; i++) w.dbg[63 * 8 + i] = (double)tph[i] / (it > 0 ? it : 1);
#endif

  // ------------------------------------------------------------------ outputs (RTI full step)
  __syncwarp();
  if (lane == 0) {
    int st;
    if (qp_status == 3) st = IHM2B_FAILURE;
    else if (qp_status == 0) st = IHM2B_SUCCESS;
    else {
      // accept an IPM that stalled above the tight tolerance but is converged to 1e-6
      const bool loose = res_g <= 1e-6 * (1.0 + mag_g) && res_b <= 1e-6 * (1.0 + mag_b) &&
                         res_d <= 1e-6 && res_m <= 1e-6;
      // acados carries on with the iterate of a QP solver that stopped at its iteration limit; the early exit for
      // infeasible QPs (4, not in HPIPM) stands in for that limit
      st = (loose || (cf.qp_tolerate_maxiter && (qp_status == 1 || qp_status == 4))) ? IHM2B_SUCCESS : IHM2B_QP_FAILURE;
    }
    double* sp = w.stats + (size_t)b * 8;
    w.last_it[b] = it;           // launch order of the handle's next QP
    if (w.sqp_it < 0) {          // RTI: one QP per call, its residuals are the statistics ([7], the cost: ocp_cost_kernel)
      sp[0] = it; sp[1] = 1.0; sp[2] = mu; sp[3] = res_g; sp[4] = res_b; sp[5] = res_d; sp[6] = res_m;
    } else {                     // SQP: the NLP residuals (sp[3..6]) are written by nlp_residual_kernel
      sp[0] = (w.sqp_it > 0 ? sp[0] : 0.0) + it; sp[1] = w.sqp_it + 1; sp[2] = mu;
      if (st != IHM2B_SUCCESS) w.active[b] = 0;   // the iteration stops at a failed QP (status 1 / 4)
      else st = IHM2B_MAXITER;                    // until the residual test of a later iteration says converged
    }
    w.status[b] = st;
    if (status_out) status_out[b] = st;
  }
  if (u0_out && lane < 2) u0_out[(size_t)b * 2 + lane] = pU[lane];
}

// ---------------------------------------------------------------------------------------
// SQP mode: KKT residuals of the NLP at the current iterate with the multipliers of the last
// QP (acados ocp_nlp_sqp, SURVEY App. C: res_stat, res_eq, res_ineq, res_comp in the
// infinity norm; all <= nlp_tol stops the instance with status 0).  Uses the fresh
// linearisation (phi, [B A]).  One warp per problem, lane (a, q) = state / row a of the
// stages q, q+4, ...  grid = B, block = 32.
// ---------------------------------------------------------------------------------------
template <bool NL>
__global__ void __launch_bounds__(32) nlp_residual_kernel(const KCfg cf, const Work w, const double* __restrict__ x0,
                                                          const double nlp_tol, const int have_mult,
                                                          int32_t* __restrict__ status_out) {
  const int lane = threadIdx.x, a = lane & 7, q = lane >> 3, b = blockIdx.x;
  if (w.active && !w.active[b]) return;
  const int N = cf.N, N1 = N + 1;
  constexpr int RS = NL ? 16 : 8;
  const double* gBA = w.BA + (size_t)b * N * 80;
  const double* phi = w.phi + (size_t)b * N * 8;
  const double* pX = w.x + (size_t)b * N1 * 8;
  const double* pU = w.u + (size_t)b * N * 2;
  const double* pPI = w.xa + (size_t)b * N1 * 8;
  const double* pLL = w.LT + (size_t)b * N1 * 4 * RS;
  const double* pLU = pLL + N1 * RS;
  const double* pLSL = NL ? w.SS + ((size_t)b * 8 + 2) * N1 * 8 : nullptr;   // multipliers of sL, sU >= 0, [N1][8]: slot = lane
  const double* pLSU = pLSL + N1 * 8;
  const double* gHC = NL ? w.HC + (size_t)b * N1 * 16 : nullptr;
  double itT = 0.0, itD = 0.0;
  if (NL) { itT = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TT]; itD = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TD]; }
  const HLane hl(cf, a, itT, itD);
  const double s0 = x0[(size_t)b * 8];
  const double hx = cf.cs * (cf.W[a] + (a == 6 ? cf.W[10] : a == 7 ? cf.W[11] : 0.0));
  const double hxN = cf.ct * cf.We[a];
  const double hu0 = cf.cs * (cf.W[8] + cf.W[10]), hu1 = cf.cs * (cf.W[9] + cf.W[11]);
  const double c10 = cf.cs * cf.W[10], c11 = cf.cs * cf.W[11];
  const double hxu0 = (a == 6) ? -c10 : 0.0, hxu1 = (a == 7) ? -c11 : 0.0;
  const double gs_scale = -cf.cs * cf.W[0], gsN = -cf.ct * cf.We[0] * (s0 + cf.s_target);
  const RowDef Rp = row_path(cf, a), Rt = row_term(cf, a);
  const CtDef ctP = ct_path(a), ctT = ct_term(a);
  double rs = 0.0, re = fabs(pX[a] - x0[(size_t)b * 8 + a]), ri = 0.0, rc = 0.0;
  const unsigned seg = 0xffu << (8 * q);   // the segments run different trip counts
  // one two-sided row at value v: residuals; returns lamL - lamU.  Soft rows: slack eliminated, s* = violation
  auto row_res = [&](bool soft, double v, double lb, double ub, int e, int es, double Z, double zz) {
    const double sL = soft ? fmax(0.0, lb - v) : 0.0, sU = soft ? fmax(0.0, v - ub) : 0.0;
    ri = nmax(ri, nmax(lb - sL - v, v - ub - sU));
    if (!have_mult) return 0.0;
    const double lamL = pLL[e], lamU = pLU[e];
    rc = nmax(rc, nmax(fabs(lamL * (v + sL - lb)), fabs(lamU * (ub + sU - v))));
    if (NL && soft) {
      const double lsL = pLSL[es], lsU = pLSU[es];
      rc = nmax(rc, nmax(fabs(lsL * sL), fabs(lsU * sU)));
      rs = nmax(rs, nmax(fabs(Z * sL + zz - lamL - lsL), fabs(Z * sU + zz - lamU - lsU)));
    }
    return lamL - lamU;
  };
  for (int k = q; k <= N; k += 4) {
    const bool term = (k == N);
    const double xa_ = pX[k * 8 + a];
    const double Z = (term ? cf.ct : cf.cs) * cf.slack_Z, zz = (term ? cf.ct : cf.cs) * cf.slack_z;
    // inequality row a of this stage
    const RowDef& R = term ? Rt : Rp;
    const bool act = term ? (a < 4) : (a < 4 || k > 0);
    double dl = 0.0, dl1 = 0.0;
    if (act) {
      const double v = row_value(R, term, pU[(term ? 0 : k) * 2 + R.iu], pX[k * 8 + R.ix]);
      dl = row_res(NL && (term ? (a == 2 || a == 3) : (a >= 6)), v, R.lb, R.ub, k * RS + a, k * 8 + a, Z, zz);
    }
    if (NL && (a < 2 || (a == 4 && !term))) {
      const double v1 = hl.value(gHC + k * 16, pX + k * 8, term ? 0.0 : pU[k * 2], term ? 0.0 : pU[k * 2 + 1], a);
      dl1 = row_res(true, v1, hl.lb, hl.ub, k * RS + 8 + a, k * 8 + a, Z, zz);
    }
    // C'(lamL - lamU): the rows of this stage live in the 8 lanes of this segment
    const CtDef& ct = term ? ctT : ctP;
    const double d1 = __shfl_sync(seg, dl, ct.i1, GROUP), d2 = __shfl_sync(seg, dl, ct.i2, GROUP);
    const double dj = __shfl_sync(seg, dl, a & 1, GROUP), dj2 = __shfl_sync(seg, dl, 2 + (a & 1), GROUP);
    double cdl = ct.c1 * d1 + ct.c2 * d2, cdu = dj + dj2;
    if (NL) {
      const double h1 = __shfl_sync(seg, dl1, hl.r1, GROUP), h2 = __shfl_sync(seg, dl1, hl.r2, GROUP);
      const double hj = __shfl_sync(seg, dl1, 2 + (a & 1), GROUP);
      const double* h = gHC + k * 16;
      cdl += h[hl.so1] * h1 + ((a < 3) ? h[4 + a] : hl.hc2r) * h2;   // h[15] = 0 for the state no row touches
      cdu += ((a & 1) ? itD : itT) * hj;
    }
    double rgx;
    if (!term) {
      const double u0 = pU[k * 2], u1 = pU[k * 2 + 1];
      re = nmax(re, fabs(phi[k * 8 + a] - pX[(k + 1) * 8 + a]));
      double atp = 0.0, btp = 0.0;
      const int j = a & 1;
      if (have_mult) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
          const double pn = pPI[(k + 1) * 8 + r];
          atp = fma(gBA[k * 80 + r * 10 + 2 + a], pn, atp);
          btp = fma(gBA[k * 80 + r * 10 + j], pn, btp);
        }
      }
      const double gx = (a == 0) ? gs_scale * (s0 + cf.s_target * k / N) : 0.0;
      rgx = hx * xa_ + hxu0 * u0 + hxu1 * u1 + gx + atp - (have_mult ? pPI[k * 8 + a] : 0.0) - cdl;
      const double rgu = (j == 0 ? hu0 * u0 - c10 * pX[k * 8 + 6] : hu1 * u1 - c11 * pX[k * 8 + 7]) + btp - cdu;
      if (a < 2) rs = nmax(rs, fabs(rgu));
    } else {
      rgx = hxN * xa_ + ((a == 0) ? gsN : 0.0) - (have_mult ? pPI[k * 8 + a] : 0.0) - cdl;
    }
    if (k > 0) rs = nmax(rs, fabs(rgx));
  }
  rs = wmax(rs); re = wmax(re); ri = fmax(wmax(ri), 0.0); rc = wmax(rc);
  if (!have_mult) { rs = __longlong_as_double(0x7ff0000000000000LL); rc = rs; }
  if (lane == 0) {
    double* sp = w.stats + (size_t)b * 8;
    sp[3] = rs; sp[4] = re; sp[5] = ri; sp[6] = rc;
    if (have_mult && rs <= nlp_tol && re <= nlp_tol && ri <= nlp_tol && rc <= nlp_tol) {
      w.status[b] = IHM2B_SUCCESS;
      if (status_out) status_out[b] = IHM2B_SUCCESS;
      w.active[b] = 0;
    }
  }
}

void launch_nlp_residual(const KCfg& cf, const Work& w, const double* x0, double nlp_tol, int have_mult,
                         int32_t* status, cudaStream_t stream) {
  if (cf.nl) nlp_residual_kernel<true><<<cf.B, 32, 0, stream>>>(cf, w, x0, nlp_tol, have_mult, status);
  else nlp_residual_kernel<false><<<cf.B, 32, 0, stream>>>(cf, w, x0, nlp_tol, have_mult, status);
}

// ---------------------------------------------------------------------------------------
// SQP mode, globalization = MERIT_BACKTRACKING (new_python/controller.py:308; acados line search restated in
// SURVEY App. C): L1 merit = cost + sum w_eq |equality defect| + sum w_lo max(0, lb - v) + w_up max(0, v - ub),
// weights from the multipliers of the QP just solved (|mult| at the first SQP iteration, then
// max(|mult|, (w + |mult|) / 2)); alpha <- 0.7 alpha until the merit decreases, not below 0.05.
// One warp per problem, lane (a, q) = state / row a of the stages q, q+4, ...
// ---------------------------------------------------------------------------------------
constexpr double LS_ALPHA_MIN = 0.05, LS_REDUCTION = 0.7;

// this lane's share of the OCP cost at (x, u): z'(H z / 2 + g) per stage, plus (NL) the penalties of the soft rows
// with their slacks eliminated (the h rows through their current linearisation Work::HC: exact at its point)
template <bool NL>
__device__ __forceinline__ double cost_lane(const KCfg& cf, const int N, const int a, const int q, const double s0,
                                            const double* __restrict__ pX, const double* __restrict__ pU,
                                            const double* __restrict__ gHC, const HLane& hl) {
  const double hx = cf.cs * (cf.W[a] + (a == 6 ? cf.W[10] : a == 7 ? cf.W[11] : 0.0));
  const double hxN = cf.ct * cf.We[a];
  const double hu = cf.cs * (a == 0 ? cf.W[8] + cf.W[10] : cf.W[9] + cf.W[11]);
  const double cxu = (a == 6) ? cf.cs * cf.W[10] : (a == 7) ? cf.cs * cf.W[11] : 0.0;
  const double gs_scale = -cf.cs * cf.W[0], gsN = -cf.ct * cf.We[0] * (s0 + cf.s_target);
  const RowDef Rp = row_path(cf, a), Rt = row_term(cf, a);
  double m = 0.0;
  for (int k = q; k <= N; k += 4) {
    const bool term = (k == N);
    const double xa_ = pX[k * 8 + a];
    if (!term) {
      const double gx = (a == 0) ? gs_scale * (s0 + cf.s_target * k / N) : 0.0;
      m += xa_ * (0.5 * hx * xa_ + gx);
      if (a >= 6) m -= cxu * pU[k * 2 + (a - 6)] * xa_;
      if (a < 2) { const double ua_ = pU[k * 2 + a]; m += 0.5 * hu * ua_ * ua_; }
    } else {
      m += xa_ * (0.5 * hxN * xa_ + ((a == 0) ? gsN : 0.0));
    }
    if (NL) {
      const double Z = (term ? cf.ct : cf.cs) * cf.slack_Z, zz = (term ? cf.ct : cf.cs) * cf.slack_z;
      if (term ? (a == 2 || a == 3) : (a >= 6 && k > 0)) {
        const RowDef& R = term ? Rt : Rp;
        const double v = row_value(R, term, 0.0, pX[k * 8 + R.ix]);
        m += soft_penalty(zz, Z, fmax(0.0, R.lb - v) + fmax(0.0, v - R.ub));
      }
      if (a < 2 || (a == 4 && !term)) {
        const double v1 = hl.value(gHC + k * 16, pX + k * 8, term ? 0.0 : pU[k * 2], term ? 0.0 : pU[k * 2 + 1], a);
        m += soft_penalty(zz, Z, fmax(0.0, hl.lb - v1) + fmax(0.0, v1 - hl.ub));
      }
    }
  }
  return m;
}

// L1 merit: cost + weighted equality defects + weighted violations of the hard inequality rows (the exact penalty of
// a soft row is part of the cost)
template <bool NL>
__device__ __forceinline__ double merit_lane(const KCfg& cf, const int N, const int a, const int q, const double s0,
                                             const double x0a, const double* __restrict__ pX,
                                             const double* __restrict__ pU, const double* __restrict__ phi,
                                             const double* __restrict__ pW, const double* __restrict__ gHC, const HLane& hl) {
  const RowDef Rp = row_path(cf, a), Rt = row_term(cf, a);
  double m = cost_lane<NL>(cf, N, a, q, s0, pX, pU, gHC, hl);
  for (int k = q; k <= N; k += 4) {
    const bool term = (k == N);
    const double xa_ = pX[k * 8 + a];
    // equality defects: x0 embedding at k = 0, dynamics of interval k - 1 otherwise
    const double defect = (k == 0) ? xa_ - x0a : phi[(k - 1) * 8 + a] - xa_;
    m += pW[k * 24 + a] * fabs(defect);
    // inequality row a of this stage
    const RowDef& R = term ? Rt : Rp;
    const bool act = term ? (a < 4) : (a < 4 || k > 0);
    const bool soft = NL && (term ? (a == 2 || a == 3) : (a >= 6));
    if (act && !soft) {
      const double v = row_value(R, term, pU[(term ? 0 : k) * 2 + R.iu], pX[k * 8 + R.ix]);
      m += pW[k * 24 + 8 + a] * fmax(0.0, R.lb - v) + pW[k * 24 + 16 + a] * fmax(0.0, v - R.ub);
    }
  }
  return m;
}

// End of a solve: cost of the OCP at the returned iterate -> stats[7]; u0 = u_0 of the iterate (SQP mode: with a
// shortened last step it is not the QP's control); an instance whose solve failed (status 1 / 3 / 4) gets the iterate
// it had before the call back, as the reference leaves x_pred / u_pred untouched when solver.solve() fails
// (new_python/controller.py:397-403).
template <bool NL>
__global__ void __launch_bounds__(32) solve_finish_kernel(const KCfg cf, const Work w, const double* __restrict__ x0,
                                                          const double* __restrict__ snap_x, const double* __restrict__ snap_u,
                                                          double* __restrict__ u0_out, const int closed_loop_mask) {
  const int lane = threadIdx.x, a = lane & 7, q = lane >> 3, b = blockIdx.x;
  if (closed_loop_mask && w.active && !w.active[b]) return;   // closed loop: retired instance
  const int N = cf.N, N1 = N + 1;
  double* pX = w.x + (size_t)b * N1 * 8;
  double* pU = w.u + (size_t)b * N * 2;
  double itT = 0.0, itD = 0.0;
  if (NL) { itT = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TT]; itD = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TD]; }
  const HLane hl(cf, a, itT, itD);
  const double c = wsum(cost_lane<NL>(cf, N, a, q, x0[(size_t)b * 8], pX, pU, NL ? w.HC + (size_t)b * N1 * 16 : nullptr, hl));
  if (lane == 0) w.stats[(size_t)b * 8 + 7] = c;
  if (u0_out && lane < 2) u0_out[(size_t)b * 2 + lane] = pU[lane];
  const int st = w.status[b];
  if (snap_x && st != IHM2B_SUCCESS && st != IHM2B_MAXITER) {
    const double* sx = snap_x + (size_t)b * N1 * 8;
    const double* su = snap_u + (size_t)b * N * 2;
    for (int i = lane; i < N1 * 8; i += 32) pX[i] = sx[i];
    for (int i = lane; i < N * 2; i += 32) pU[i] = su[i];
  }
}

void launch_solve_finish(const KCfg& cf, const Work& w, const double* x0, const double* snap_x, const double* snap_u,
                         double* u0, int closed_loop_mask, cudaStream_t stream) {
  if (cf.nl) solve_finish_kernel<true><<<cf.B, 32, 0, stream>>>(cf, w, x0, snap_x, snap_u, u0, closed_loop_mask);
  else solve_finish_kernel<false><<<cf.B, 32, 0, stream>>>(cf, w, x0, snap_x, snap_u, u0, closed_loop_mask);
}

// after the QP: weights, merit at the base iterate (w.phi still holds the base linearisation), the full step
template <bool NL>
__global__ void __launch_bounds__(32) merit_base_kernel(const KCfg cf, const Work w, const LineSearch ls,
                                                        const double* __restrict__ x0, const int first) {
  const int lane = threadIdx.x, a = lane & 7, q = lane >> 3, b = blockIdx.x;
  if (!w.active[b]) {   // converged before this QP, or the QP failed
    if (lane == 0) ls.mask[b] = 0;
    return;
  }
  const int N = cf.N, N1 = N + 1;
  constexpr int RS = NL ? 16 : 8;
  const double* pPI = w.xa + (size_t)b * N1 * 8;
  const double* pLL = w.LT + (size_t)b * N1 * 4 * RS;
  const double* pLU = pLL + N1 * RS;
  double* pW = ls.wgt + (size_t)b * N1 * 24;
  const double* gHC = NL ? w.HC + (size_t)b * N1 * 16 : nullptr;
  double itT = 0.0, itD = 0.0;
  if (NL) { itT = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TT]; itD = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TD]; }
  const HLane hl(cf, a, itT, itD);
  for (int k = q; k <= N; k += 4) {
    const bool term = (k == N);
    const bool act = term ? (a < 4) : (a < 4 || k > 0);
    double pe = pPI[k * 8 + a];
    if (k == 0) {
      // multiplier of the embedding x_0 = x0 (the QP eliminates x_0 and never forms it): stationarity in x_0 at
      // the QP solution, pi_0 = (H z + g)_x + A_0' pi_1 - C_x'(lamL - lamU); only the two rate rows touch x_0
      const double* qX = w.x + (size_t)b * N1 * 8;
      const double* qU = w.u + (size_t)b * N * 2;
      const double* gBA = w.BA + (size_t)b * N * 80;
      const double hx = cf.cs * (cf.W[a] + (a == 6 ? cf.W[10] : a == 7 ? cf.W[11] : 0.0));
      pe = hx * qX[a] + ((a == 0) ? -cf.cs * cf.W[0] * x0[(size_t)b * 8] : 0.0);
      if (a >= 6) pe -= cf.cs * cf.W[4 + a] * qU[a - 6] - (pLL[a - 4] - pLU[a - 4]);
      if (NL)   // the h rows of stage 0 depend on x_0 too
        pe -= gHC[hl.so1] * (pLL[8 + hl.r1] - pLU[8 + hl.r1]) + ((a < 3) ? gHC[4 + a] : hl.hc2r) * (pLL[8 + hl.r2] - pLU[8 + hl.r2]);
      for (int r = 0; r < 8; r++) pe = fma(gBA[r * 10 + 2 + a], pPI[8 + r], pe);
    }
    const double m[3] = {fabs(pe), act ? fabs(pLL[k * RS + a]) : 0.0, act ? fabs(pLU[k * RS + a]) : 0.0};
#pragma unroll
    for (int j = 0; j < 3; j++) {
      double* wp = pW + k * 24 + j * 8 + a;
      *wp = first ? m[j] : fmax(m[j], 0.5 * (*wp + m[j]));
    }
  }
  __syncwarp();
  const double* bx = ls.base_x + (size_t)b * N1 * 8;
  const double* bu = ls.base_u + (size_t)b * N * 2;
  const double m0 = wsum(merit_lane<NL>(cf, N, a, q, x0[(size_t)b * 8], x0[(size_t)b * 8 + a], bx, bu,
                                        w.phi + (size_t)b * N * 8, pW, gHC, hl));
  const double* pX = w.x + (size_t)b * N1 * 8;
  const double* pU = w.u + (size_t)b * N * 2;
  double* sx = ls.step_x + (size_t)b * N1 * 8;
  double* su = ls.step_u + (size_t)b * N * 2;
  for (int i = lane; i < N1 * 8; i += 32) sx[i] = pX[i] - bx[i];
  for (int i = lane; i < N * 2; i += 32) su[i] = pU[i] - bu[i];
  if (lane == 0) { ls.m0[b] = m0; ls.alpha[b] = 1.0; ls.mask[b] = 1; }
}

// after the linearisation at the trial iterate: accept, or shorten the step and write the next trial iterate
template <bool NL>
__global__ void __launch_bounds__(32) merit_trial_kernel(const KCfg cf, const Work w, const LineSearch ls,
                                                         const double* __restrict__ x0, int32_t* __restrict__ sqp_active,
                                                         int32_t* __restrict__ status_out) {
  const int lane = threadIdx.x, a = lane & 7, q = lane >> 3, b = blockIdx.x;
  if (!ls.mask[b]) return;
  const int N = cf.N, N1 = N + 1;
  double* pX = w.x + (size_t)b * N1 * 8;
  double* pU = w.u + (size_t)b * N * 2;
  double itT = 0.0, itD = 0.0;
  if (NL) { itT = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TT]; itD = 1.0 / w.params[(size_t)b * NP + IHM2B_P_TD]; }
  const HLane hl(cf, a, itT, itD);
  const double m1 = wsum(merit_lane<NL>(cf, N, a, q, x0[(size_t)b * 8], x0[(size_t)b * 8 + a], pX, pU,
                                        w.phi + (size_t)b * N * 8, ls.wgt + (size_t)b * N1 * 24,
                                        NL ? w.HC + (size_t)b * N1 * 16 : nullptr, hl));
  const double alpha = ls.alpha[b];
  __syncwarp();
  if (m1 < ls.m0[b] || alpha <= LS_ALPHA_MIN) {
    if (lane == 0) {
      ls.mask[b] = 0;
      w.stats[(size_t)b * 8 + 2] = alpha;   // statistics slot [2] carries the accepted step length in this mode
      if (!(m1 == m1)) {                    // the integrator failed at the shortest step
        w.status[b] = IHM2B_FAILURE;
        if (status_out) status_out[b] = IHM2B_FAILURE;
        sqp_active[b] = 0;
      }
    }
    return;
  }
  const double an = fmax(LS_REDUCTION * alpha, LS_ALPHA_MIN);
  const double* bx = ls.base_x + (size_t)b * N1 * 8;
  const double* bu = ls.base_u + (size_t)b * N * 2;
  const double* sx = ls.step_x + (size_t)b * N1 * 8;
  const double* su = ls.step_u + (size_t)b * N * 2;
  for (int i = lane; i < N1 * 8; i += 32) pX[i] = fma(an, sx[i], bx[i]);
  for (int i = lane; i < N * 2; i += 32) pU[i] = fma(an, su[i], bu[i]);
  if (lane == 0) ls.alpha[b] = an;
}

void launch_merit_base(const KCfg& cf, const Work& w, const LineSearch& ls, const double* x0, int first,
                       cudaStream_t stream) {
  if (cf.nl) merit_base_kernel<true><<<cf.B, 32, 0, stream>>>(cf, w, ls, x0, first);
  else merit_base_kernel<false><<<cf.B, 32, 0, stream>>>(cf, w, ls, x0, first);
}
void launch_merit_trial(const KCfg& cf, const Work& w, const LineSearch& ls, const double* x0, int32_t* sqp_active,
                        int32_t* status, cudaStream_t stream) {
  if (cf.nl) merit_trial_kernel<true><<<cf.B, 32, 0, stream>>>(cf, w, ls, x0, sqp_active, status);
  else merit_trial_kernel<false><<<cf.B, 32, 0, stream>>>(cf, w, ls, x0, sqp_active, status);
}

// ---------------------------------------------------------------------------------------
// Launch order for the next solve of this handle: problems sorted by decreasing IPM iteration count of
// the solve just finished (counting sort, counts 0..63).  A solve is a serial chain whose length is its
// iteration count (13 .. 50 here), and a batch is only 2-3 waves of resident warps: with the long problems
// first the last wave is not left waiting for a 50-iteration straggler.  One block.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) build_order_kernel(int B, const int32_t* __restrict__ last_it, int32_t* __restrict__ order) {
  __shared__ int cnt[64], start[64];
  const int tid = threadIdx.x;
  if (tid < 64) cnt[tid] = 0;
  __syncthreads();
  for (int b = tid; b < B; b += blockDim.x) {
    const int key = min(63, max(0, last_it[b]));
    atomicAdd(&cnt[key], 1);
  }
  __syncthreads();
  if (tid == 0) {
    int acc = 0;
    for (int k = 63; k >= 0; k--) { start[k] = acc; acc += cnt[k]; }
  }
  __syncthreads();
  for (int b = tid; b < B; b += blockDim.x) {
    const int key = min(63, max(0, last_it[b]));
    order[atomicAdd(&start[key], 1)] = b;
  }
}

void launch_build_order(int B, const int32_t* last_it, int32_t* order, cudaStream_t stream) {
  build_order_kernel<<<1, 1024, 0, stream>>>(B, last_it, order);
}

int qp_smem_bytes(int N, int nl) { return (N + 1) * (nl ? PER_STAGE_NL : PER_STAGE) * (int)sizeof(double); }   // per problem

namespace {
// the dynamic shared-memory limit is a per-device attribute of each kernel instance: remembered per (instance, device)
template <int WPB, class K>
int launch_qp_instance(K kernel, int slot, const KCfg& cf, const Work& w, const double* x0, double* u0, int32_t* status,
                       cudaStream_t stream) {
  static int configured[4][64] = {};
  const int bytes = WPB * qp_smem_bytes(cf.N, cf.nl);
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 1;
  if (dev < 0 || dev >= 64 || configured[slot][dev] < bytes) {
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) { cudaGetLastError(); return 1; }
    if (dev >= 0 && dev < 64) configured[slot][dev] = bytes;
  }
  kernel<<<(cf.B + WPB - 1) / WPB, 32 * WPB, bytes, stream>>>(cf, w, x0, u0, status);
  return 0;
}
}  // namespace

int launch_qp_rti(const KCfg& cf, const Work& w, const double* x0, double* u0, int32_t* status,
                  cudaStream_t stream) {
  if (cf.nl) {
    if (cf.N == 40) return launch_qp_instance<QP_WPB>(qp_rti_kernel<40, true, QP_WPB>, 2, cf, w, x0, u0, status, stream);
    return launch_qp_instance<1>(qp_rti_kernel<0, true, 1>, 3, cf, w, x0, u0, status, stream);
  }
  if (cf.N == 40) return launch_qp_instance<QP_WPB>(qp_rti_kernel<40, false, QP_WPB>, 0, cf, w, x0, u0, status, stream);
  return launch_qp_instance<1>(qp_rti_kernel<0, false, 1>, 1, cf, w, x0, u0, status, stream);
}

}  // namespace ihm2b
