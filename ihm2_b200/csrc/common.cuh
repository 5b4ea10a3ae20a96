// common.cuh -- shared definitions of libihm2b.so (device side).
//
// Product code: nothing here includes, links or calls anything under oracle/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ihm2b.h"

namespace ihm2b {

constexpr int NX = IHM2B_NX, NU = IHM2B_NU, NZ = NX + NU, NP = IHM2B_NP;
constexpr int NROW = 8;        // two-sided inequality rows per stage (padded)
constexpr int GROUP = 8;       // lanes per problem in the QP kernel (lane = state row)

// Kernel-side copy of the scalar options of ihm2b_config (passed by value).
struct KCfg {
  int B, N, n_grid;
  int integrator, qp_max_iter, qp_tolerate_maxiter, lin_fp32;
  double dt, s_target;
  double n_max, v_x_max, T_max, delta_max, g_T_max, g_delta_max;
  double W[IHM2B_NY], We[IHM2B_NYN];
  double cs, ct;               // cost scaling: stage, terminal
  double mu0, tol, tol_comp, thr0;
  int split_step, infeas_exit;
  double irk_A[16], irk_b[4];
  // soft nonlinear constraints h(x, u) (old/generate_acaods_interface.py:312-325, 351-449), nl = 1
  int nl;
  double half_L, half_W, a_lat_max, T_dot_max, delta_dot_max;
  double slack_z, slack_Z;     // L1 / L2 slack penalties before the cost scaling
};

// Device workspace of one handle.  Batch-major, one contiguous block per problem and per
// stage inside it, so that the 8 lanes that own a problem read 64..640 contiguous bytes.
struct Work {
  // problem data
  const double* s_ref;      // [n_tracks][n_grid]
  const double* kappa_ref;  // [n_tracks][n_grid]
  const double* wr;         // [n_tracks][n_grid] right / left track widths (nl only)
  const double* wl;
  const int32_t* track;     // [B]
  const double* params;     // [B][NP]
  const double* actw;       // [B][8]  (I + h/t_T A)^-1 1, (I + h/t_delta A)^-1 1 (controller tableau)
  const double* actw_sim;   // [B][8]  same for the plant tableau and sub-step
  // iterate (python: x_pred / u_pred), reference layout
  double* x;                // [B][N+1][8]
  double* u;                // [B][N][2]
  // linearisation: BA[k][i][0..1] = B_k row i, BA[k][i][2..9] = A_k row i ; c_k = phi - A x - B u
  double* BA;               // [B][N][8][10]
  double* c;                // [B][N][8]
  double* phi;              // [B][N][8]
  // interior point state: lamL, lamU, tL, tU per row (RS = 8 row slots per stage, 16 with nl)
  double* LT;               // [B][N+1][4][RS]
  double* D;                // [B][N+1][2][RS] second-order terms dlam_aff * dt_aff (L, U) / first-order hand-over
  double* xa;               // [B][N+1][8]     costate pi
  double* atp;              // [B][N][8]       A_k' pi_{k+1}, carried along the interior-point iterations (qp.cu, phase B)
  // soft nonlinear constraints (nl = 1; the interior-point arrays above then carry 16 row slots per stage:
  // slots 0..7 the rows of python/mpc.py, slots 8..12 the rows of h)
  double* HC;               // [B][N+1][16] linearised h rows: [0..2] d h0 / d (s, n, psi), [3] e0; [4..6], [7] the same
                            //              for h1; [8..11] d h4 / d (v_x, v_y, T, delta), [12] e4; [13] e2, [14] e3
                            //              (row value = coefficients . z + e; h2, h3 have the constant coefficients -+1/t)
  double* SS;               // [B][8][N+1][8] slacks of the soft rows: sL, sU, their multipliers, second-order terms / steps
                            //              of both, slack steps; 8 slots per stage (a lane has at most one soft row)
  double* CAB;              // [B][N+1][2][16] corrector terms per state / control: chi = CA - mu_t CB
  // outputs
  double* stats;            // [B][8]
  int32_t* last_it;         // [B] IPM iterations of the last QP (launch order of the next one)
  int32_t* status;          // [B]
  double* dbg;              // [64][8] per-iteration trace of problem 0 (development aid), may be NULL
  // SQP mode (nlp_solver = SQP): instances still iterating, index of the current SQP iteration
  int32_t* active;          // [B], NULL in RTI mode (all instances run)
  int sqp_it;               // -1 in RTI mode
  // launch order of the QP kernel: block i solves problem order[i] (longest first by the iteration
  // counts of the handle's previous solve), NULL = identity
  const int32_t* order;
};

// SQP merit-function line search (globalization = MERIT_BACKTRACKING): per-instance state
struct LineSearch {
  double* base_x;   // [B][N+1][8]  iterate the step starts from
  double* base_u;   // [B][N][2]
  double* step_x;   // [B][N+1][8]  full QP step
  double* step_u;   // [B][N][2]
  double* wgt;      // [B][N+1][24] merit weights per stage: 8 equality (k = 0: x0 embedding, k > 0: dynamics of
                    //              interval k-1), 8 lower-side and 8 upper-side inequality rows
  double* m0;       // [B] merit at the base iterate
  double* alpha;    // [B] current trial step length
  int32_t* mask;    // [B] 1 while the instance is still searching
};

}  // namespace ihm2b
