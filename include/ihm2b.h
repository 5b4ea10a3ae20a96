/* ihm2b.h -- C ABI of libihm2b.so, the B200-native batched NMPC engine for the ihm2
 * path-parametric bicycle-model controller.
 *
 * This is the drop-in boundary for the one hot path named by BASELINE.json: what the
 * reference reaches through `AcadosOcpSolver` (ctypes over the acados-generated
 * libacados_ocp_solver_ihm2_fkin6.so) and, in C++, through the generated
 * `ihm2_fkin6_acados_*` capsule API.  Each entry point cites the reference call it
 * replaces (paths relative to the reference tree).
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, int return (0 = IHM2B_OK); no exceptions.
 *  - `_dev` pointers are CUDA device pointers owned by the caller (PyTorch tensors);
 *    `_host` pointers are host memory.  The handle owns all workspace.
 *  - `stream` is a cudaStream_t passed as void*; every `_dev` call is asynchronous on it.
 *  - One handle per (device, stream); a handle is not thread-safe (same as an acados capsule,
 *    src/ihm2/src/mpc_control_node.cpp:274-304).
 *  - Batch-major row layouts at the boundary: x [B][N+1][8], u [B][N][2], x0 [B][8] ...
 *    (the reference keeps one row per stage, mpc_control_node.cpp:71); the library
 *    transposes to its internal structure-of-arrays layout on the device.
 *  - Per-instance NLP status codes follow acados (dpc/main.py:287-293):
 *    0 success, 1 failure (NaN), 2 max iterations, 3 min step, 4 QP failure.
 */
#ifndef IHM2B_H
#define IHM2B_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IHM2B_NX 8   /* IHM2_FKIN6_NX */
#define IHM2B_NU 2   /* IHM2_FKIN6_NU */
#define IHM2B_NY 12  /* IHM2_FKIN6_NY  (python/mpc.py:48) */
#define IHM2B_NYN 8  /* IHM2_FKIN6_NYN (python/mpc.py:62) */
#define IHM2B_NP 28  /* per-instance vehicle parameters, see ihm2b_param_index */
#define IHM2B_NX10 15 /* states of the dyn10 plant (python/models.py:609-801) */
#define IHM2B_NU10 5  /* its controls: four wheel torque commands and the steering command */

enum ihm2b_error { IHM2B_OK = 0, IHM2B_EINVAL = 1, IHM2B_ECUDA = 2, IHM2B_ENOMEM = 3 };
enum ihm2b_integrator { IHM2B_ERK4 = 0, IHM2B_IRK_GL4 = 1 };
enum ihm2b_nlp_solver { IHM2B_SQP_RTI = 0, IHM2B_SQP = 1 };
enum ihm2b_globalization { IHM2B_FIXED_STEP = 0, IHM2B_MERIT_BACKTRACKING = 1 };
enum ihm2b_sim_model { IHM2B_SIM_FKIN6 = 0, IHM2B_SIM_FDYN6 = 1, IHM2B_SIM_KIN6_DYN6 = 2,
  /* Cartesian plants of the ROS sim node (python/models.py:168-229 kin6, :310-452 dyn6): state (X, Y, phi, v_x, v_y, r,
   * T, delta); KIN6C_DYN6C switches on hypot(v_x, v_y) < sim_switch (v_dyn, sim_node.cpp:199) and all three apply the
   * node's no-reverse clamp after the step (sim_node.cpp:256-260) */
  IHM2B_SIM_KIN6C = 3, IHM2B_SIM_DYN6C = 4, IHM2B_SIM_KIN6C_DYN6C = 5 };
enum ihm2b_status {
  IHM2B_SUCCESS = 0, IHM2B_FAILURE = 1, IHM2B_MAXITER = 2, IHM2B_MINSTEP = 3, IHM2B_QP_FAILURE = 4
};

/* per-instance vehicle parameter vector (python/constants.py:43-111) */
enum ihm2b_param_index {
  IHM2B_P_M = 0, IHM2B_P_IZ, IHM2B_P_LR, IHM2B_P_LF, IHM2B_P_CM0, IHM2B_P_CR0, IHM2B_P_CR1,
  IHM2B_P_CR2, IHM2B_P_TT, IHM2B_P_TD, IHM2B_P_ZCG, IHM2B_P_TRACKW, IHM2B_P_CDOWN,
  IHM2B_P_KTV, IHM2B_P_BA, IHM2B_P_CA, IHM2B_P_DA, IHM2B_P_EA, IHM2B_P_G,
  IHM2B_P_CROSS, /* 1: literal slip-angle cross assignment of python/models.py:543-546 */
  /* dyn10 plant only (python/constants.py:85-101): wheel radius / inertia, dynamic / static friction of the wheel
   * bearing, longitudinal Pacejka coefficients */
  IHM2B_P_RW, IHM2B_P_IW, IHM2B_P_KD, IHM2B_P_KS, IHM2B_P_BS, IHM2B_P_CS, IHM2B_P_DS, IHM2B_P_ES
};

/* Everything `IHM2Controller.__init__` fixes at construction
 * (new_python/controller.py:250-366, python/mpc.py:29-101). */
typedef struct ihm2b_config {
  int32_t batch;          /* B independent problems */
  int32_t N;              /* horizon Nf (controller.py:254) */
  int32_t n_tracks;       /* number of (s_ref, kappa_ref) tables */
  int32_t n_grid;         /* entries per table (3 * NUMBER_SPLINE_INTERVALS = 1500) */
  int32_t integrator;     /* ihm2b_integrator (controller.py:305 / old/generate.py:24) */
  int32_t nlp_solver;     /* ihm2b_nlp_solver (controller.py:301 / old/generate.py:21) */
  int32_t nlp_max_iter;   /* nlp_solver_max_iter (controller.py:302); ignored for RTI */
  int32_t globalization;  /* ihm2b_globalization (controller.py:308) */
  int32_t qp_max_iter;    /* interior-point iteration cap */
  int32_t nonlinear_constraints; /* 1: add soft h(x,u) of old/generate_acaods_interface.py:190-212 */
  int32_t sim_model;      /* ihm2b_sim_model (python/main.py:346, 475-502) */
  int32_t sim_steps;      /* integrator sub-steps of the plant (python/main.py:398: 100) */
  int32_t schedule_by_history; /* 1 (default of the Python layer): the QP kernel launches the problems of a solve
                             longest first, by the IPM iteration counts of the handle's previous solve (in a
                             closed loop: the previous control step).  Scheduling only; results are unchanged. */
  int32_t qp_tolerate_maxiter; /* 1: a QP whose interior point stops at qp_max_iter without reaching the tolerances is
                             used as it is and the solve goes on / reports success, as acados' SQP / SQP_RTI do with
                             ACADOS_MAXITER from the QP solver (a vanishing step or NaN stay QP failures);
                             0 (default): such a QP ends the solve with status 4 unless it is converged to 1e-6 */
  double dt;              /* sampling time (controller.py:255) */
  double s_target;        /* controller.py:256 */
  double n_max, v_x_max, T_max, delta_max; /* python/mpc.py:77-89 */
  double g_T_max;         /* t_T * T_dot_max      (python/mpc.py:98-99) */
  double g_delta_max;     /* t_delta * delta_dot_max */
  double a_lat_max;       /* only with nonlinear_constraints */
  double W[IHM2B_NY];     /* diag of stage W (controller.py:326-343) */
  double W_e[IHM2B_NYN];  /* diag of terminal W (controller.py:350-363) */
  double cost_scale_stage;    /* acados multiplies stage costs by tf/N */
  double cost_scale_terminal; /* and leaves the terminal cost unscaled */
  double qp_mu0;          /* IPM start barrier parameter */
  double qp_tol;          /* IPM tolerance on stationarity / equality / inequality residuals (relative) */
  double qp_tol_comp;     /* IPM tolerance on complementarity; the barrier parameter is floored at 0.1 x this */
  double nlp_tol;         /* SQP KKT tolerance (SQP mode) */
  double irk_A[16], irk_b[4]; /* 4-stage collocation tableau of the controller integrator */
  double sim_A[16], sim_b[4]; /* 4-stage tableau of the plant integrator (Radau IIA) */
  double sim_switch;      /* fkin6 -> fdyn6 when v^2 sin(beta)/l_R > sim_switch (main.py:482-489) */
  double track_half_L, track_half_W; /* car half length / width for the track constraint */
  double slack_z, slack_Z; /* L1 / L2 slack penalties (old/...interface.py:383-390) */
  double T_dot_max, delta_dot_max; /* bounds of the T_dot / delta_dot rows of h (old/...interface.py:418-436) */
  /* ---- interior-point strategy (HPIPM mode presets are filled in by the host layer, see
   * ihm2_b200/_abi.py:hpipm_mode; new_python/controller.py:304 selects SPEED_ABS) ---- */
  int32_t qp_split_step;  /* 1: separate primal and dual step lengths (HPIPM split_step, on in the SPEED modes) */
  int32_t qp_infeas_exit; /* 1: stop a QP whose primal infeasibility stagnates while the barrier parameter has
                             collapsed (status 4 after a few iterations instead of qp_max_iter); HPIPM has no such
                             test, 0 reproduces it */
  double qp_thr0;         /* floor of the initial slacks t (HPIPM thr0 = 0.1) */
  int32_t linearize_fp32; /* 1: mixed precision -- the IRK collocation system, its Newton solves and the forward sensitivities
                             in fp32 (the iterate, [B A], c and the QP stay fp64); north_star's reduced-precision bar is
                             1e-4 relative per SQP-RTI step.  IRK integrator only. */
  int32_t reserved_;
} ihm2b_config;

typedef struct ihm2b_handle ihm2b_handle;

/* ---- life cycle ----------------------------------------------------------------------
 * replaces ihm2_fkin6_acados_create_capsule / _create / _free / _free_capsule
 * (mpc_control_node.cpp:274-284, 399-411) and AcadosOcpSolver(...) (python/mpc.py:111). */
int ihm2b_create(const ihm2b_config* cfg, ihm2b_handle** out);
int ihm2b_destroy(ihm2b_handle* h);
const char* ihm2b_last_error(void);
int ihm2b_version(void);

/* ---- problem data --------------------------------------------------------------------
 * replaces solver.set(i, "p", p) for all stages (controller.py:320-323,
 * ihm2_fkin6_acados_update_params, mpc_control_node.cpp:360-364).
 * s_ref / kappa_ref: host, [n_tracks][n_grid]; widths (optional, may be NULL): right/left
 * track widths, same shape.  track_index: host int32 [B].  params: host [B][IHM2B_NP]. */
int ihm2b_set_tracks(ihm2b_handle* h, const double* s_ref_host, const double* kappa_ref_host,
                     const double* right_width_host, const double* left_width_host);
int ihm2b_set_track_index(ihm2b_handle* h, const int32_t* track_index_host);
int ihm2b_set_params(ihm2b_handle* h, const double* params_host);

/* ---- iterate (x_pred / u_pred) -------------------------------------------------------
 * replaces solver.set(j, "x"|"u", v) and solver.get(i, "x"|"u")
 * (controller.py:388-393, 402-403; ocp_nlp_out_get, mpc_control_node.cpp:202-206). */
int ihm2b_set_iterate_dev(ihm2b_handle* h, const double* x_dev, const double* u_dev, void* stream);
int ihm2b_get_iterate_dev(ihm2b_handle* h, double* x_dev, double* u_dev, void* stream);
/* controller.py:313-317: x_j = (-6 + j dt, 0...), u_j = (T_max, 0) */
int ihm2b_reset_iterate(ihm2b_handle* h, void* stream);
/* controller.py:388-393 shift warm start (literal: x_{N-1} = x_N = old x_N, u_{N-1} = 0) */
int ihm2b_shift_iterate(ihm2b_handle* h, void* stream);

/* ---- the hot entry point -------------------------------------------------------------
 * replaces the body of IHM2Controller.compute_control after the shift
 * (controller.py:370-385 x0 embedding + yref, :396 solver.solve(), :405 get(0,"u");
 * ihm2_fkin6_acados_solve, mpc_control_node.cpp:177-206).
 * x0_dev [B][8] in; u0_dev [B][2], status_dev [B] out.
 * nlp_solver = IHM2B_SQP_RTI: one linearisation + one QP + full step (status 0 / 1 / 4).
 * nlp_solver = IHM2B_SQP (acados ocp_nlp_sqp as configured by controller.py:301-302): up to
 * nlp_max_iter rounds of { linearise, KKT residual test against nlp_tol with the multipliers
 * of the last QP, QP, full step }; instances drop out as they converge (status 0) or their QP
 * fails (1 / 4), the rest end with status 2 (ACADOS_MAXITER, tolerated by the reference,
 * mpc_control_node.cpp:190-197).  stats: [qp iterations (sum), sqp iterations, mu, res_stat,
 * res_eq, res_ineq, res_comp, cost of the OCP at the returned iterate].
 * nonlinear_constraints = 1 adds the soft rows h(x, u) of the gen-1 OCP (right / left track boundary with the car
 * footprint, T_dot, delta_dot, a_lat: old/generate_acaods_interface.py:312-325, bounds / slack sets :351-449), makes the
 * state bounds on T and delta soft (idxsbx = [2, 3]) and eliminates the slacks inside the interior point; it needs the
 * width tables of ihm2b_set_tracks.
 * globalization = IHM2B_MERIT_BACKTRACKING (controller.py:308): each step is scaled by alpha = 1, 0.7, 0.49, ...
 * (not below 0.05) until the L1 merit function (cost + multiplier-weighted equality defects and inequality
 * violations) of the trial iterate is below that of the current one; multipliers take the full step; stats[2]
 * then carries the accepted alpha of the last iteration instead of mu. */
int ihm2b_solve_dev(ihm2b_handle* h, const double* x0_dev, double* u0_dev, int32_t* status_dev,
                    void* stream);
/* Same call with HOST buffers: H2D of x0, shift (if do_shift), solve, D2H of u0/status,
 * synchronous.  This is the shape of the reference's numpy-in / numpy-out call. */
int ihm2b_compute_control_host(ihm2b_handle* h, const double* x0_host, double* u0_host,
                               int32_t* status_host, int do_shift);

/* x_pred / u_pred to HOST buffers (the reference reads every stage back after each solve,
 * new_python/controller.py:402-403: solver.get(i, "x" | "u")): x_host [B][N+1][8], u_host [B][N][2], either may be
 * NULL; synchronous.  Page-locked destination buffers copy at the full PCIe rate. */
int ihm2b_get_iterate_host(ihm2b_handle* h, double* x_host, double* u_host);

/* ---- pieces of the path, exposed for parity tests ------------------------------------
 * integrator + forward sensitivities at the current iterate (acados sim_erk / sim_irk inside
 * the SQP preparation phase).  Outputs batch-major: A [B][N][8][8], Bm [B][N][8][2],
 * phi [B][N][8] (phi_k = integrator(x_k, u_k)). */
int ihm2b_linearize_dev(ihm2b_handle* h, double* A_dev, double* B_dev, double* phi_dev,
                        void* stream);
/* per-instance statistics [B][8]: qp iterations, sqp iterations, final mu, res_stat, res_eq,
 * res_ineq, res_comp, cost */
int ihm2b_get_stats_dev(ihm2b_handle* h, double* stats_dev, void* stream);

/* ---- multipliers of the last QP (KKT certificates) -------------------------------------
 * replaces solver.get(i, "pi" | "lam") of AcadosOcpSolver.  RS = 8 inequality row slots per stage (16 with
 * nonlinear_constraints: slots 8..12 are the rows of h); slot order of a path stage: (u_T, u_delta, u_T - T, u_delta - delta,
 * n, v_x, T, delta), terminal stage: idxbx_e = (n, v_x, v_y, r); absent rows (slots 4..7 at stage 0, >= 4 at stage N,
 * the padding) carry zeros.  pi_dev [B][N+1][8]: pi_k multiplies x_k = A x_{k-1} + B u_{k-1} + c (k >= 1; pi_0 is not
 * formed, x_0 is eliminated).  lamL_dev / lamU_dev [B][N+1][RS]: multipliers of the lower / upper side of each row.
 * Sign convention: stationarity reads H z + g + [B' pi_{k+1}; A' pi_{k+1} - pi_k] - C'(lamL - lamU) = 0. */
int ihm2b_get_multipliers_dev(ihm2b_handle* h, double* pi_dev, double* lamL_dev, double* lamU_dev, void* stream);
/* slacks of the soft rows and the multipliers of s >= 0, [B][N+1][8] each (nonlinear_constraints only; any pointer
 * may be NULL).  Slot j of a path stage: j = 0, 1 the track rows h0, h1; 4 a_lat; 6, 7 the soft bounds on T, delta
 * (stages >= 1); terminal stage: 0, 1 the track rows, 2, 3 the soft bounds on v_y, r; other slots unused. */
int ihm2b_get_slacks_dev(ihm2b_handle* h, double* sL_dev, double* sU_dev, double* lamSL_dev, double* lamSU_dev,
                         void* stream);
/* the linearised rows of h at the last linearisation point, [B][N+1][16] (layout: csrc/common.cuh, Work::HC) */
int ihm2b_get_h_rows_dev(ihm2b_handle* h, double* hc_dev, void* stream);

/* ---- plant ---------------------------------------------------------------------------
 * replaces AcadosSimSolver.simulate(x, u) with the model switch of python/main.py:475-502.
 * x_dev [B][8], u_dev [B][2] -> xnext_dev [B][8]; model_dev (optional) gets the model used. */
int ihm2b_sim_step_dev(ihm2b_handle* h, const double* x_dev, const double* u_dev,
                       double* xnext_dev, int32_t* model_dev, void* stream);

/* The same plant step for callers that cannot allocate device memory around the call (CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Its scratch -- the launch order of the kinematic / dynamic model partition, [B] + 2 counters of int32 --
 * lives in a static device array inside the library (256 KB, held by every GPU that loads the library until the process
 * exits) instead of a buffer of the handle; the call itself makes no cudaMalloc / cudaFree or other allocating call.
 * Limit: cfg.batch <= IHM2B_SIM_STATIC_MAX_BATCH, larger handles get IHM2B_EINVAL.  Same arguments and results as
 * ihm2b_sim_step_dev (the caller passes inputs and outputs; model_dev optional).  All calls on a device share the array:
 * calls that overlap on the GPU (different streams, or two handles from two threads) race -- serialise them.  Nothing
 * is carried from call to call (counters cleared, the order rewritten each time). */
#define IHM2B_SIM_STATIC_MAX_BATCH 65536
int ihm2b_sim_step_static_dev(ihm2b_handle* h, const double* x_dev, const double* u_dev,
                              double* xnext_dev, int32_t* model_dev, void* stream);

/* The dyn10 plant (`fdyn10_model`, python/models.py:609-801): Frenet pose, body velocities, four wheel
 * speeds with slip-ratio (longitudinal Pacejka) and slip-angle (lateral Pacejka) tyre forces on load-transfer vertical
 * loads, four torque lags and the steering lag.  x10_dev [B][15] = (s, n, psi, v_x, v_y, r, omega_FL, omega_FR,
 * omega_RL, omega_RR, tau_FL, tau_FR, tau_RL, tau_RR, delta); u5_dev [B][5] = (u_tau_FL, u_tau_FR, u_tau_RL, u_tau_RR,
 * u_delta) (python/main.py:490-500 feeds 0.25 u_T to each wheel).  Same integrator as the other plants: Radau IIA,
 * 4 stages, cfg.sim_steps sub-steps over cfg.dt.  NaN rows where the integrator failed. */
int ihm2b_sim_step_dyn10_dev(ihm2b_handle* h, const double* x10_dev, const double* u5_dev, double* xnext10_dev,
                             void* stream);

/* ---- closed loop ---------------------------------------------------------------------
 * python/main.py:448-517 with the NMPC call enabled: n_steps iterations of
 * shift -> solve -> plant step on the device, no host round trips.
 * x_dev [B][8] in/out (current state); lap_length_host [n_tracks].
 * Optional logs (may be NULL): xlog_dev [n_steps+1][B][8], ulog_dev [n_steps][B][2].
 * done_dev [B]: step index at which s > lap_length + 1 (or -1), fail_dev [B]: step index of
 * the first failure (solver status not in {0,2} or NaN in the plant), or -1.
 * Optional per-instance results (may be NULL): lap_time_dev [B], the lap time as python/main.py:561-570 computes it
 * from the state log -- dt x (second - first index at which mod(s, lap_length) jumps by more than 1 between
 * consecutive states), NaN while fewer than two jumps have been seen; cost_sum_dev [B], the OCP cost at the solution
 * summed over the control steps taken ("mean cost"). */
int ihm2b_closed_loop_dev(ihm2b_handle* h, int32_t n_steps, double* x_dev,
                          const double* lap_length_host, double* xlog_dev, double* ulog_dev,
                          int32_t* done_dev, int32_t* fail_dev, double* lap_time_dev, double* cost_sum_dev,
                          void* stream);

/* ---- DPC rollout ---------------------------------------------------------
 * the differentiable-predictive-control loss of dpc/main.py: kin4 model (:56-84), one RK4 step of dt per
 * stage (:87-97), IHM1.5 tracking cost (:374-409; stage 0 has no tracking term, the last stage uses the _f
 * weights), and its gradient in the controls -- what a policy trained through the rollout receives (the
 * reference leaves the policy itself empty, dpc/main.py:617-635).  No handle: the model constants are those of
 * dpc/main.py:23-29.  Device buffers: x0_dev [B][4] = (X, Y, phi, v), u_dev [B][Nf][2] = (T, delta),
 * ref_dev [B][Nf+1][4] = (X_ref, Y_ref, phi_ref, v_ref) -> cost_dev [B], grad_dev [B][Nf][2],
 * xtraj_dev [B][Nf+1][4] (optional, may be NULL).  weights_host [10] = (q_lon, q_lat, q_phi, q_v, r_T, r_delta,
 * q_lon_f, q_lat_f, q_phi_f, q_v_f).  Nf must be 40 (dpc/main.py:31) or 80. */
int ihm2b_dpc_rollout_grad_dev(int64_t batch, int32_t Nf, double dt, const double* x0_dev, const double* u_dev,
                               const double* ref_dev, const double* weights_host, double* cost_dev,
                               double* grad_dev, double* xtraj_dev, void* stream);

/* fp32 variant of the call above (the optional reduced-precision path; north_star tolerance 1e-4 relative): same
 * layout with float buffers, weights still double on the host */
int ihm2b_dpc_rollout_grad_f32_dev(int64_t batch, int32_t Nf, double dt, const float* x0_dev, const float* u_dev,
                                   const float* ref_dev, const double* weights_host, float* cost_dev,
                                   float* grad_dev, float* xtraj_dev, void* stream);

/* ---- Cartesian pose -> Frenet state ------------------------------------------
 * Track::project (src/ihm2/src/common/tracks.cpp:200-288) followed by the control node's lateral offset and
 * heading error (src/ihm2/src/mpc_control_node.cpp:142-157), for a batch of cars.  No handle.  Device buffers:
 * tables s_ref / X_ref / Y_ref / phi_ref [n_tracks][n_grid] (the columns of the track CSV, tracks.cpp:132-149),
 * track_index_dev [B] (NULL: track 0), pose_dev [B][3] = (X, Y, phi), s_guess_dev [B]; s_tol is the half-width of
 * the search window in metres (2.0 in the node).  Outputs: frenet_dev [B][3] = (s, n, psi); optional (NULL to skip)
 * ref_dev [B][3] = (X_ref, Y_ref, phi_ref) at the projection and flag_dev [B] = 1 where the node would throw
 * "tangent x normal is too small". */
int ihm2b_project_dev(int64_t batch, int32_t n_grid, const double* s_ref_dev, const double* X_ref_dev,
                      const double* Y_ref_dev, const double* phi_ref_dev, const int32_t* track_index_dev,
                      const double* pose_dev, const double* s_guess_dev, double s_tol, double* frenet_dev,
                      double* ref_dev, int32_t* flag_dev, void* stream);

/* ---- diagnostics / measurement aids ------------------------------------------------------
 * number of kernel launches issued by this handle since creation (bench bookkeeping) */
int64_t ihm2b_launch_count(const ihm2b_handle* h);
/* per-kernel device times of ihm2b_solve_dev (RTI mode): when enabled every solve synchronises on its own events;
 * ihm2b_profile_read returns the mean linearisation / QP kernel time in ms since the last enable (ms2[2]) */
int ihm2b_profile_enable(ihm2b_handle* h, int on);
int ihm2b_profile_read(ihm2b_handle* h, double* ms2_host);
/* per-iteration interior-point trace of problem 0 of the last QP: host [64][8] =
 * (res_g, res_b, res_d, res_m, mu, alpha, mu_target, -) per iteration */
int ihm2b_debug_trace(ihm2b_handle* h, double* host64x8);
/* FP64 throughput of the current device in TFLOP/s (the roofline denominator MEASURED_PEAKS.json does not carry):
 * the larger of a DFMA chain (CUDA cores) and a DMMA chain (mma.m8n8k4.f64) */
int ihm2b_measure_fp64_tflops(double* tflops_host);

#ifdef __cplusplus
}
#endif
#endif /* IHM2B_H */
