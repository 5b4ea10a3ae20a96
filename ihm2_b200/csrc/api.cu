// api.cu -- C ABI of libihm2b.so (include/ihm2b.h): handle life cycle, problem data,
// iterate management and the solve / closed-loop entry points.
//
// The handle plays the role of the acados capsule (ihm2_fkin6_acados_create_capsule ...,
// src/ihm2/src/mpc_control_node.cpp:274-304 in the reference): it owns every workspace
// buffer; callers own the I/O buffers they pass in.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

#include "common.cuh"

namespace ihm2b {
int launch_linearize(const KCfg& cf, const Work& w, cudaStream_t stream);
int launch_qp_rti(const KCfg& cf, const Work& w, const double* x0, double* u0, int32_t* status,
                  cudaStream_t stream);
int qp_smem_bytes(int N, int nl);
void launch_build_order(int B, const int32_t* last_it, int32_t* order, cudaStream_t stream);
void launch_nlp_residual(const KCfg& cf, const Work& w, const double* x0, double nlp_tol, int have_mult,
                         int32_t* status, cudaStream_t stream);
cudaError_t launch_dpc_rollout_grad(long long B, int Nf, double dt, const double* weights, const double* x0,
                                    const double* u, const double* ref, double* cost, double* grad, double* xtraj,
                                    cudaStream_t stream);
cudaError_t launch_dpc_rollout_grad_f32(long long B, int Nf, double dt, const double* weights, const float* x0,
                                        const float* u, const float* ref, float* cost, float* grad, float* xtraj,
                                        cudaStream_t stream);
cudaError_t launch_project(long long B, int n_grid, const double* s_ref, const double* X_ref, const double* Y_ref,
                           const double* phi_ref, const int32_t* track, const double* pose, const double* s_guess,
                           double s_tol, double* frenet, double* ref, int32_t* flag, cudaStream_t stream);
void launch_merit_base(const KCfg& cf, const Work& w, const LineSearch& ls, const double* x0, int first,
                       cudaStream_t stream);
void launch_merit_trial(const KCfg& cf, const Work& w, const LineSearch& ls, const double* x0, int32_t* sqp_active,
                        int32_t* status, cudaStream_t stream);
void launch_sim_step(const ihm2b_config& cfg, const KCfg& cf, const Work& w, const double* x,
                     const double* u, double* xn, int32_t* model, int32_t* order_scratch, cudaStream_t stream);
void launch_sim_step_dyn10(const ihm2b_config& cfg, const KCfg& cf, const Work& w, const double* x, const double* u, double* xn,
                           cudaStream_t stream);
cudaError_t launch_sim_step_static(const ihm2b_config& cfg, const KCfg& cf, const Work& w, const double* x, const double* u,
                                   double* xn, int32_t* model, cudaStream_t stream);
int sim_static_max_batch();
void launch_closed_loop_post(int B, int step, double dt, const int32_t* track, const double* lap, const int32_t* status,
                             const double* u0, const double* xn, const double* stats, double* xs, double* xlog,
                             double* ulog, int32_t* done, int32_t* fail, int32_t* active_mask, int32_t* lap_state,
                             double* lap_time, double* cost_sum, cudaStream_t stream);
void launch_solve_finish(const KCfg& cf, const Work& w, const double* x0, const double* snap_x, const double* snap_u,
                         double* u0, int closed_loop_mask, cudaStream_t stream);
}  // namespace ihm2b

using namespace ihm2b;

static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}
#define CU(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) return fail(IHM2B_ECUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
  } while (0)

struct ihm2b_handle {
  ihm2b_config cfg;
  KCfg k;
  Work w;
  int device;
  std::vector<void*> allocs;
  double *s_ref, *kappa_ref, *wr, *wl, *params, *lap_len, *actw, *actw_sim;
  int32_t* track;
  // scratch for the host entry point and the closed loop
  double *x0_dev, *u0_dev, *xn_dev;
  int32_t* status_dev;
  int32_t* active_dev;   // SQP mode: instances still iterating
  int32_t* lap_state_dev;  // [B][2] closed loop: jumps of mod(s, lap_length) seen, step of the first one
  int32_t* sim_order_dev;  // [B + 2] plant step: instances grouped by the model they take, two counters
  int32_t* order_dev;    // launch order of the QP kernel (longest first by the previous solve's iteration counts)
  bool have_order;
  double *snap_x, *snap_u;  // iterate at the start of a solve (restored for the instances whose solve fails)
  LineSearch ls;         // allocated when globalization = MERIT_BACKTRACKING
  double *x0_pin, *u0_pin;
  int32_t* status_pin;
  bool have_tracks, have_widths, have_params;
  bool profile;
  cudaEvent_t ev[3];
  double prof_ms[2];   // accumulated linearise / QP kernel time
  int64_t prof_n;
  int64_t launches;
};

// every entry point runs on the device the handle was created on, whatever the caller's current device is
static int on_device(ihm2b_handle* h) {
  if (!h) return fail(IHM2B_EINVAL, "null handle");
  CU(cudaSetDevice(h->device));
  return IHM2B_OK;
}

namespace {

template <typename T>
int dev_alloc(ihm2b_handle* h, T** p, size_t n) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, n * sizeof(T));
  if (e != cudaSuccess) return fail(IHM2B_ENOMEM, "cudaMalloc(%zu bytes): %s", n * sizeof(T), cudaGetErrorString(e));
  e = cudaMemset(q, 0, n * sizeof(T));
  if (e != cudaSuccess) return fail(IHM2B_ECUDA, "cudaMemset: %s", cudaGetErrorString(e));
  h->allocs.push_back(q);
  *p = static_cast<T*>(q);
  return IHM2B_OK;
}

__global__ void reset_iterate_kernel(int B, int N, double dt, double T_max, double* x, double* u) {
  // controller.py:313-317: x_j = (-6 + j dt, 0, ...), u_j = (T_max, 0)
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * (N + 1)) return;
  const int b = gid / (N + 1), j = gid - b * (N + 1);
  double* xp = x + (size_t)gid * 8;
  xp[0] = -6.0 + j * dt;
  for (int i = 1; i < 8; i++) xp[i] = 0.0;
  if (j < N) { u[((size_t)b * N + j) * 2] = T_max; u[((size_t)b * N + j) * 2 + 1] = 0.0; }
}

// controller.py:388-393 (literal): x_j <- x_{j+1}, u_j <- u_{j+1} for j < N-1;
// x_{N-1} <- x_N, x_N unchanged, u_{N-1} <- 0.  One thread per (problem, component),
// marching over the stages so the in-place shift needs no second buffer.
__global__ void shift_iterate_kernel(int B, int N, double* x, double* u, const int32_t* active) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * 10) return;
  const int b = gid / 10, c = gid - b * 10;
  if (active && !active[b]) return;
  if (c < 8) {
    double* xp = x + (size_t)b * (N + 1) * 8 + c;
    for (int j = 0; j < N; j++) xp[j * 8] = xp[(j + 1) * 8];
  } else {
    double* up = u + (size_t)b * N * 2 + (c - 8);
    for (int j = 0; j < N - 1; j++) up[j * 2] = up[(j + 1) * 2];
    up[(N - 1) * 2] = 0.0;
  }
}

__global__ void fill_i32_kernel(int n, int32_t* p, int32_t v) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid < n) p[gid] = v;
}

void fill_kcfg(const ihm2b_config& c, KCfg& k) {
  memset(&k, 0, sizeof k);
  k.B = c.batch; k.N = c.N; k.n_grid = c.n_grid;
  k.integrator = c.integrator; k.qp_max_iter = c.qp_max_iter; k.qp_tolerate_maxiter = c.qp_tolerate_maxiter;
  k.dt = c.dt; k.s_target = c.s_target;
  k.n_max = c.n_max; k.v_x_max = c.v_x_max; k.T_max = c.T_max; k.delta_max = c.delta_max;
  k.g_T_max = c.g_T_max; k.g_delta_max = c.g_delta_max;
  memcpy(k.W, c.W, sizeof k.W); memcpy(k.We, c.W_e, sizeof k.We);
  k.cs = c.cost_scale_stage; k.ct = c.cost_scale_terminal;
  k.mu0 = c.qp_mu0; k.tol = c.qp_tol; k.tol_comp = c.qp_tol_comp;
  k.thr0 = c.qp_thr0 > 0.0 ? c.qp_thr0 : 0.1;
  k.split_step = c.qp_split_step; k.infeas_exit = c.qp_infeas_exit;
  k.nl = c.nonlinear_constraints ? 1 : 0;
  k.lin_fp32 = c.linearize_fp32 ? 1 : 0;
  k.half_L = c.track_half_L; k.half_W = c.track_half_W; k.a_lat_max = c.a_lat_max;
  k.T_dot_max = c.T_dot_max; k.delta_dot_max = c.delta_dot_max;
  k.slack_z = c.slack_z; k.slack_Z = c.slack_Z;
  memcpy(k.irk_A, c.irk_A, sizeof k.irk_A); memcpy(k.irk_b, c.irk_b, sizeof k.irk_b);
}

}  // namespace

extern "C" {

const char* ihm2b_last_error(void) { return g_err; }
int ihm2b_version(void) { return 100; }

int ihm2b_create(const ihm2b_config* cfg, ihm2b_handle** out) {
  if (!cfg || !out) return fail(IHM2B_EINVAL, "null argument");
  if (cfg->batch <= 0 || cfg->N < 2 || cfg->N > 512 || cfg->n_tracks <= 0 || cfg->n_grid < 2)
    return fail(IHM2B_EINVAL, "bad dimensions: batch=%d N=%d n_tracks=%d n_grid=%d", cfg->batch, cfg->N,
                cfg->n_tracks, cfg->n_grid);
  if (cfg->integrator != IHM2B_ERK4 && cfg->integrator != IHM2B_IRK_GL4)
    return fail(IHM2B_EINVAL, "unknown integrator %d", cfg->integrator);
  if (cfg->nlp_solver != IHM2B_SQP_RTI && cfg->nlp_solver != IHM2B_SQP)
    return fail(IHM2B_EINVAL, "unknown nlp_solver %d", cfg->nlp_solver);
  if (cfg->nonlinear_constraints && !(cfg->slack_Z > 0.0 && cfg->slack_z >= 0.0 && cfg->a_lat_max > 0.0 &&
                                      cfg->T_dot_max > 0.0 && cfg->delta_dot_max > 0.0))
    return fail(IHM2B_EINVAL, "nonlinear_constraints: slack_Z, a_lat_max, T_dot_max and delta_dot_max must be positive");
  int ndev = 0, dev = 0;
  CU(cudaGetDeviceCount(&ndev));
  if (ndev == 0) return fail(IHM2B_ECUDA, "no CUDA device");
  CU(cudaGetDevice(&dev));
  *out = nullptr;
  ihm2b_handle* h = new (std::nothrow) ihm2b_handle();
  if (!h) return fail(IHM2B_ENOMEM, "host allocation failed");
  h->cfg = *cfg;
  fill_kcfg(*cfg, h->k);
  h->launches = 0;
  h->have_tracks = h->have_widths = h->have_params = false;
  h->profile = false; h->prof_ms[0] = h->prof_ms[1] = 0.0; h->prof_n = 0;
  for (int i = 0; i < 3; i++) h->ev[i] = nullptr;
  h->device = dev;
  h->x0_pin = h->u0_pin = nullptr; h->status_pin = nullptr;
  const size_t B = cfg->batch, N = cfg->N;
  Work& w = h->w;
  memset(&w, 0, sizeof w);
  int rc = IHM2B_OK;
#define ALLOC(p, n) if (rc == IHM2B_OK) rc = dev_alloc(h, &(p), (n))
  ALLOC(h->s_ref, (size_t)cfg->n_tracks * cfg->n_grid);
  ALLOC(h->kappa_ref, (size_t)cfg->n_tracks * cfg->n_grid);
  ALLOC(h->params, B * NP);
  ALLOC(h->actw, B * 8);
  ALLOC(h->actw_sim, B * 8);
  ALLOC(h->track, B);
  ALLOC(h->lap_len, (size_t)cfg->n_tracks);
  ALLOC(w.x, B * (N + 1) * 8);
  ALLOC(w.u, B * N * 2);
  ALLOC(w.BA, B * N * 80);
  ALLOC(w.c, B * N * 8);
  ALLOC(w.phi, B * N * 8);
  const size_t RS = cfg->nonlinear_constraints ? 16 : 8;   // inequality row slots per stage
  ALLOC(w.LT, B * (N + 1) * 4 * RS);
  ALLOC(w.D, B * (N + 1) * 2 * RS);
  ALLOC(w.xa, B * (N + 1) * 8);
  ALLOC(w.atp, B * N * 8);
  if (cfg->nonlinear_constraints) {
    ALLOC(h->wr, (size_t)cfg->n_tracks * cfg->n_grid);
    ALLOC(h->wl, (size_t)cfg->n_tracks * cfg->n_grid);
    ALLOC(w.HC, B * (N + 1) * 16);
    ALLOC(w.SS, B * (N + 1) * 8 * 8);
    ALLOC(w.CAB, B * (N + 1) * 2 * 16);
  }
  ALLOC(w.stats, B * 8);
  ALLOC(w.last_it, B);
  ALLOC(w.status, B);
  ALLOC(w.dbg, 64 * 8);
  ALLOC(h->x0_dev, B * 8);
  ALLOC(h->u0_dev, B * 2);
  ALLOC(h->xn_dev, B * 8);
  ALLOC(h->status_dev, B);
  ALLOC(h->active_dev, B);
  ALLOC(h->order_dev, B);
  ALLOC(h->sim_order_dev, B + 2);
  ALLOC(h->lap_state_dev, B * 2);
  ALLOC(h->snap_x, B * (N + 1) * 8);
  ALLOC(h->snap_u, B * N * 2);
  memset(&h->ls, 0, sizeof h->ls);
  if (cfg->nlp_solver == IHM2B_SQP && cfg->globalization == IHM2B_MERIT_BACKTRACKING) {
    ALLOC(h->ls.base_x, B * (N + 1) * 8);
    ALLOC(h->ls.base_u, B * N * 2);
    ALLOC(h->ls.step_x, B * (N + 1) * 8);
    ALLOC(h->ls.step_u, B * N * 2);
    ALLOC(h->ls.wgt, B * (N + 1) * 24);
    ALLOC(h->ls.m0, B);
    ALLOC(h->ls.alpha, B);
    ALLOC(h->ls.mask, B);
  }
#undef ALLOC
  w.active = nullptr; w.sqp_it = -1; w.order = nullptr;
  h->have_order = false;
  if (rc == IHM2B_OK) {
    cudaError_t e = cudaMallocHost(&h->x0_pin, B * 8 * sizeof(double));
    if (e == cudaSuccess) e = cudaMallocHost(&h->u0_pin, B * 2 * sizeof(double));
    if (e == cudaSuccess) e = cudaMallocHost(&h->status_pin, B * sizeof(int32_t));
    if (e != cudaSuccess) rc = fail(IHM2B_ENOMEM, "cudaMallocHost: %s", cudaGetErrorString(e));
  }
  if (rc != IHM2B_OK) { ihm2b_destroy(h); return rc; }
  w.s_ref = h->s_ref; w.kappa_ref = h->kappa_ref; w.wr = h->wr; w.wl = h->wl; w.params = h->params; w.track = h->track;
  w.actw = h->actw; w.actw_sim = h->actw_sim;
  rc = ihm2b_reset_iterate(h, nullptr);
  if (rc != IHM2B_OK) { ihm2b_destroy(h); return rc; }
  *out = h;
  return IHM2B_OK;
}

int ihm2b_destroy(ihm2b_handle* h) {
  if (!h) return IHM2B_OK;
  for (void* p : h->allocs) cudaFree(p);
  for (int i = 0; i < 3; i++) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  if (h->x0_pin) cudaFreeHost(h->x0_pin);
  if (h->u0_pin) cudaFreeHost(h->u0_pin);
  if (h->status_pin) cudaFreeHost(h->status_pin);
  delete h;
  return IHM2B_OK;
}

int ihm2b_set_tracks(ihm2b_handle* h, const double* s_ref, const double* kappa_ref, const double* rw,
                     const double* lw) {
  if (!h || !s_ref || !kappa_ref) return fail(IHM2B_EINVAL, "null argument");
  if (h->cfg.nonlinear_constraints && (!rw || !lw))
    return fail(IHM2B_EINVAL, "nonlinear_constraints: the track constraint needs right_width and left_width");
  const size_t n = (size_t)h->cfg.n_tracks * h->cfg.n_grid;
  for (int t = 0; t < h->cfg.n_tracks; t++)
    for (int i = 1; i < h->cfg.n_grid; i++)
      if (!(s_ref[(size_t)t * h->cfg.n_grid + i] > s_ref[(size_t)t * h->cfg.n_grid + i - 1]))
        return fail(IHM2B_EINVAL, "s_ref of track %d is not strictly increasing at %d", t, i);
  CU(cudaMemcpy(h->s_ref, s_ref, n * sizeof(double), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(h->kappa_ref, kappa_ref, n * sizeof(double), cudaMemcpyHostToDevice));
  if (h->cfg.nonlinear_constraints) {
    CU(cudaMemcpy(h->wr, rw, n * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->wl, lw, n * sizeof(double), cudaMemcpyHostToDevice));
    h->have_widths = true;
  }
  h->have_tracks = true;
  return IHM2B_OK;
}

int ihm2b_set_track_index(ihm2b_handle* h, const int32_t* idx) {
  if (!h || !idx) return fail(IHM2B_EINVAL, "null argument");
  for (int i = 0; i < h->cfg.batch; i++)
    if (idx[i] < 0 || idx[i] >= h->cfg.n_tracks) return fail(IHM2B_EINVAL, "track_index[%d] = %d out of range", i, idx[i]);
  CU(cudaMemcpy(h->track, idx, (size_t)h->cfg.batch * sizeof(int32_t), cudaMemcpyHostToDevice));
  return IHM2B_OK;
}

// w = (I + z A)^-1 1 for a 4-stage tableau A (row-major): the collocation solution of the
// first-order lag x' = (u - x) / t with z = h / t is K = w (u - x) / t.
static bool lag_weights(const double* A, double z, double* w) {
  double M[4][5];
  for (int i = 0; i < 4; i++) {
    for (int j = 0; j < 4; j++) M[i][j] = (i == j ? 1.0 : 0.0) + z * A[i * 4 + j];
    M[i][4] = 1.0;
  }
  for (int c = 0; c < 4; c++) {
    int pr = c;
    for (int r = c + 1; r < 4; r++) if (fabs(M[r][c]) > fabs(M[pr][c])) pr = r;
    if (!(fabs(M[pr][c]) > 0.0)) return false;
    if (pr != c) for (int j = 0; j < 5; j++) { double t = M[c][j]; M[c][j] = M[pr][j]; M[pr][j] = t; }
    for (int r = c + 1; r < 4; r++) {
      const double l = M[r][c] / M[c][c];
      for (int j = c; j < 5; j++) M[r][j] -= l * M[c][j];
    }
  }
  for (int r = 3; r >= 0; r--) {
    double v = M[r][4];
    for (int c = r + 1; c < 4; c++) v -= M[r][c] * w[c];
    w[r] = v / M[r][r];
  }
  return true;
}

int ihm2b_set_params(ihm2b_handle* h, const double* params) {
  if (!h || !params) return fail(IHM2B_EINVAL, "null argument");
  const size_t B = h->cfg.batch;
  std::vector<double> wc(B * 8), ws(B * 8);
  const double hs = h->cfg.dt / (h->cfg.sim_steps > 0 ? h->cfg.sim_steps : 1);
  for (size_t i = 0; i < B; i++) {
    const double tT = params[i * NP + IHM2B_P_TT], tD = params[i * NP + IHM2B_P_TD];
    if (!(tT > 0.0) || !(tD > 0.0) || !(params[i * NP + IHM2B_P_M] > 0.0))
      return fail(IHM2B_EINVAL, "params[%zu]: t_T, t_delta and m must be positive", i);
    if (!lag_weights(h->cfg.irk_A, h->cfg.dt / tT, &wc[i * 8]) || !lag_weights(h->cfg.irk_A, h->cfg.dt / tD, &wc[i * 8 + 4]) ||
        !lag_weights(h->cfg.sim_A, hs / tT, &ws[i * 8]) || !lag_weights(h->cfg.sim_A, hs / tD, &ws[i * 8 + 4]))
      return fail(IHM2B_EINVAL, "params[%zu]: singular collocation system for the actuator lags", i);
  }
  CU(cudaMemcpy(h->actw, wc.data(), B * 8 * sizeof(double), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(h->actw_sim, ws.data(), B * 8 * sizeof(double), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(h->params, params, (size_t)h->cfg.batch * NP * sizeof(double), cudaMemcpyHostToDevice));
  h->have_params = true;
  return IHM2B_OK;
}

int ihm2b_set_iterate_dev(ihm2b_handle* h, const double* x, const double* u, void* stream) {
  { const int rc_ = on_device(h); if (rc_) return rc_; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t B = h->cfg.batch, N = h->cfg.N;
  if (x) CU(cudaMemcpyAsync(h->w.x, x, B * (N + 1) * 8 * sizeof(double), cudaMemcpyDeviceToDevice, s));
  if (u) CU(cudaMemcpyAsync(h->w.u, u, B * N * 2 * sizeof(double), cudaMemcpyDeviceToDevice, s));
  return IHM2B_OK;
}

int ihm2b_get_iterate_dev(ihm2b_handle* h, double* x, double* u, void* stream) {
  { const int rc_ = on_device(h); if (rc_) return rc_; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t B = h->cfg.batch, N = h->cfg.N;
  if (x) CU(cudaMemcpyAsync(x, h->w.x, B * (N + 1) * 8 * sizeof(double), cudaMemcpyDeviceToDevice, s));
  if (u) CU(cudaMemcpyAsync(u, h->w.u, B * N * 2 * sizeof(double), cudaMemcpyDeviceToDevice, s));
  return IHM2B_OK;
}

int ihm2b_get_iterate_host(ihm2b_handle* h, double* x, double* u) {
  { const int rc_ = on_device(h); if (rc_) return rc_; }
  const size_t B = h->cfg.batch, N = h->cfg.N;
  cudaStream_t s = nullptr;
  if (x) CU(cudaMemcpyAsync(x, h->w.x, B * (N + 1) * 8 * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (u) CU(cudaMemcpyAsync(u, h->w.u, B * N * 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  return IHM2B_OK;
}

int ihm2b_reset_iterate(ihm2b_handle* h, void* stream) {
  { const int rc_ = on_device(h); if (rc_) return rc_; }
  const int total = h->cfg.batch * (h->cfg.N + 1);
  reset_iterate_kernel<<<(total + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      h->cfg.batch, h->cfg.N, h->cfg.dt, h->cfg.T_max, h->w.x, h->w.u);
  h->launches++;
  CU(cudaGetLastError());
  return IHM2B_OK;
}

int ihm2b_shift_iterate(ihm2b_handle* h, void* stream) {
  { const int rc_ = on_device(h); if (rc_) return rc_; }
  const int total = h->cfg.batch * 10;
  shift_iterate_kernel<<<(total + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      h->cfg.batch, h->cfg.N, h->w.x, h->w.u, h->w.active);
  h->launches++;
  CU(cudaGetLastError());
  return IHM2B_OK;
}

static int check_ready(ihm2b_handle* h) {
  if (!h) return fail(IHM2B_EINVAL, "null handle");
  CU(cudaSetDevice(h->device));
  if (!h->have_tracks) return fail(IHM2B_EINVAL, "ihm2b_set_tracks has not been called");
  if (h->cfg.nonlinear_constraints && !h->have_widths) return fail(IHM2B_EINVAL, "nonlinear_constraints: no track widths set");
  if (!h->have_params) return fail(IHM2B_EINVAL, "ihm2b_set_params has not been called");
  return IHM2B_OK;
}

int ihm2b_linearize_dev(ihm2b_handle* h, double* A, double* Bm, double* phi, void* stream);

int ihm2b_solve_dev(ihm2b_handle* h, const double* x0, double* u0, int32_t* status, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!x0) return fail(IHM2B_EINVAL, "null x0");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  {
    const size_t B = h->cfg.batch, N = h->cfg.N;
    CU(cudaMemcpyAsync(h->snap_x, h->w.x, B * (N + 1) * 8 * sizeof(double), cudaMemcpyDeviceToDevice, s));
    CU(cudaMemcpyAsync(h->snap_u, h->w.u, B * N * 2 * sizeof(double), cudaMemcpyDeviceToDevice, s));
  }
  if (h->cfg.nlp_solver == IHM2B_SQP) {
    // acados ocp_nlp_sqp restated (configured by new_python/controller.py:301-302):
    // repeat { linearise; KKT residual test with the multipliers of the last QP; QP; full step }.
    // Instances that converged (status 0) or whose QP failed (status 1 / 4) drop out through
    // the `active` mask; the others end with status 2 (ACADOS_MAXITER), which the reference
    // tolerates (mpc_control_node.cpp:190-197).  Step length: fixed 1.0, or (controller.py:308) backtracking on the
    // L1 merit function: the trial iterate is linearised, its merit compared with the base iterate's, and the
    // instances still searching shorten their step; the linearisation of the accepted trial point is the one
    // the next SQP iteration starts from.
    const bool line_search = h->cfg.globalization == IHM2B_MERIT_BACKTRACKING;
    if (h->cfg.globalization != IHM2B_FIXED_STEP && !line_search)
      return fail(IHM2B_EINVAL, "unknown globalization %d", h->cfg.globalization);
    if (h->cfg.nlp_max_iter < 1) return fail(IHM2B_EINVAL, "nlp_max_iter %d < 1", h->cfg.nlp_max_iter);
    fill_i32_kernel<<<(h->cfg.batch + 255) / 256, 256, 0, s>>>(h->cfg.batch, h->active_dev, 1);
    fill_i32_kernel<<<(h->cfg.batch + 255) / 256, 256, 0, s>>>(h->cfg.batch, h->w.status, IHM2B_MAXITER);
    if (status) fill_i32_kernel<<<(h->cfg.batch + 255) / 256, 256, 0, s>>>(h->cfg.batch, status, IHM2B_MAXITER);
    h->launches += status ? 3 : 2;
    Work w = h->w;
    w.active = h->active_dev;
    w.order = nullptr;
    for (int it = 0; it < h->cfg.nlp_max_iter; it++) {
      w.sqp_it = it;
      if (!(line_search && it > 0)) h->launches += launch_linearize(h->k, w, s);
      launch_nlp_residual(h->k, w, x0, h->cfg.nlp_tol, it > 0, status, s);
      if (line_search) {
        const size_t B = h->cfg.batch, N = h->cfg.N;
        CU(cudaMemcpyAsync(h->ls.base_x, w.x, B * (N + 1) * 8 * sizeof(double), cudaMemcpyDeviceToDevice, s));
        CU(cudaMemcpyAsync(h->ls.base_u, w.u, B * N * 2 * sizeof(double), cudaMemcpyDeviceToDevice, s));
      }
      if (launch_qp_rti(h->k, w, x0, u0, status, s) != 0)
        return fail(IHM2B_ECUDA, "QP kernel needs %d bytes of shared memory per problem (N = %d too large)",
                    qp_smem_bytes(h->cfg.N, h->k.nl), h->cfg.N);
      h->launches += 2;
      if (line_search) {
        launch_merit_base(h->k, w, h->ls, x0, it == 0, s);
        h->launches++;
        Work wt = w;
        wt.active = h->ls.mask;
        // alpha = 1, 0.7, ..., 0.7^8 = 0.058, then alpha_min = 0.05: ten trial points at most
        for (int j = 0; j < 10; j++) {
          h->launches += launch_linearize(h->k, wt, s);
          launch_merit_trial(h->k, w, h->ls, x0, h->active_dev, status, s);
          h->launches++;
        }
      }
      if (h->cfg.schedule_by_history && it + 1 < h->cfg.nlp_max_iter) {
        // the QP of the next SQP iteration needs about as many IPM iterations as this one: longest first
        launch_build_order(h->cfg.batch, h->w.last_it, h->order_dev, s);
        w.order = h->order_dev;
        h->launches++;
      }
    }
    // all instances, at the iterate and the h linearisation the solve ends with
    launch_solve_finish(h->k, w, x0, h->snap_x, h->snap_u, u0, 0, s);
    h->launches++;
    CU(cudaGetLastError());
    return IHM2B_OK;
  }
  // RTI preparation phase: linearise at the (shifted) iterate; the measured state only
  // enters the feedback phase, as the fixed x_0 of the QP.
  if (h->profile) cudaEventRecord(h->ev[0], s);
  h->launches += launch_linearize(h->k, h->w, s);
  if (h->profile) cudaEventRecord(h->ev[1], s);
  h->w.order = (h->have_order && h->cfg.schedule_by_history) ? h->order_dev : nullptr;
  if (launch_qp_rti(h->k, h->w, x0, u0, status, s) != 0)
    return fail(IHM2B_ECUDA, "QP kernel needs %d bytes of shared memory per problem (N = %d too large)",
                qp_smem_bytes(h->cfg.N, h->k.nl), h->cfg.N);
  if (h->profile) cudaEventRecord(h->ev[2], s);
  launch_solve_finish(h->k, h->w, x0, h->snap_x, h->snap_u, nullptr, 1, s);
  h->launches++;
  if (h->cfg.schedule_by_history) {
    // next solve of this handle (closed loop: the next control step) launches its long problems first
    launch_build_order(h->cfg.batch, h->w.last_it, h->order_dev, s);
    h->have_order = true;
    h->launches++;
  }
  if (h->profile) {
    cudaEventSynchronize(h->ev[2]);
    float a = 0.f, b = 0.f;
    cudaEventElapsedTime(&a, h->ev[0], h->ev[1]);
    cudaEventElapsedTime(&b, h->ev[1], h->ev[2]);
    h->prof_ms[0] += a; h->prof_ms[1] += b; h->prof_n++;
  }
  h->launches += 1;
  CU(cudaGetLastError());
  return IHM2B_OK;
}

int ihm2b_compute_control_host(ihm2b_handle* h, const double* x0, double* u0, int32_t* status, int do_shift) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!x0 || !u0) return fail(IHM2B_EINVAL, "null argument");
  const size_t B = h->cfg.batch;
  cudaStream_t s = nullptr;
  memcpy(h->x0_pin, x0, B * 8 * sizeof(double));
  CU(cudaMemcpyAsync(h->x0_dev, h->x0_pin, B * 8 * sizeof(double), cudaMemcpyHostToDevice, s));
  if (do_shift) { rc = ihm2b_shift_iterate(h, s); if (rc) return rc; }
  rc = ihm2b_solve_dev(h, h->x0_dev, h->u0_dev, h->status_dev, s);
  if (rc) return rc;
  CU(cudaMemcpyAsync(h->u0_pin, h->u0_dev, B * 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
  CU(cudaMemcpyAsync(h->status_pin, h->status_dev, B * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  memcpy(u0, h->u0_pin, B * 2 * sizeof(double));
  if (status) memcpy(status, h->status_pin, B * sizeof(int32_t));
  return IHM2B_OK;
}

namespace {
// [B][N][8][10] (B | A) -> A [B][N][8][8], Bm [B][N][8][2]
__global__ void split_ba_kernel(size_t rows, const double* BA, double* A, double* Bm) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= rows) return;
  const double* r = BA + gid * 10;
  if (Bm) { Bm[gid * 2] = r[0]; Bm[gid * 2 + 1] = r[1]; }
  if (A) for (int j = 0; j < 8; j++) A[gid * 8 + j] = r[2 + j];
}
}  // namespace

int ihm2b_linearize_dev(ihm2b_handle* h, double* A, double* Bm, double* phi, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  h->launches += launch_linearize(h->k, h->w, s);
  const size_t rows = (size_t)h->cfg.batch * h->cfg.N * 8;
  if (A || Bm) {
    split_ba_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(rows, h->w.BA, A, Bm);
    h->launches++;
  }
  if (phi) CU(cudaMemcpyAsync(phi, h->w.phi, rows * sizeof(double), cudaMemcpyDeviceToDevice, s));
  CU(cudaGetLastError());
  return IHM2B_OK;
}

int ihm2b_get_stats_dev(ihm2b_handle* h, double* stats, void* stream) {
  if (!h || !stats) return fail(IHM2B_EINVAL, "null argument");
  CU(cudaMemcpyAsync(stats, h->w.stats, (size_t)h->cfg.batch * 8 * sizeof(double), cudaMemcpyDeviceToDevice,
                     static_cast<cudaStream_t>(stream)));
  return IHM2B_OK;
}

int ihm2b_get_multipliers_dev(ihm2b_handle* h, double* pi, double* lamL, double* lamU, void* stream) {
  { const int rc_ = on_device(h); if (rc_) return rc_; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t B = h->cfg.batch, N1 = h->cfg.N + 1, RS = h->k.nl ? 16 : 8;
  if (pi) CU(cudaMemcpyAsync(pi, h->w.xa, B * N1 * 8 * sizeof(double), cudaMemcpyDeviceToDevice, s));
  const size_t row = N1 * RS * sizeof(double);   // one problem's block of one quantity; LT holds 4 of them per problem
  if (lamL) CU(cudaMemcpy2DAsync(lamL, row, h->w.LT, 4 * row, row, B, cudaMemcpyDeviceToDevice, s));
  if (lamU) CU(cudaMemcpy2DAsync(lamU, row, h->w.LT + N1 * RS, 4 * row, row, B, cudaMemcpyDeviceToDevice, s));
  return IHM2B_OK;
}

int ihm2b_get_slacks_dev(ihm2b_handle* h, double* sL, double* sU, double* lamSL, double* lamSU, void* stream) {
  { const int rc_ = on_device(h); if (rc_) return rc_; }
  if (!h->k.nl) return fail(IHM2B_EINVAL, "no soft rows without nonlinear_constraints");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t B = h->cfg.batch, N1 = h->cfg.N + 1;
  const size_t row = N1 * 8 * sizeof(double);
  double* out[4] = {sL, sU, lamSL, lamSU};
  for (int i = 0; i < 4; i++)
    if (out[i]) CU(cudaMemcpy2DAsync(out[i], row, h->w.SS + i * N1 * 8, 8 * row, row, B, cudaMemcpyDeviceToDevice, s));
  return IHM2B_OK;
}

int ihm2b_get_h_rows_dev(ihm2b_handle* h, double* hc, void* stream) {
  if (!h || !hc) return fail(IHM2B_EINVAL, "null argument");
  if (!h->k.nl) return fail(IHM2B_EINVAL, "no h rows without nonlinear_constraints");
  CU(cudaMemcpyAsync(hc, h->w.HC, (size_t)h->cfg.batch * (h->cfg.N + 1) * 16 * sizeof(double), cudaMemcpyDeviceToDevice,
                     static_cast<cudaStream_t>(stream)));
  return IHM2B_OK;
}

int ihm2b_sim_step_dev(ihm2b_handle* h, const double* x, const double* u, double* xn, int32_t* model, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!x || !u || !xn) return fail(IHM2B_EINVAL, "null argument");
  if (h->cfg.sim_model < IHM2B_SIM_FKIN6 || h->cfg.sim_model > IHM2B_SIM_KIN6C_DYN6C) return fail(IHM2B_EINVAL, "unknown sim_model %d", h->cfg.sim_model);
  launch_sim_step(h->cfg, h->k, h->w, x, u, xn, model, h->sim_order_dev, static_cast<cudaStream_t>(stream));
  h->launches++;
  CU(cudaGetLastError());
  return IHM2B_OK;
}

// The plant step with its scratch (launch order of the model partition) in static device memory: allocates nothing.
int ihm2b_sim_step_static_dev(ihm2b_handle* h, const double* x, const double* u, double* xn, int32_t* model, void* stream) {
  int rc = check_ready(h);   // selects the handle's device
  if (rc) return rc;
  if (!x || !u || !xn) return fail(IHM2B_EINVAL, "null argument");
  if (h->cfg.sim_model < IHM2B_SIM_FKIN6 || h->cfg.sim_model > IHM2B_SIM_KIN6C_DYN6C) return fail(IHM2B_EINVAL, "unknown sim_model %d", h->cfg.sim_model);
  if (h->cfg.batch > sim_static_max_batch())
    return fail(IHM2B_EINVAL, "batch %d exceeds the %d instances the static scratch of ihm2b_sim_step_static_dev holds; use ihm2b_sim_step_dev",
                (int)h->cfg.batch, sim_static_max_batch());
  const cudaError_t e = launch_sim_step_static(h->cfg, h->k, h->w, x, u, xn, model, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail(IHM2B_ECUDA, "ihm2b_sim_step_static_dev: %s", cudaGetErrorString(e));
  h->launches++;
  return IHM2B_OK;
}

int ihm2b_sim_step_dyn10_dev(ihm2b_handle* h, const double* x, const double* u, double* xn, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!x || !u || !xn) return fail(IHM2B_EINVAL, "null argument");
  if (h->cfg.sim_steps < 1) return fail(IHM2B_EINVAL, "sim_steps %d < 1", h->cfg.sim_steps);
  launch_sim_step_dyn10(h->cfg, h->k, h->w, x, u, xn, static_cast<cudaStream_t>(stream));
  h->launches++;
  CU(cudaGetLastError());
  return IHM2B_OK;
}

int ihm2b_closed_loop_dev(ihm2b_handle* h, int32_t n_steps, double* x, const double* lap_length, double* xlog,
                          double* ulog, int32_t* done, int32_t* failp, double* lap_time, double* cost_sum, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (!x || !lap_length || !done || !failp || n_steps < 0) return fail(IHM2B_EINVAL, "bad argument");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t B = h->cfg.batch;
  CU(cudaMemcpyAsync(h->lap_len, lap_length, (size_t)h->cfg.n_tracks * sizeof(double), cudaMemcpyHostToDevice, s));
  CU(cudaMemsetAsync(done, 0xFF, B * sizeof(int32_t), s));   // -1
  CU(cudaMemsetAsync(failp, 0xFF, B * sizeof(int32_t), s));
  CU(cudaMemsetAsync(h->lap_state_dev, 0, 2 * B * sizeof(int32_t), s));
  if (lap_time) CU(cudaMemsetAsync(lap_time, 0xFF, B * sizeof(double), s));   // NaN until two jumps have been seen
  if (cost_sum) CU(cudaMemsetAsync(cost_sum, 0, B * sizeof(double), s));
  if (xlog) CU(cudaMemcpyAsync(xlog, x, B * 8 * sizeof(double), cudaMemcpyDeviceToDevice, s));
  // instances that finished their lap or stopped on a solver failure cost nothing in later steps: every
  // kernel of the step honours the mask (RTI mode only; the SQP loop owns the mask during a solve)
  const bool masked = (h->cfg.nlp_solver == IHM2B_SQP_RTI);
  if (masked) {
    fill_i32_kernel<<<(unsigned)((B + 255) / 256), 256, 0, s>>>((int)B, h->active_dev, 1);
    h->w.active = h->active_dev;
    h->launches++;
  }
  for (int step = 0; step < n_steps; step++) {
    rc = ihm2b_shift_iterate(h, stream);
    if (!rc) rc = ihm2b_solve_dev(h, x, h->u0_dev, h->status_dev, stream);
    if (rc) { h->w.active = nullptr; return rc; }
    launch_sim_step(h->cfg, h->k, h->w, x, h->u0_dev, h->xn_dev, nullptr, h->sim_order_dev, s);
    launch_closed_loop_post((int)B, step, h->cfg.dt, h->track, h->lap_len, h->status_dev, h->u0_dev, h->xn_dev, h->w.stats, x,
                            xlog, ulog, done, failp, masked ? h->active_dev : nullptr, h->lap_state_dev, lap_time, cost_sum, s);
    h->launches += 2;
  }
  h->w.active = nullptr;
  CU(cudaGetLastError());
  return IHM2B_OK;
}

/* per-iteration IPM trace of problem 0 */
int ihm2b_debug_trace(ihm2b_handle* h, double* host64x8) {
  if (!h || !host64x8) return IHM2B_EINVAL;
  return cudaMemcpy(host64x8, h->w.dbg, 64 * 8 * sizeof(double), cudaMemcpyDeviceToHost) == cudaSuccess ? IHM2B_OK : IHM2B_ECUDA;
}

/* measurement aids ----------------------------------------------------------------------
 * per-kernel device times of ihm2b_solve_dev: when enabled every solve synchronises on its
 * own events; ihm2b_profile_read returns the mean linearise / QP kernel time in ms. */
int ihm2b_profile_enable(ihm2b_handle* h, int on) {
  if (!h) return IHM2B_EINVAL;
  if (on && !h->ev[0]) for (int i = 0; i < 3; i++) if (cudaEventCreate(&h->ev[i]) != cudaSuccess) return IHM2B_ECUDA;
  h->profile = on != 0; h->prof_ms[0] = h->prof_ms[1] = 0.0; h->prof_n = 0;
  return IHM2B_OK;
}
int ihm2b_profile_read(ihm2b_handle* h, double* ms2) {
  if (!h || !ms2) return IHM2B_EINVAL;
  const double n = h->prof_n > 0 ? (double)h->prof_n : 1.0;
  ms2[0] = h->prof_ms[0] / n; ms2[1] = h->prof_ms[1] / n;
  return IHM2B_OK;
}

namespace {
__global__ void dfma_peak_kernel(double* out, int iters) {
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double m = 1.0000001, c = 1e-9;
  for (int i = 0; i < iters; i++) {
    a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
    a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
  }
  if (a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 == 12345.678) out[0] = a0;
}
// the same with mma.sync.m8n8k4.f64 (the QP kernel's sweeps run on this pipe): 8 independent accumulator pairs
__global__ void dmma_peak_kernel(double* out, int iters) {
  const double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  double c[8][2] = {};
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[j][0]), "+d"(c[j][1]) : "d"(a), "d"(b));
  }
  double s = 0.0;
  for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1];
  if (s == 12345.678) out[0] = s;
}
}  // namespace
/* FP64 throughput of the current device in TFLOP/s, the roofline denominator that MEASURED_PEAKS.json
 * does not carry: the larger of a DFMA chain (CUDA cores) and a DMMA chain (mma.m8n8k4.f64). */
int ihm2b_measure_fp64_tflops(double* tflops) {
  if (!tflops) return IHM2B_EINVAL;
  cudaDeviceProp prop;
  int dev = 0;
  CU(cudaGetDevice(&dev));
  CU(cudaGetDeviceProperties(&prop, dev));
  double* d = nullptr;
  CU(cudaMalloc(&d, 8));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 15;
  double best = 0.0;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    dfma_peak_kernel<<<blocks, threads>>>(d, iters);
    cudaEventRecord(e1);
    CU(cudaEventSynchronize(e1));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double tf = 2.0 * 8.0 * iters * (double)blocks * threads / (ms * 1e-3) * 1e-12;
    if (rep > 0 && tf > best) best = tf;
  }
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    dmma_peak_kernel<<<blocks, threads>>>(d, iters / 8);
    cudaEventRecord(e1);
    CU(cudaEventSynchronize(e1));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double tf = 2.0 * 256.0 * 8.0 * (iters / 8) * (double)blocks * (threads / 32) / (ms * 1e-3) * 1e-12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  *tflops = best;
  return IHM2B_OK;
}

int ihm2b_dpc_rollout_grad_dev(int64_t batch, int32_t Nf, double dt, const double* x0, const double* u,
                               const double* ref, const double* weights, double* cost, double* grad, double* xtraj,
                               void* stream) {
  if (!x0 || !u || !ref || !weights || !cost || !grad) return fail(IHM2B_EINVAL, "null argument");
  if (batch < 0 || !(dt > 0.0)) return fail(IHM2B_EINVAL, "bad batch %lld or dt %g", (long long)batch, dt);
  if (Nf != 40 && Nf != 80) return fail(IHM2B_EINVAL, "Nf = %d: the rollout kernel is built for Nf = 40 and 80", Nf);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(IHM2B_ECUDA, "no CUDA device");
  if (batch == 0) return IHM2B_OK;
  CU(launch_dpc_rollout_grad(batch, Nf, dt, weights, x0, u, ref, cost, grad, xtraj, static_cast<cudaStream_t>(stream)));
  return IHM2B_OK;
}

int ihm2b_dpc_rollout_grad_f32_dev(int64_t batch, int32_t Nf, double dt, const float* x0, const float* u, const float* ref,
                                   const double* weights, float* cost, float* grad, float* xtraj, void* stream) {
  if (!x0 || !u || !ref || !weights || !cost || !grad) return fail(IHM2B_EINVAL, "null argument");
  if (batch < 0 || !(dt > 0.0)) return fail(IHM2B_EINVAL, "bad batch %lld or dt %g", (long long)batch, dt);
  if (Nf != 40 && Nf != 80) return fail(IHM2B_EINVAL, "Nf = %d: the rollout kernel is built for Nf = 40 and 80", Nf);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(IHM2B_ECUDA, "no CUDA device");
  if (batch == 0) return IHM2B_OK;
  CU(launch_dpc_rollout_grad_f32(batch, Nf, dt, weights, x0, u, ref, cost, grad, xtraj, static_cast<cudaStream_t>(stream)));
  return IHM2B_OK;
}

int ihm2b_project_dev(int64_t batch, int32_t n_grid, const double* s_ref, const double* X_ref, const double* Y_ref,
                      const double* phi_ref, const int32_t* track, const double* pose, const double* s_guess,
                      double s_tol, double* frenet, double* ref, int32_t* flag, void* stream) {
  if (!s_ref || !X_ref || !Y_ref || !phi_ref || !pose || !s_guess || !frenet) return fail(IHM2B_EINVAL, "null argument");
  if (batch < 0 || n_grid < 3 || !(s_tol > 0.0)) return fail(IHM2B_EINVAL, "bad batch %lld, n_grid %d or s_tol %g", (long long)batch, n_grid, s_tol);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(IHM2B_ECUDA, "no CUDA device");
  if (batch == 0) return IHM2B_OK;
  CU(launch_project(batch, n_grid, s_ref, X_ref, Y_ref, phi_ref, track, pose, s_guess, s_tol, frenet, ref, flag,
                    static_cast<cudaStream_t>(stream)));
  return IHM2B_OK;
}

int64_t ihm2b_launch_count(const ihm2b_handle* h) { return h ? h->launches : 0; }

}  // extern "C"
