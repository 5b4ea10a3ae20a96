// linearize.cu -- integrator of the controller model with forward sensitivities.
//
// Replaces acados' sim_erk inside the SQP preparation phase (configured in the reference at
// old/generate.py:24-26: ERK, 4 stages, 1 step = classic RK4) for all N shooting intervals
// of all B problems at once: one thread per (problem, interval).  The variational equations
// are propagated with the compile-time sparsity of the fkin6 Jacobian (model.cuh), column
// by column so that only the previous stage's sensitivities and the running sum are live.
#include "model.cuh"

namespace ihm2b {

__device__ __forceinline__ bool bit(unsigned m, int j) { return (m >> j) & 1u; }

// One classic RK4 step with forward sensitivities.  S[i][j] = d xn_i / d (x,u)_j.
__device__ __forceinline__ void erk4_sens(const double* __restrict__ x, double u0, double u1,
                                          const double* __restrict__ p, const TrackView& tr,
                                          double h, double* __restrict__ xn, double (*S)[NZ]) {
  double kprev[NX], acc[NX], xs[NX], k[NX];
  double J[NX][NZ];
  double dkp[NX][NZ];  // sensitivities of the previous stage derivative (SMASK entries)
#pragma unroll
  for (int i = 0; i < NX; i++) {
    acc[i] = 0.0;
    kprev[i] = 0.0;
#pragma unroll
    for (int j = 0; j < NZ; j++) { S[i][j] = 0.0; dkp[i][j] = 0.0; }
  }
#pragma unroll
  for (int st = 0; st < 4; st++) {
    const double a = (st == 0) ? 0.0 : (st == 3 ? 1.0 : 0.5);
    const double bw = (st == 0 || st == 3) ? (1.0 / 6.0) : (1.0 / 3.0);
    const double ah = a * h;
#pragma unroll
    for (int i = 0; i < NX; i++) xs[i] = x[i] + ah * kprev[i];
    fkin6<true>(xs, u0, u1, p, tr, k, J);
#pragma unroll
    for (int i = 0; i < NX; i++) { acc[i] += bw * k[i]; kprev[i] = k[i]; }
    // dk[:, j] = Jx (e_j + ah dkp[:, j]) + Ju[:, j]
#pragma unroll
    for (int j = 0; j < NZ; j++) {
      double t[NX], nk[NX];
#pragma unroll
      for (int l = 0; l < NX; l++) t[l] = (st > 0 && bit(SMASK[l], j)) ? ah * dkp[l][j] : 0.0;
#pragma unroll
      for (int i = 0; i < NX; i++) {
        if (!bit(SMASK[i], j)) continue;
        // direct term: identity part of the stage state (j < 8) or d f / d u (j >= 8)
        double v = bit(JMASK[i], j) ? J[i][j] : 0.0;
#pragma unroll
        for (int l = 0; l < NX; l++)
          if (st > 0 && bit(JMASK[i], l) && bit(SMASK[l], j)) v = fma(J[i][l], t[l], v);
        nk[i] = v;
      }
#pragma unroll
      for (int i = 0; i < NX; i++)
        if (bit(SMASK[i], j)) { dkp[i][j] = nk[i]; S[i][j] = fma(bw, nk[i], S[i][j]); }
    }
  }
#pragma unroll
  for (int i = 0; i < NX; i++) {
    xn[i] = x[i] + h * acc[i];
#pragma unroll
    for (int j = 0; j < NZ; j++) S[i][j] = ((i == j) ? 1.0 : 0.0) + (bit(SMASK[i], j) ? h * S[i][j] : 0.0);
  }
}

// grid: ceil(B*N / blockDim), one thread per (problem, interval)
__global__ void __launch_bounds__(128) linearize_erk4_kernel(const KCfg cf, const Work w) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= cf.B * cf.N) return;
  const int b = gid / cf.N, k = gid - b * cf.N;
  if (w.active && !w.active[b]) return;   // SQP mode: instance already converged or failed
  const int t = w.track[b];
  TrackView tr{cf.n_grid, w.s_ref + (size_t)t * cf.n_grid, w.kappa_ref + (size_t)t * cf.n_grid};
  double p[NP];
  {
    const double2* pp = reinterpret_cast<const double2*>(w.params + (size_t)b * NP);
#pragma unroll
    for (int i = 0; i < NP / 2; i++) { double2 v = __ldg(pp + i); p[2 * i] = v.x; p[2 * i + 1] = v.y; }
  }
  double x[NX];
  {
    const double2* xp = reinterpret_cast<const double2*>(w.x + ((size_t)b * (cf.N + 1) + k) * NX);
#pragma unroll
    for (int i = 0; i < NX / 2; i++) { double2 v = xp[i]; x[2 * i] = v.x; x[2 * i + 1] = v.y; }
  }
  const double2 uu = *reinterpret_cast<const double2*>(w.u + ((size_t)b * cf.N + k) * NU);
  double xn[NX], S[NX][NZ];
  erk4_sens(x, uu.x, uu.y, p, tr, cf.dt, xn, S);

  const size_t sk = (size_t)b * cf.N + k;
  double2* ba = reinterpret_cast<double2*>(w.BA + sk * NX * NZ);
  double c[NX];
#pragma unroll
  for (int i = 0; i < NX; i++) {
    double ci = xn[i] - S[i][8] * uu.x - S[i][9] * uu.y;
#pragma unroll
    for (int j = 0; j < NX; j++) ci = fma(-S[i][j], x[j], ci);
    c[i] = ci;
    ba[i * 5 + 0] = make_double2(S[i][8], S[i][9]);
#pragma unroll
    for (int j = 0; j < 4; j++) ba[i * 5 + 1 + j] = make_double2(S[i][2 * j], S[i][2 * j + 1]);
  }
  double2* cp = reinterpret_cast<double2*>(w.c + sk * NX);
  double2* pp = reinterpret_cast<double2*>(w.phi + sk * NX);
#pragma unroll
  for (int i = 0; i < NX / 2; i++) {
    cp[i] = make_double2(c[2 * i], c[2 * i + 1]);
    pp[i] = make_double2(xn[2 * i], xn[2 * i + 1]);
  }
}

// linearised rows of h(x, u) at the current iterate, all N + 1 stages: one thread per (problem, stage)
__global__ void __launch_bounds__(128) con_h_kernel(const KCfg cf, const Work w) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int N1 = cf.N + 1;
  if (gid >= cf.B * N1) return;
  const int b = gid / N1, k = gid - b * N1;
  if (w.active && !w.active[b]) return;
  const int t = w.track[b];
  const size_t to = (size_t)t * cf.n_grid;
  TrackView tr{cf.n_grid, w.s_ref + to, w.kappa_ref + to};
  double p[NP];
  {
    const double2* pp = reinterpret_cast<const double2*>(w.params + (size_t)b * NP);
#pragma unroll
    for (int i = 0; i < NP / 2; i++) { double2 v = __ldg(pp + i); p[2 * i] = v.x; p[2 * i + 1] = v.y; }
  }
  double x[NX];
  {
    const double2* xp = reinterpret_cast<const double2*>(w.x + (size_t)gid * NX);
#pragma unroll
    for (int i = 0; i < NX / 2; i++) { double2 v = xp[i]; x[2 * i] = v.x; x[2 * i + 1] = v.y; }
  }
  const bool term = (k == cf.N);
  double2 uu = make_double2(0.0, 0.0);
  if (!term) uu = *reinterpret_cast<const double2*>(w.u + ((size_t)b * cf.N + k) * NU);
  double out[16];
  con_h_rows(x, uu.x, uu.y, p, tr, w.wr + to, w.wl + to, cf.half_L, cf.half_W, term, out);
  double2* hp = reinterpret_cast<double2*>(w.HC + (size_t)gid * 16);
#pragma unroll
  for (int i = 0; i < 8; i++) hp[i] = make_double2(out[2 * i], out[2 * i + 1]);
}

void launch_linearize_irk(const KCfg& cf, const Work& w, cudaStream_t stream);

// returns the number of kernels launched
int launch_linearize(const KCfg& cf, const Work& w, cudaStream_t stream) {
  int n = 1;
  if (cf.nl) {
    const int total = cf.B * (cf.N + 1);
    con_h_kernel<<<(total + 127) / 128, 128, 0, stream>>>(cf, w);
    n = 2;
  }
  if (cf.integrator == IHM2B_IRK_GL4) { launch_linearize_irk(cf, w, stream); return n; }
  const int total = cf.B * cf.N;
  linearize_erk4_kernel<<<(total + 127) / 128, 128, 0, stream>>>(cf, w);
  return n;
}

}  // namespace ihm2b
