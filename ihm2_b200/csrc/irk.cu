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
// One thread per (problem, shooting interval); the small dense factors live in local memory.
#include "model.cuh"

namespace ihm2b {

namespace {

constexpr int NS = 4;
constexpr int MAX_NEWTON = 12;

// in-place LU without pivoting (the matrices are I - h A (x) J, close to the identity)
template <int n>
__device__ __forceinline__ void lu_factor(double* M) {
#pragma unroll 1
  for (int c = 0; c < n; c++) {
    const double inv = 1.0 / M[c * n + c];
    M[c * n + c] = inv;  // store the reciprocal pivot
    for (int r = c + 1; r < n; r++) {
      const double l = M[r * n + c] * inv;
      M[r * n + c] = l;
      for (int j = c + 1; j < n; j++) M[r * n + j] = fma(-l, M[c * n + j], M[r * n + j]);
    }
  }
}
template <int n>
__device__ __forceinline__ void lu_solve(const double* M, double* b) {
#pragma unroll 1
  for (int r = 1; r < n; r++) {
    double v = b[r];
    for (int c = 0; c < r; c++) v = fma(-M[r * n + c], b[c], v);
    b[r] = v;
  }
#pragma unroll 1
  for (int r = n - 1; r >= 0; r--) {
    double v = b[r];
    for (int c = r + 1; c < n; c++) v = fma(-M[r * n + c], b[c], v);
    b[r] = v * M[r * n + r];
  }
}

// quantities of a collocation point that depend only on (T_i, delta_i, d delta_i / dt)
struct VPoint {
  double cb, sb, bdot, cdb, Fm;
};

}  // namespace

// x[8], u -> xn[8], S[8][10].  wT / wD: (I + (h / t) A)^-1 1 for the two actuator lags.
__device__ void irk_gl4_sens(const double* __restrict__ x, double u0, double u1,
                             const double* __restrict__ p, const TrackView& tr, double h,
                             const double* __restrict__ A4, const double* __restrict__ b4,
                             const double* __restrict__ wT, const double* __restrict__ wD,
                             double* __restrict__ xn, double (*S)[NZ], bool& ok) {
  const double m = p[IHM2B_P_M], lR = p[IHM2B_P_LR], lF = p[IHM2B_P_LF];
  const double C = lR / (lR + lF), im = 1.0 / m;
  const double itT = 1.0 / p[IHM2B_P_TT], itD = 1.0 / p[IHM2B_P_TD];
  double K[NS][NX];
  // ---- actuators: K = w (u - x) / t
  double Ti[NS], Di[NS];
  for (int i = 0; i < NS; i++) { K[i][6] = wT[i] * (u0 - x[6]) * itT; K[i][7] = wD[i] * (u1 - x[7]) * itD; }
  for (int i = 0; i < NS; i++) {
    double a = x[6], d = x[7];
    for (int j = 0; j < NS; j++) { a = fma(h * A4[i * NS + j], K[j][6], a); d = fma(h * A4[i * NS + j], K[j][7], d); }
    Ti[i] = a; Di[i] = d;
  }
  // ---- (v_x, v_y): Newton on 8 unknowns
  VPoint vp[NS];
  for (int i = 0; i < NS; i++) {
    double sd, cd;
    sincos(Di[i], &sd, &cd);
    const double td = sd / cd, Ctd = C * td, den = 1.0 + Ctd * Ctd;
    vp[i].cb = rsqrt(den); vp[i].sb = Ctd * vp[i].cb;
    vp[i].bdot = C * (1.0 + td * td) / den * K[i][7];
    vp[i].cdb = cd * vp[i].cb + sd * vp[i].sb;
    vp[i].Fm = p[IHM2B_P_CM0] * Ti[i];
  }
  double Mv[8 * 8], Rv[8];
  auto veval = [&](int i, double vx, double vy, double& f3, double& f4, double& dvd) {
    const double th = tanh(10.0 * vx);
    const double poly = p[IHM2B_P_CR0] + p[IHM2B_P_CR1] * vx + p[IHM2B_P_CR2] * vx * vx;
    const double FRx = 0.5 * vp[i].Fm - poly * th;
    const double vdot = (FRx * vp[i].cb + 0.5 * vp[i].Fm * vp[i].cdb) * im;
    f3 = vdot * vp[i].cb - vp[i].bdot * vy;
    f4 = vdot * vp[i].sb + vp[i].bdot * vx;
    dvd = (-(p[IHM2B_P_CR1] + 2.0 * p[IHM2B_P_CR2] * vx) * th - poly * 10.0 * (1.0 - th * th)) * vp[i].cb * im;
  };
  for (int i = 0; i < NS; i++) { double d; veval(i, x[3], x[4], K[i][3], K[i][4], d); }
  ok = true;
#pragma unroll 1
  for (int it = 0;; it++) {
    for (int i = 0; i < NS; i++) {
      double vx = x[3], vy = x[4];
      for (int j = 0; j < NS; j++) { vx = fma(h * A4[i * NS + j], K[j][3], vx); vy = fma(h * A4[i * NS + j], K[j][4], vy); }
      double f3, f4, dvd;
      veval(i, vx, vy, f3, f4, dvd);
      Rv[2 * i] = K[i][3] - f3; Rv[2 * i + 1] = K[i][4] - f4;
      const double J33 = dvd * vp[i].cb, J34 = -vp[i].bdot, J43 = dvd * vp[i].sb + vp[i].bdot;
      for (int j = 0; j < NS; j++) {
        const double ha = h * A4[i * NS + j], e = (i == j) ? 1.0 : 0.0;
        Mv[(2 * i) * 8 + 2 * j] = e - ha * J33;      Mv[(2 * i) * 8 + 2 * j + 1] = -ha * J34;
        Mv[(2 * i + 1) * 8 + 2 * j] = -ha * J43;     Mv[(2 * i + 1) * 8 + 2 * j + 1] = e;
      }
    }
    lu_factor<8>(Mv);
    lu_solve<8>(Mv, Rv);
    double dmax = 0.0, kmax = 0.0;
    for (int i = 0; i < NS; i++) {
      K[i][3] -= Rv[2 * i]; K[i][4] -= Rv[2 * i + 1];
      dmax = fmax(dmax, fmax(fabs(Rv[2 * i]), fabs(Rv[2 * i + 1])));
      kmax = fmax(kmax, fmax(fabs(K[i][3]), fabs(K[i][4])));
    }
    if (!(dmax == dmax)) { ok = false; break; }
    if (dmax <= 1e-14 * (1.0 + kmax) || it >= MAX_NEWTON) break;
  }
  // ---- r: explicit, r' = l_R v_y' - beta'   (python/models.py:304)
  double vxi[NS], vyi[NS], ri[NS];
  for (int i = 0; i < NS; i++) K[i][5] = lR * K[i][4] - vp[i].bdot;
  for (int i = 0; i < NS; i++) {
    double a = x[3], b = x[4], c = x[5];
    for (int j = 0; j < NS; j++) {
      const double ha = h * A4[i * NS + j];
      a = fma(ha, K[j][3], a); b = fma(ha, K[j][4], b); c = fma(ha, K[j][5], c);
    }
    vxi[i] = a; vyi[i] = b; ri[i] = c;
  }
  // ---- (s, n, psi): Newton on 12 unknowns
  double Mp[12 * 12], Rp[12];
  {
    double kap, dkap, sp, cp;
    kappa_lookup(tr, x[0], kap, dkap);
    sincos(x[2], &sp, &cp);
    const double sdot = (x[3] * cp - x[4] * sp) / (1.0 + kap * x[1]);
    for (int i = 0; i < NS; i++) { K[i][0] = sdot; K[i][1] = x[3] * sp + x[4] * cp; K[i][2] = x[5] - kap * sdot; }
  }
#pragma unroll 1
  for (int it = 0;; it++) {
    for (int i = 0; i < NS; i++) {
      double s = x[0], n = x[1], psi = x[2];
      for (int j = 0; j < NS; j++) {
        const double ha = h * A4[i * NS + j];
        s = fma(ha, K[j][0], s); n = fma(ha, K[j][1], n); psi = fma(ha, K[j][2], psi);
      }
      double kap, dkap, sp, cp;
      kappa_lookup(tr, s, kap, dkap);
      sincos(psi, &sp, &cp);
      const double iD = 1.0 / (1.0 + kap * n);
      const double num = vxi[i] * cp - vyi[i] * sp;
      const double sdot = num * iD;
      Rp[3 * i] = K[i][0] - sdot;
      Rp[3 * i + 1] = K[i][1] - (vxi[i] * sp + vyi[i] * cp);
      Rp[3 * i + 2] = K[i][2] - (ri[i] - kap * sdot);
      double J[3][3];
      J[0][0] = -sdot * dkap * n * iD; J[0][1] = -sdot * kap * iD; J[0][2] = -(vxi[i] * sp + vyi[i] * cp) * iD;
      J[1][0] = 0.0; J[1][1] = 0.0; J[1][2] = num;
      J[2][0] = -dkap * sdot - kap * J[0][0]; J[2][1] = -kap * J[0][1]; J[2][2] = -kap * J[0][2];
      for (int j = 0; j < NS; j++) {
        const double ha = h * A4[i * NS + j];
        for (int a = 0; a < 3; a++)
          for (int c = 0; c < 3; c++)
            Mp[(3 * i + a) * 12 + 3 * j + c] = ((i == j && a == c) ? 1.0 : 0.0) - ha * J[a][c];
      }
    }
    lu_factor<12>(Mp);
    lu_solve<12>(Mp, Rp);
    double dmax = 0.0, kmax = 0.0;
    for (int i = 0; i < NS; i++)
      for (int a = 0; a < 3; a++) {
        K[i][a] -= Rp[3 * i + a];
        dmax = fmax(dmax, fabs(Rp[3 * i + a]));
        kmax = fmax(kmax, fabs(K[i][a]));
      }
    if (!(dmax == dmax)) { ok = false; break; }
    if (dmax <= 1e-14 * (1.0 + kmax) || it >= MAX_NEWTON) break;
  }
  for (int a = 0; a < NX; a++) {
    double v = x[a];
    for (int i = 0; i < NS; i++) v = fma(h * b4[i], K[i][a], v);
    xn[a] = v;
  }
  if (!S) return;

  // ---- sensitivities: (I - h A (x) J) dK = J_x E + J_u, block forward substitution,
  //      one column of d / d(x,u) at a time.  J at the converged collocation points.
  double J4[NS][NX][NZ];
  for (int i = 0; i < NS; i++) {
    double xs[NX], f[NX];
    for (int a = 0; a < NX; a++) {
      double v = x[a];
      for (int j = 0; j < NS; j++) v = fma(h * A4[i * NS + j], K[j][a], v);
      xs[a] = v;
    }
    fkin6<true>(xs, u0, u1, p, tr, f, J4[i]);
  }
#pragma unroll 1
  for (int col = 0; col < NZ; col++) {
    double dK[NS][NX], dX[NS][NX];  // dX excludes the identity part
    for (int i = 0; i < NS; i++) {
      dK[i][6] = (col == 6) ? -wT[i] * itT : (col == 8) ? wT[i] * itT : 0.0;
      dK[i][7] = (col == 7) ? -wD[i] * itD : (col == 9) ? wD[i] * itD : 0.0;
    }
    auto stage_sum = [&](int a) {
      for (int i = 0; i < NS; i++) {
        double v = 0.0;
        for (int j = 0; j < NS; j++) v = fma(h * A4[i * NS + j], dK[j][a], v);
        dX[i][a] = v;
      }
    };
    stage_sum(6); stage_sum(7);
    const bool vcol = (SMASK[3] >> col) & 1u;
    if (vcol) {
      for (int i = 0; i < NS; i++)
        for (int a = 0; a < 2; a++) {
          double v = ((JMASK[3 + a] >> col) & 1u) ? J4[i][3 + a][col] : 0.0;
          v = fma(J4[i][3 + a][6], dX[i][6], v);
          v = fma(J4[i][3 + a][7], dX[i][7], v);
          Rv[2 * i + a] = v;
        }
      lu_solve<8>(Mv, Rv);
      for (int i = 0; i < NS; i++) { dK[i][3] = Rv[2 * i]; dK[i][4] = Rv[2 * i + 1]; }
    } else {
      for (int i = 0; i < NS; i++) dK[i][3] = dK[i][4] = 0.0;
    }
    stage_sum(3); stage_sum(4);
    for (int i = 0; i < NS; i++) {
      double v = ((JMASK[5] >> col) & 1u) ? J4[i][5][col] : 0.0;
      v = fma(J4[i][5][3], dX[i][3], v);
      v = fma(J4[i][5][6], dX[i][6], v);
      v = fma(J4[i][5][7], dX[i][7], v);
      dK[i][5] = vcol ? v : 0.0;
    }
    stage_sum(5);
    for (int i = 0; i < NS; i++)
      for (int a = 0; a < 3; a++) {
        double v = ((JMASK[a] >> col) & 1u) ? J4[i][a][col] : 0.0;
        if ((JMASK[a] >> 3) & 1u) v = fma(J4[i][a][3], dX[i][3], v);
        if ((JMASK[a] >> 4) & 1u) v = fma(J4[i][a][4], dX[i][4], v);
        if ((JMASK[a] >> 5) & 1u) v = fma(J4[i][a][5], dX[i][5], v);
        Rp[3 * i + a] = v;
      }
    lu_solve<12>(Mp, Rp);
    for (int i = 0; i < NS; i++)
      for (int a = 0; a < 3; a++) dK[i][a] = Rp[3 * i + a];
    for (int a = 0; a < NX; a++) {
      double v = (a == col) ? 1.0 : 0.0;
      for (int i = 0; i < NS; i++) v = fma(h * b4[i], dK[i][a], v);
      S[a][col] = v;
    }
  }
}

// grid: ceil(B*N / 128), one thread per (problem, interval)
#ifndef LIN_MIN_BLOCKS
#define LIN_MIN_BLOCKS 2
#endif
__global__ void __launch_bounds__(128, LIN_MIN_BLOCKS) linearize_irk_kernel(const KCfg cf, const Work w) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= cf.B * cf.N) return;
  const int b = gid / cf.N, k = gid - b * cf.N;
  if (w.active && !w.active[b]) return;   // SQP mode: instance already converged or failed
  const int t = w.track[b];
  TrackView tr{cf.n_grid, w.s_ref + (size_t)t * cf.n_grid, w.kappa_ref + (size_t)t * cf.n_grid};
  double p[NP], x[NX], wa[8];
  for (int i = 0; i < NP; i++) p[i] = __ldg(w.params + (size_t)b * NP + i);
  for (int i = 0; i < 8; i++) wa[i] = __ldg(w.actw + (size_t)b * 8 + i);
  for (int i = 0; i < NX; i++) x[i] = w.x[((size_t)b * (cf.N + 1) + k) * NX + i];
  const double u0 = w.u[((size_t)b * cf.N + k) * NU], u1 = w.u[((size_t)b * cf.N + k) * NU + 1];
  double xn[NX], S[NX][NZ];
  bool ok;
  irk_gl4_sens(x, u0, u1, p, tr, cf.dt, cf.irk_A, cf.irk_b, wa, wa + 4, xn, S, ok);
  const size_t sk = (size_t)b * cf.N + k;
  double* ba = w.BA + sk * NX * NZ;
  for (int i = 0; i < NX; i++) {
    double ci = ok ? xn[i] - S[i][8] * u0 - S[i][9] * u1 : nan("");
    for (int j = 0; j < NX; j++) ci = fma(-S[i][j], x[j], ci);
    w.c[sk * NX + i] = ci;
    w.phi[sk * NX + i] = ok ? xn[i] : nan("");
    ba[i * NZ] = S[i][8]; ba[i * NZ + 1] = S[i][9];
    for (int j = 0; j < NX; j++) ba[i * NZ + 2 + j] = S[i][j];
  }
}

void launch_linearize_irk(const KCfg& cf, const Work& w, cudaStream_t stream) {
  const int total = cf.B * cf.N;
  linearize_irk_kernel<<<(total + 127) / 128, 128, 0, stream>>>(cf, w);
}

}  // namespace ihm2b
