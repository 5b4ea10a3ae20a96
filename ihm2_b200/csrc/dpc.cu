// dpc.cu -- DPC rollout of the kin4 bicycle model and the gradient of its tracking cost in the controls
// (BASELINE.json configs[4]).
//
// Reference: dpc/main.py:56-84 (kin4 ODE), :87-97 (one RK4 step of dt = 1/20), :374-409 (IHM1.5 cost: tracking
// terms on stages 1..Nf-1, control terms on stages 0..Nf-1, terminal terms with the _f weights; doc/dpc.typ:44-56),
// constants dpc/main.py:23-37.  The reference leaves the policy itself undefined (DPCController is an empty class,
// dpc/main.py:617-635); what it fixes is the rollout and the loss, so this kernel returns the loss and d loss / d u
// for a batch of control sequences -- the quantity any policy trained through the rollout back-propagates.
//
// One thread per rollout.  Forward: RK4 over the horizon, the stage cost and its state gradient are formed on the
// fly; per stage the thread keeps (phi, v, dJ/dX, dJ/dY, dJ/dphi, dJ/dv) in local memory (interleaved per thread by
// the hardware, so coalesced).  Backward: the four RK4 stage points of a step are recomputed from (phi, v) -- X and Y
// never enter the right-hand side -- and the adjoint is propagated through them by hand; the X / Y adjoints are plain
// suffix sums.  Controls and gradients move as 32-byte pairs of stages so every sector touched is used in full.
#include "common.cuh"

namespace ihm2b {

namespace {
// dpc/main.py:23-29
constexpr double DPC_M = 230.0, DPC_WHEELBASE = 1.5706, DPC_CM0 = 4.950, DPC_CR0 = 297.030, DPC_CR1 = 16.665,
                 DPC_CR2 = 0.6784;
constexpr double DPC_INV_M = 1.0 / DPC_M, DPC_INV_LR = 2.0 / DPC_WHEELBASE;

template <typename T>
struct DpcWeights {
  T q[4], r[2], qf[4], dt;
};

// right-hand side at one RK4 stage point; only (phi, v) matter (dpc/main.py:69-83)
template <typename T>
struct Kin4Eval {
  T s, c, dres;  // sin / cos of phi + beta, d(resistance)/d v_x
  T k0, k1, k2, k3;
};

// the reference computes in fp64; the fp32 instantiation is the optional reduced-precision path of SURVEY 8(b)
__device__ __forceinline__ void sincos_(double x, double* s, double* c) { sincos(x, s, c); }
__device__ __forceinline__ void sincos_(float x, float* s, float* c) { sincosf(x, s, c); }
__device__ __forceinline__ double tanh_(double x) { return tanh(x); }
__device__ __forceinline__ float tanh_(float x) { return tanhf(x); }
template <typename T> struct Vec4;
template <> struct Vec4<double> { using type = double4; };
template <> struct Vec4<float> { using type = float4; };

#ifndef DPC_NOINLINE
#define DPC_NOINLINE 0
#endif
#ifndef DPC_UNROLL_H
#define DPC_UNROLL_H 1
#endif
#ifndef DPC_MIN_BLOCKS
#define DPC_MIN_BLOCKS 8
#endif
#ifndef DPC_BLOCK
#define DPC_BLOCK 128
#endif
#if DPC_NOINLINE
#define DPC_INLINE __noinline__
#else
#define DPC_INLINE __forceinline__
#endif
#define DPC_PRAGMA_(x) _Pragma(#x)
#define DPC_PRAGMA(x) DPC_PRAGMA_(x)

template <typename T, bool WITH_DERIV>
__device__ DPC_INLINE Kin4Eval<T> kin4(T phi, T v, T Tq, T beta, T sb, T cb) {
  constexpr T CR0 = T(DPC_CR0), CR1 = T(DPC_CR1), CR2 = T(DPC_CR2), CM0 = T(DPC_CM0), INV_M = T(DPC_INV_M),
              INV_LR = T(DPC_INV_LR);
  Kin4Eval<T> e;
  sincos_(phi + beta, &e.s, &e.c);
  const T vx = v * cb;
  const T th = tanh_(T(10) * vx);
  const T poly = fma(vx, fma(vx, CR2, CR1), CR0);
  e.k0 = v * e.c;
  e.k1 = v * e.s;
  e.k2 = v * sb * INV_LR;
  e.k3 = (CM0 * Tq - poly * th) * INV_M;
  if (WITH_DERIV) e.dres = fma(fma(T(2) * CR2, vx, CR1), th, poly * T(10) * (T(1) - th * th));
  return e;
}

// tracking term of one stage and its gradient in the state (dpc/main.py:377-388)
template <typename T, typename V4>
__device__ __forceinline__ T track_term(const T x[4], const V4 r, const T* q, T g[4]) {
  T sp, cp;
  sincos_(r.z, &sp, &cp);
  const T dX = x[0] - r.x, dY = x[1] - r.y;
  const T elon = cp * dX + sp * dY, elat = cp * dY - sp * dX;
  const T ephi = x[2] - r.z, ev = x[3] - r.w;
  const T wl = T(2) * q[0] * elon, wt = T(2) * q[1] * elat;
  g[0] = wl * cp - wt * sp;
  g[1] = wl * sp + wt * cp;
  g[2] = T(2) * q[2] * ephi;
  g[3] = T(2) * q[3] * ev;
  return q[0] * elon * elon + q[1] * elat * elat + q[2] * ephi * ephi + q[3] * ev * ev;
}

template <int NF, typename T>
__global__ void __launch_bounds__(DPC_BLOCK, DPC_MIN_BLOCKS)
dpc_rollout_grad_kernel(long long B, DpcWeights<T> w, const T* __restrict__ x0, const T* __restrict__ u,
                        const T* __restrict__ ref, T* __restrict__ cost, T* __restrict__ grad, T* __restrict__ xtraj) {
  static_assert(NF % 2 == 0, "stages move in pairs");
  using V4 = typename Vec4<T>::type;
  constexpr T CM0 = T(DPC_CM0), INV_M = T(DPC_INV_M), INV_LR = T(DPC_INV_LR);
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const T dt = w.dt;
  const V4* up = reinterpret_cast<const V4*>(u + i * (NF * 2));
  const V4* rp = reinterpret_cast<const V4*>(ref + i * ((NF + 1) * 4));
  V4* xo = xtraj ? reinterpret_cast<V4*>(xtraj + i * ((NF + 1) * 4)) : nullptr;

  // per stage k: (phi_k, v_k) and the state gradient of the tracking term at stage k + 1
  T sv[NF][2], sg[NF][4];
  T x[4];
  {
    const V4 xi = *reinterpret_cast<const V4*>(x0 + i * 4);
    x[0] = xi.x; x[1] = xi.y; x[2] = xi.z; x[3] = xi.w;
  }
  auto put = [](V4* p, const T* v) { V4 o; o.x = v[0]; o.y = v[1]; o.z = v[2]; o.w = v[3]; *p = o; };
  if (xo) put(xo, x);
  T J = T(0);
#pragma unroll 1
  for (int kk = 0; kk < NF; kk += 2) {
    const V4 u2 = up[kk >> 1];
DPC_PRAGMA(unroll DPC_UNROLL_H)
    for (int h = 0; h < 2; h++) {
      const int k = kk + h;
      const T Tq = h ? u2.z : u2.x, delta = h ? u2.w : u2.y;
      const T beta = T(0.5) * delta;
      T sb, cb;
      sincos_(beta, &sb, &cb);
      sv[k][0] = x[2];
      sv[k][1] = x[3];
      const Kin4Eval<T> e1 = kin4<T, false>(x[2], x[3], Tq, beta, sb, cb);
      const Kin4Eval<T> e2 = kin4<T, false>(fma(T(0.5) * dt, e1.k2, x[2]), fma(T(0.5) * dt, e1.k3, x[3]), Tq, beta, sb, cb);
      const Kin4Eval<T> e3 = kin4<T, false>(fma(T(0.5) * dt, e2.k2, x[2]), fma(T(0.5) * dt, e2.k3, x[3]), Tq, beta, sb, cb);
      const Kin4Eval<T> e4 = kin4<T, false>(fma(dt, e3.k2, x[2]), fma(dt, e3.k3, x[3]), Tq, beta, sb, cb);
      const T h6 = dt * T(1.0 / 6.0);
      x[0] = fma(h6, e1.k0 + T(2) * e2.k0 + T(2) * e3.k0 + e4.k0, x[0]);
      x[1] = fma(h6, e1.k1 + T(2) * e2.k1 + T(2) * e3.k1 + e4.k1, x[1]);
      x[2] = fma(h6, e1.k2 + T(2) * e2.k2 + T(2) * e3.k2 + e4.k2, x[2]);
      x[3] = fma(h6, e1.k3 + T(2) * e2.k3 + T(2) * e3.k3 + e4.k3, x[3]);
      J += w.r[0] * Tq * Tq + w.r[1] * delta * delta;
      J += track_term<T, V4>(x, rp[k + 1], (k + 1 == NF) ? w.qf : w.q, sg[k]);
      if (xo) put(xo + k + 1, x);
    }
  }
  cost[i] = J;

  // backward: a = d J / d x_{k+1}
  V4* gp = reinterpret_cast<V4*>(grad + i * (NF * 2));
  T a[4] = {T(0), T(0), T(0), T(0)};
#pragma unroll 1
  for (int kk = NF - 2; kk >= 0; kk -= 2) {
    const V4 u2 = up[kk >> 1];
    V4 g2;
DPC_PRAGMA(unroll DPC_UNROLL_H)
    for (int h = 1; h >= 0; h--) {
      const int k = kk + h;
      const T Tq = h ? u2.z : u2.x, delta = h ? u2.w : u2.y;
      const T beta = T(0.5) * delta;
      T sb, cb;
      sincos_(beta, &sb, &cb);
#pragma unroll
      for (int c = 0; c < 4; c++) a[c] += sg[k][c];
      const T phi = sv[k][0], v = sv[k][1];
      const Kin4Eval<T> e1 = kin4<T, true>(phi, v, Tq, beta, sb, cb);
      const T p2 = fma(T(0.5) * dt, e1.k2, phi), v2 = fma(T(0.5) * dt, e1.k3, v);
      const Kin4Eval<T> e2 = kin4<T, true>(p2, v2, Tq, beta, sb, cb);
      const T p3 = fma(T(0.5) * dt, e2.k2, phi), v3 = fma(T(0.5) * dt, e2.k3, v);
      const Kin4Eval<T> e3 = kin4<T, true>(p3, v3, Tq, beta, sb, cb);
      const T p4 = fma(dt, e3.k2, phi), v4 = fma(dt, e3.k3, v);
      const Kin4Eval<T> e4 = kin4<T, true>(p4, v4, Tq, beta, sb, cb);
      // adjoint of one stage evaluation: kb = d J / d k (4), returns the (phi, v) part of J_x' kb and adds J_u' kb
      T gT = T(0), gd = T(0);
      auto vjp = [&](const Kin4Eval<T>& e, T vs, T kb0, T kb1, T kb2, T kb3, T& xphi, T& xv) {
        const T wv = kb1 * e.c - kb0 * e.s;
        const T fr = kb3 * e.dres * INV_M;
        xphi = vs * wv;
        xv = kb0 * e.c + kb1 * e.s + kb2 * sb * INV_LR - fr * cb;
        gT += kb3 * (CM0 * INV_M);
        gd += T(0.5) * vs * (wv + kb2 * cb * INV_LR + fr * sb);
      };
      const T h6 = dt * T(1.0 / 6.0), h3 = dt * T(1.0 / 3.0);
      T xp4, xv4, xp3, xv3, xp2, xv2, xp1, xv1;
      vjp(e4, v4, h6 * a[0], h6 * a[1], h6 * a[2], h6 * a[3], xp4, xv4);
      vjp(e3, v3, h3 * a[0], h3 * a[1], fma(dt, xp4, h3 * a[2]), fma(dt, xv4, h3 * a[3]), xp3, xv3);
      vjp(e2, v2, h3 * a[0], h3 * a[1], fma(T(0.5) * dt, xp3, h3 * a[2]), fma(T(0.5) * dt, xv3, h3 * a[3]), xp2, xv2);
      vjp(e1, v, h6 * a[0], h6 * a[1], fma(T(0.5) * dt, xp2, h6 * a[2]), fma(T(0.5) * dt, xv2, h6 * a[3]), xp1, xv1);
      a[2] += xp1 + xp2 + xp3 + xp4;
      a[3] += xv1 + xv2 + xv3 + xv4;
      gT = fma(T(2) * w.r[0], Tq, gT);
      gd = fma(T(2) * w.r[1], delta, gd);
      if (h) { g2.z = gT; g2.w = gd; } else { g2.x = gT; g2.y = gd; }
    }
    gp[kk >> 1] = g2;
  }
}

template <typename T>
cudaError_t launch_dpc_t(long long B, int Nf, double dt, const double* weights, const T* x0, const T* u, const T* ref,
                         T* cost, T* grad, T* xtraj, cudaStream_t stream) {
  DpcWeights<T> w;
  for (int c = 0; c < 4; c++) { w.q[c] = T(weights[c]); w.qf[c] = T(weights[6 + c]); }
  w.r[0] = T(weights[4]);
  w.r[1] = T(weights[5]);
  w.dt = T(dt);
  const unsigned grid = (unsigned)((B + DPC_BLOCK - 1) / DPC_BLOCK);
  if (Nf == 40) dpc_rollout_grad_kernel<40, T><<<grid, DPC_BLOCK, 0, stream>>>(B, w, x0, u, ref, cost, grad, xtraj);
  else if (Nf == 80) dpc_rollout_grad_kernel<80, T><<<grid, DPC_BLOCK, 0, stream>>>(B, w, x0, u, ref, cost, grad, xtraj);
  else return cudaErrorInvalidValue;
  return cudaGetLastError();
}
}  // namespace

cudaError_t launch_dpc_rollout_grad(long long B, int Nf, double dt, const double* weights, const double* x0,
                                    const double* u, const double* ref, double* cost, double* grad, double* xtraj,
                                    cudaStream_t stream) {
  return launch_dpc_t<double>(B, Nf, dt, weights, x0, u, ref, cost, grad, xtraj, stream);
}
cudaError_t launch_dpc_rollout_grad_f32(long long B, int Nf, double dt, const double* weights, const float* x0,
                                        const float* u, const float* ref, float* cost, float* grad, float* xtraj,
                                        cudaStream_t stream) {
  return launch_dpc_t<float>(B, Nf, dt, weights, x0, u, ref, cost, grad, xtraj, stream);
}

}  // namespace ihm2b
