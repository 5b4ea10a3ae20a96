// Micro-benchmark: dependent-chain latency and per-SM throughput of mma.sync.m8n8k4.f64 (DMMA),
// DFMA, shared-memory loads and warp shuffles on sm_100a.  nvcc -arch=sm_100a -O3 -o dmma_lat dmma_lat.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int chains> __global__ void k_dmma(double* out, long long* cyc, int iters) {
  double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  double c[8][2] = {};
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++) if (j < chains) dmma(c[j][0], c[j][1], a, b);
  }
  long long t1 = clock64();
  double s = 0; for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int chains> __global__ void k_dfma(double* out, long long* cyc, int iters) {
  double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
  double c[8] = {};
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++) if (j < chains) c[j] = fma(c[j], a, b);
  }
  long long t1 = clock64();
  double s = 0; for (int j = 0; j < 8; j++) s += c[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_lds(double* out, long long* cyc, int iters) {
  __shared__ double sm[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = (double)((i * 7 + 3) & 1023);
  __syncthreads();
  int idx = threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) idx = (int)sm[idx & 1023];
  long long t1 = clock64();
  out[threadIdx.x] = idx;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_shfl(double* out, long long* cyc, int iters) {
  double v = threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) v = __shfl_xor_sync(0xffffffffu, v, 1) + 1.0;
  long long t1 = clock64();
  out[threadIdx.x] = v;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  double* out; long long* cyc; cudaMalloc(&out, 1 << 24); cudaMalloc(&cyc, 8);
  long long h; const int it = 4096;
#define RUN1(K, C, name) { K<C><<<1, 32>>>(out, cyc, it); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); \
    printf(name " 1 warp, %d chains: %.2f cycles per round, %.2f per instr\n", C, (double)h / it, (double)h / it / C); }
  RUN1(k_dmma, 1, "DMMA") RUN1(k_dmma, 2, "DMMA") RUN1(k_dmma, 4, "DMMA") RUN1(k_dmma, 8, "DMMA")
  RUN1(k_dfma, 1, "DFMA") RUN1(k_dfma, 2, "DFMA") RUN1(k_dfma, 4, "DFMA") RUN1(k_dfma, 8, "DFMA")
  k_lds<<<1, 32>>>(out, cyc, it); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("LDS.64 dependent (+cvt): %.2f cycles\n", (double)h / it);
  k_shfl<<<1, 32>>>(out, cyc, it); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("SHFL.64 + DADD dependent: %.2f cycles\n", (double)h / it);
  // throughput: many warps per SM
  for (int warps : {4, 8, 16, 32}) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_dmma<8><<<148, 32 * warps>>>(out, cyc, it);
    cudaEventRecord(e0); k_dmma<8><<<148, 32 * warps>>>(out, cyc, it * 4); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double flops = 148.0 * warps * it * 4 * 8 * (2.0 * 8 * 8 * 4);
    printf("DMMA throughput %d warps/SM: %.2f TFLOP/s\n", warps, flops / ms * 1e-9);
    k_dfma<8><<<148, 32 * warps>>>(out, cyc, it);
    cudaEventRecord(e0); k_dfma<8><<<148, 32 * warps>>>(out, cyc, it * 4); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    flops = 148.0 * warps * 32 * it * 4 * 8 * 2.0;
    printf("DFMA throughput %d warps/SM: %.2f TFLOP/s\n", warps, flops / ms * 1e-9);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
