// Dependent-chain latencies on sm_100a (diagnostic micro-benchmark, not part of the library).
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma_884(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int MODE> __global__ void k(double *out, long long *cyc, int iters, double seed) {
  double x = seed + threadIdx.x * 1e-9, y = 1.0000001, c0 = x, c1 = y;
  __shared__ double sm[64];
  sm[threadIdx.x & 63] = x;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) {
    if (MODE == 0) { x = fma(x, y, 1e-9); }                       // dependent DFMA
    else if (MODE == 1) { x = __drcp_rn(x) + 0.5; }               // drcp + dadd
    else if (MODE == 2) { dmma_884(c0, c1, x, y); }               // dependent DMMA (accumulator chain)
    else if (MODE == 3) { x = sm[(threadIdx.x + (int)x) & 63]; x = x * 0.0 + 1.0 + threadIdx.x; } // LDS + 1 op
    else if (MODE == 4) { x = fma(x, y, 1e-9); __syncthreads(); } // DFMA + barrier
    else if (MODE == 5) { x = 1.0 / x + 0.5; }                    // IEEE division
  }
  long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = x + c0 + c1;
}
int main() {
  double *out; long long *cyc, h;
  cudaMalloc(&out, 1 << 24); cudaMalloc(&cyc, 8);
  const int iters = 4096;
  const char *names[] = {"dep DFMA", "drcp_rn + dadd", "dep DMMA m8n8k4", "LDS + 2 ops", "DFMA + __syncthreads", "1.0/x + dadd"};
  for (int threads : {32, 160, 480}) {
    for (int blocks : {1, 148 * 1}) {
#define RUN(M) k<M><<<blocks, threads>>>(out, cyc, iters, 1.5); cudaDeviceSynchronize(); k<M><<<blocks, threads>>>(out, cyc, iters, 1.5); cudaDeviceSynchronize(); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("threads %3d blocks %3d  %-22s %.1f cycles/iter\n", threads, blocks, names[M], (double)h / iters);
      RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5)
    }
  }
  return 0;
}
