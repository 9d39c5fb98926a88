// fp64_peak.cu — FP64 micro-benchmarks that give the roofline denominator for the
// compute-bound configs (MEASURED_PEAKS.json records no FP64 figure).
//   kind 0: DFMA register loop (8 independent accumulators per thread)
//   kind 1: DMMA  mma.sync.aligned.m8n8k4.row.col.f64  (4 independent accumulator tiles)
// tcgen05 has no FP64 kind, so these two are the only FP64 pipes on sm_100a.
#include <cuda_runtime.h>
#include "kernels.cuh"

namespace garb200 {

__global__ void __launch_bounds__(256) dfma_kernel(double *out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
      x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
  }
  const double s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) dmma_kernel(double *out, int iters, double a, double b) {
  double c0[2] = {0, 0}, c1[2] = {1, 1}, c2[2] = {2, 2}, c3[2] = {3, 3};
  const double fa = a + threadIdx.x * 1e-9, fb = b;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0[0]), "+d"(c0[1]) : "d"(fa), "d"(fb));
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c1[0]), "+d"(c1[1]) : "d"(fa), "d"(fb));
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c2[0]), "+d"(c2[1]) : "d"(fa), "d"(fb));
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c3[0]), "+d"(c3[1]) : "d"(fa), "d"(fb));
    }
  }
  const double s = c0[0] + c0[1] + c1[0] + c1[1] + c2[0] + c2[1] + c3[0] + c3[1];
  if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int fp64_peak(int kind, double *tflops) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -2;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -2;
  double *out = nullptr;
  const int grid = sms * 8, block = 256, iters = 4096;
  if (cudaMalloc(&out, (size_t)grid * block * sizeof(double)) != cudaSuccess) return -2;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0.0;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    if (kind == 0) dfma_kernel<<<grid, block>>>(out, iters, 0.999999, 1e-7);
    else dmma_kernel<<<grid, block>>>(out, iters, 0.999999, 1e-7);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return -2; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    double flops;
    if (kind == 0) flops = 2.0 * 64.0 * iters * (double)grid * block;          // 64 FMA / iter / thread
    else flops = 2.0 * 16.0 * 256.0 * iters * ((double)grid * block / 32.0);    // 16 MMA(8x8x4) / iter / warp
    const double tf = flops / (ms * 1e-3) / 1e12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(out);
  *tflops = best;
  return 0;
}

} // namespace garb200
