// peaks.cu -- FMA-issue microbenchmark.  MEASURED_PEAKS.json carries only HBM and bf16-tensor peaks, but the
// rollout and update kernels are bound by CUDA-core FMA issue, so bench.py measures that
// denominator live: 8 independent dependent-FMA chains per thread, enough warps to saturate every SMSP.
#include "engine.h"

namespace b2 {

template <typename T>
__global__ void __launch_bounds__(256) k_fma_peak(T* out, int iters, T seed) {
  T a0 = seed, a1 = seed + 1, a2 = seed + 2, a3 = seed + 3, a4 = seed + 4, a5 = seed + 5, a6 = seed + 6, a7 = seed + 7;
  const T m = (T)0.999, c = (T)1e-3;
#pragma unroll 1
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      a0 = a0 * m + c; a1 = a1 * m + c; a2 = a2 * m + c; a3 = a3 * m + c;
      a4 = a4 * m + c; a5 = a5 * m + c; a6 = a6 * m + c; a7 = a7 * m + c;
    }
  }
  T s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (s == (T)123456789) out[0] = s;  // never true: keeps the chains alive
}

cudaError_t measure_fma_peak(bool fp64, int sm_count, cudaStream_t st, double* tflops) {
  const int grid = sm_count * 8, block = 256, iters = fp64 ? 4096 : 8192;
  void* scratch = nullptr;
  cudaError_t e = cudaMalloc(&scratch, 64);
  if (e != cudaSuccess) return e;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0, st);
    if (fp64) k_fma_peak<double><<<grid, block, 0, st>>>((double*)scratch, iters, 1.0);
    else k_fma_peak<float><<<grid, block, 0, st>>>((float*)scratch, iters, 1.0f);
    cudaEventRecord(e1, st);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * 64.0 * iters * (double)grid * block;
    if (rep > 0 && ms > 0) best = flops / (ms * 1e-3) / 1e12 > best ? flops / (ms * 1e-3) / 1e12 : best;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(scratch);
  *tflops = best;
  return cudaGetLastError();
}

}  // namespace b2
