// Microbenchmark: does fma.rn.f32x2 (packed FP32 FMA, sm_100) raise the FP32 FMA ceiling or only halve issue slots?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 ffma2.cu -o ffma2 && ./ffma2
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k_ffma(float* out, int iters, float seed) {
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; i++) a[i] = seed + i;
  const float m = 0.999f, c = 1e-3f;
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++)
#pragma unroll
      for (int i = 0; i < 16; i++) a[i] = fmaf(a[i], m, c);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += a[i];
  if (s == 123456789.f) out[0] = s;
}

__global__ void __launch_bounds__(256) k_ffma2(float* out, int iters, float seed) {
  unsigned long long a[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    float lo = seed + 2 * i, hi = seed + 2 * i + 1;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a[i]) : "f"(lo), "f"(hi));
  }
  unsigned long long m, c;
  asm("mov.b64 %0, {%1, %1};" : "=l"(m) : "f"(0.999f));
  asm("mov.b64 %0, {%1, %1};" : "=l"(c) : "f"(1e-3f));
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++)
#pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[i]) : "l"(m), "l"(c));
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a[i]));
    s += lo + hi;
  }
  if (s == 123456789.f) out[0] = s;
}

// FFMA2 with one shared-memory 128-bit load per 8 packed FMAs (the register-tile inner loop shape)
__global__ void __launch_bounds__(256) k_mix(float* out, int iters, float seed, int packed) {
  __shared__ float4 sm[256];
  sm[threadIdx.x] = make_float4(seed, seed + 1, seed + 2, seed + 3);
  __syncthreads();
  unsigned long long a[8];
  float f[16];
#pragma unroll
  for (int i = 0; i < 16; i++) f[i] = seed + i;
#pragma unroll
  for (int i = 0; i < 8; i++) asm("mov.b64 %0, {%1, %2};" : "=l"(a[i]) : "f"(f[2 * i]), "f"(f[2 * i + 1]));
  int idx = threadIdx.x & 31;
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const float4 w = sm[(idx + u) & 255];
      if (packed) {
        unsigned long long w01, w23;
        asm("mov.b64 %0, {%1, %2};" : "=l"(w01) : "f"(w.x), "f"(w.y));
        asm("mov.b64 %0, {%1, %2};" : "=l"(w23) : "f"(w.z), "f"(w.w));
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
          asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[i]) : "l"(w01), "l"(w23));
          asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[i + 1]) : "l"(w23), "l"(w01));
        }
      } else {
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
          f[i] = fmaf(f[i], w.x, w.y); f[i + 1] = fmaf(f[i + 1], w.y, w.z);
          f[i + 2] = fmaf(f[i + 2], w.z, w.w); f[i + 3] = fmaf(f[i + 3], w.w, w.x);
        }
      }
    }
    idx += 4;
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a[i]));
    s += lo + hi;
  }
#pragma unroll
  for (int i = 0; i < 16; i++) s += f[i];
  if (s == 123456789.f) out[0] = s;
}

template <class F>
static double run(F launch, double flops) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0;
  for (int r = 0; r < 5; r++) {
    cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double t = flops / (ms * 1e-3) / 1e12;
    if (r && t > best) best = t;
  }
  return best;
}

int main() {
  float* out; cudaMalloc(&out, 64);
  const int grid = 148 * 8, block = 256, iters = 4096;
  const double fl = 2.0 * 64 * iters * (double)grid * block;
  printf("FFMA   scalar : %.2f TFLOP/s\n", run([&] { k_ffma<<<grid, block>>>(out, iters, 1.f); }, fl));
  printf("FFMA2  packed : %.2f TFLOP/s\n", run([&] { k_ffma2<<<grid, block>>>(out, iters, 1.f); }, fl));
  printf("mix scalar FFMA + LDS.128 per 16 FMA : %.2f TFLOP/s\n", run([&] { k_mix<<<grid, block>>>(out, iters, 1.f, 0); }, fl));
  printf("mix packed FFMA2 + LDS.128 per 16 FMA: %.2f TFLOP/s\n", run([&] { k_mix<<<grid, block>>>(out, iters, 1.f, 1); }, fl));
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
