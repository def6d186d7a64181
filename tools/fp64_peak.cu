// tools/fp64_peak.cu — measures the FP64 FMA rate of the GPU (the roof the evaluation kernels actually hit:
// tcgen05 has no f64 kind, so FP64 work runs on the DFMA pipe).  nvcc -O3 -arch=sm_100a tools/fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void dfma_kernel(double* out, int iters) {
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double b = 1.0000001, c = 1e-9;
  for (int i = 0; i < iters; ++i) {
    a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
    a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// mode 0: DMMA only; mode 1: DMMA and DFMA interleaved (do the two paths share one pipe?)
__global__ void dmma_kernel(double* out, int iters, int mode) {
  double c[8][2];
  for (int i = 0; i < 8; ++i) { c[i][0] = threadIdx.x * 1e-9; c[i][1] = i; }
  double f0 = 1.0, f1 = 2.0, f2 = 3.0, f3 = 4.0, f4 = 5, f5 = 6, f6 = 7, f7 = 8;
  const double a = 1.0000001, b = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dmma(c[i][0], c[i][1], a, b);
    if (mode == 1) {
      f0 = fma(f0, a, b); f1 = fma(f1, a, b); f2 = fma(f2, a, b); f3 = fma(f3, a, b);
      f4 = fma(f4, a, b); f5 = fma(f5, a, b); f6 = fma(f6, a, b); f7 = fma(f7, a, b);
    }
  }
  double s = f0 + f1 + f2 + f3 + f4 + f5 + f6 + f7;
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int blocks = sms * 8, threads = 256, iters = 1 << 16;
  double* out;
  cudaMalloc(&out, sizeof(double) * blocks * threads);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  dfma_kernel<<<blocks, threads>>>(out, 1024);
  cudaDeviceSynchronize();
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    dfma_kernel<<<blocks, threads>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fmas = (double)blocks * threads * iters * 8;
    printf("{\"sms\": %d, \"dfma_per_s\": %.4e, \"fp64_tflops\": %.2f, \"ms\": %.3f}\n", sms, fmas / (ms * 1e-3),
           2 * fmas / (ms * 1e-3) / 1e12, ms);
  }
  for (int mode = 0; mode < 2; ++mode) {
    dmma_kernel<<<blocks, threads>>>(out, 256, mode);
    cudaDeviceSynchronize();
    const int it2 = 1 << 14;
    cudaEventRecord(e0);
    dmma_kernel<<<blocks, threads>>>(out, it2, mode);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double warps = (double)blocks * threads / 32;
    const double mma_fma = warps * it2 * 8 * 256.0;                 // m8n8k4 = 256 FMA per warp instruction
    const double vec_fma = mode ? (double)blocks * threads * it2 * 8 : 0.0;
    printf("{\"mode\": \"%s\", \"dmma_tflops\": %.2f, \"dfma_tflops\": %.2f, \"ms\": %.3f}\n", mode ? "dmma+dfma" : "dmma",
           2 * mma_fma / (ms * 1e-3) / 1e12, 2 * vec_fma / (ms * 1e-3) / 1e12, ms);
  }
  return 0;
}
