// tools/overlap_bw.cu — do FP64 arithmetic and shared-memory traffic overlap on one SM?  Half of the warps of every
// CTA run a DFMA loop, the other half an LDS.128 / STS.128 loop; each is timed alone and together.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k(double* out, int fp_iters, int mem_iters, int mode) {   // mode: 1 fp only, 2 mem only, 3 both
  extern __shared__ __align__(16) unsigned char raw[];
  double2* s = reinterpret_cast<double2*>(raw);
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) s[i] = make_double2(0.0, 0.0);
  __syncthreads();
  const int warp = threadIdx.x >> 5;
  double r = 0;
  if ((warp & 1) == 0) {
    if (mode & 1) {
      double a[8];
      for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-9 + i;
      for (int it = 0; it < fp_iters; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = __fma_rn(a[i], 1.0000001, 1e-9);
      for (int i = 0; i < 8; ++i) r += a[i];
    }
  } else if (mode & 2) {
    double2 acc = make_double2(0.0, 0.0);
    int idx = threadIdx.x;
    for (int it = 0; it < mem_iters; ++it) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const double2 v = s[idx];
        acc.x = __longlong_as_double(__double_as_longlong(acc.x) ^ __double_as_longlong(v.x));
        acc.y = __longlong_as_double(__double_as_longlong(acc.y) ^ __double_as_longlong(v.y));
        s[idx ^ 1024] = acc;
        idx = (idx + 256) & 1023;
      }
    }
    r = acc.x + acc.y;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

int main() {
  const int blocks = 148 * 2, threads = 256;
  double* out;
  cudaMalloc(&out, sizeof(double) * blocks * threads);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int fp_iters = 40000, mem_iters = 10000;     // alone: 8 fp warps/SM -> 40000*8*2*2 cyc/SMSP...; tuned to similar times
  for (int rep = 0; rep < 2; ++rep)
    for (int mode = 1; mode <= 3; ++mode) {
      cudaEventRecord(e0);
      k<<<blocks, threads, 32768>>>(out, fp_iters, mem_iters, mode);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep) printf("mode %d (%s): %.3f ms\n", mode, mode == 1 ? "fp64 only" : mode == 2 ? "smem only" : "both", ms);
    }
  return 0;
}
