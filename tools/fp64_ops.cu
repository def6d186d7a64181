// tools/fp64_ops.cu — issue rate of DADD, DMUL and DFMA (and mixes) on the FP64 pipe: thread-instructions per second.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/fp64_ops tools/fp64_ops.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(double* out, int iters) {
  double a[8];
  for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-9 + i;
  const double b = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) a[i] = __fma_rn(a[i], b, c);
      else if (MODE == 1) a[i] = __dadd_rn(a[i], c);
      else if (MODE == 2) a[i] = __dmul_rn(a[i], b);
      else if (MODE == 3) a[i] = (i & 1) ? __dadd_rn(a[i], c) : __fma_rn(a[i], b, c);       // half adds, half fmas
      else if (MODE == 4) a[i] = __fma_rn(a[i], 1.0, c);                                    // an add written as an fma
      else if (MODE == 5) a[i] = __fma_rn(a[i], b, -0.0);                                   // a mul written as an fma
    }
  }
  double s = 0;
  for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE> void run(const char* name) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int blocks = sms * 8, threads = 256, iters = 1 << 15;
  double* out;
  cudaMalloc(&out, sizeof(double) * blocks * threads);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<blocks, threads>>>(out, 256);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(out, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double n = (double)blocks * threads * iters * 8;
  printf("{\"op\": \"%s\", \"thread_inst_per_s\": %.4e, \"per_clk_per_sm\": %.1f, \"ms\": %.3f}\n", name, n / (ms * 1e-3),
         n / (ms * 1e-3) / sms / 1.965e9, ms);
  cudaFree(out);
}

int main() {
  run<0>("dfma"); run<1>("dadd"); run<2>("dmul"); run<3>("dadd+dfma 1:1"); run<4>("fma(a,1,c)"); run<5>("fma(a,b,-0)");
  return 0;
}
