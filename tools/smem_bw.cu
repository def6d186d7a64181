// tools/smem_bw.cu — shared-memory load/store throughput per SM for 4-, 8- and 16-byte conflict-free accesses
// (which access width the FFT stages should use).  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/smem_bw tools/smem_bw.cu
#include <cstdio>
#include <cuda_runtime.h>

template <typename T> struct Zero;
template <> struct Zero<float> { static __device__ float v() { return 0.f; } };
template <> struct Zero<double> { static __device__ double v() { return 0.0; } };
template <> struct Zero<double2> { static __device__ double2 v() { return make_double2(0.0, 0.0); } };
__device__ float acc(float a, float b) { return a + b; }
__device__ double acc(double a, double b) { return __longlong_as_double(__double_as_longlong(a) ^ __double_as_longlong(b)); }
__device__ double2 acc(double2 a, double2 b) {
  return make_double2(__longlong_as_double(__double_as_longlong(a.x) ^ __double_as_longlong(b.x)),
                      __longlong_as_double(__double_as_longlong(a.y) ^ __double_as_longlong(b.y)));
}

template <typename T, bool STORE>
__global__ void k(T* out, int iters, long long* cyc) {
  extern __shared__ __align__(16) unsigned char raw[];
  T* s = reinterpret_cast<T*>(raw);
  const int n = 32768 / sizeof(T);           // 32 KB window
  for (int i = threadIdx.x; i < n; i += blockDim.x) s[i] = Zero<T>::v();
  __syncthreads();
  T a = Zero<T>::v();
  const long long t0 = clock64();
  int idx = threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      if (STORE) s[idx] = a; else a = acc(a, s[idx]);
      idx = (idx + blockDim.x) & (n - 1);
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  if (!STORE) out[blockIdx.x * blockDim.x + threadIdx.x] = a;
  else if (iters < 0) out[0] = s[idx];
}

template <typename T, bool STORE> void run(const char* name, int threads, int ctas_per_sm) {
  int sms = 148;
  T* out; long long* cyc;
  cudaMalloc(&out, sizeof(T) * sms * ctas_per_sm * threads);
  cudaMalloc(&cyc, 8 * sms * ctas_per_sm);
  const int iters = 2000;
  cudaFuncSetAttribute(k<T, STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  k<T, STORE><<<sms * ctas_per_sm, threads, 32768>>>(out, iters, cyc);
  k<T, STORE><<<sms * ctas_per_sm, threads, 32768>>>(out, iters, cyc);
  cudaDeviceSynchronize();
  long long h[148 * 8];
  cudaMemcpy(h, cyc, 8 * sms * ctas_per_sm, cudaMemcpyDeviceToHost);
  double mean = 0;
  for (int i = 0; i < sms * ctas_per_sm; ++i) mean += h[i];
  mean /= sms * ctas_per_sm;
  const double bytes_per_sm = (double)iters * 16 * threads * sizeof(T) * ctas_per_sm;
  printf("%-10s %s  %4d thr x %d CTA/SM: %.1f B/clk/SM  (%.2f clk per warp-instruction)\n", name, STORE ? "store" : "load ",
         threads, ctas_per_sm, bytes_per_sm / mean, mean / ((double)iters * 16 * threads / 32 * ctas_per_sm));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int cfg = 0; cfg < 3; ++cfg) {
    const int threads = cfg == 0 ? 256 : cfg == 1 ? 512 : 1024, ctas = cfg == 0 ? 2 : 1;
    run<float, false>("4 B", threads, ctas);   run<double, false>("8 B", threads, ctas);   run<double2, false>("16 B", threads, ctas);
    run<float, true>("4 B", threads, ctas);    run<double, true>("8 B", threads, ctas);    run<double2, true>("16 B", threads, ctas);
  }
  return 0;
}
