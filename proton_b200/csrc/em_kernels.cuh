// em_kernels.cuh — packing (B*u), forward DCT, constant tables, short-line and direct evaluators.
//
// Reference functions replaced: Eigen SpMV in form_analytical_solution
// (bin/EMtool_analytical@0xc46c), dct1d (@0xc156), and for short lines calc_analytical_solution +
// idct1d (@0xc68c, @0xc2a0) plus the max / threshold consumers (PROTON_cli.py:1058-1066).
#pragma once
#include "em_fft.cuh"

namespace em {

// ---------------------------------------------------------------------------------------------
// (a) Packing: r = B*u per line, with Eigen's semantics for a column-major SparseMatrix built by
// setFromTriplets: duplicates are summed first (in input order), then for every stored (row, col),
// columns ascending, res[row] += value * u[col] with two roundings (no FMA).  Output is the sparse
// vector r as sorted unique rows; r_idx / r_val use the triplet offsets as capacity.
// One CTA per line; ranks are found by counting (O(nnz^2 / 128), nnz = 2S is a few hundred at most).
// ---------------------------------------------------------------------------------------------
struct PackParams {
  long long num_lines;
  const int* num_nodes;
  const long long* trip_off;
  const int* trip_row;
  const int* trip_col;
  const double* trip_val;
  const long long* u_off;
  const double* curden;
  int* perm;          // [ntrip] scratch: sorted position -> triplet
  double* prod;       // [ntrip] scratch
  int* r_idx;         // [ntrip]
  double* r_val;      // [ntrip]
  int* r_cnt;         // [L]
  int* error_flag;    // set to 1 + line on a bad index
};

__device__ __forceinline__ bool trip_less(int ra, int ca, int ia, int rb, int cb, int ib) {
  if (ra != rb) return ra < rb;
  if (ca != cb) return ca < cb;
  return ia < ib;
}

__global__ void __launch_bounds__(128) em_pack_kernel(const PackParams p) {
  __shared__ int bad;
  for (long long line = blockIdx.x; line < p.num_lines; line += gridDim.x) {
    const long long t0 = p.trip_off[line];
    const int nt = (int)(p.trip_off[line + 1] - t0);
    const int N = p.num_nodes[line];
    const int S = (int)(p.u_off[line + 1] - p.u_off[line]);
    const int* row = p.trip_row + t0;
    const int* col = p.trip_col + t0;
    const double* val = p.trip_val + t0;
    const double* u = p.curden + p.u_off[line];
    int* perm = p.perm + t0;
    double* prod = p.prod + t0;
    if (threadIdx.x == 0) bad = (N <= 0) ? 1 : 0;
    __syncthreads();
    // 1. rank by (row, col, input order)
    for (int i = threadIdx.x; i < nt; i += blockDim.x) {
      const int ri = row[i], ci = col[i];
      if (ri < 0 || ri >= N || ci < 0 || ci >= S) bad = 1;
      int rank = 0;
      for (int j = 0; j < nt; ++j) rank += trip_less(row[j], col[j], j, ri, ci, i) ? 1 : 0;
      perm[rank] = i;
    }
    __syncthreads();
    if (bad) {                       // uniform: every thread reads the same shared flag
      if (threadIdx.x == 0) { p.r_cnt[line] = 0; atomicMax(p.error_flag, (int)min(line + 1, (long long)0x7fffffff)); }
      __syncthreads();
      continue;
    }
    // 2. per (row, col) group: sum duplicates in input order, multiply by u[col]
    for (int s = threadIdx.x; s < nt; s += blockDim.x) {
      const int i = perm[s];
      const bool start = (s == 0) || row[perm[s - 1]] != row[i] || col[perm[s - 1]] != col[i];
      if (start) {
        double v = val[i];
        for (int q = s + 1; q < nt && row[perm[q]] == row[i] && col[perm[q]] == col[i]; ++q) v = __dadd_rn(v, val[perm[q]]);
        prod[s] = __dmul_rn(v, u[col[i]]);
      }
    }
    __syncthreads();
    // 3. per row: accumulate its products in ascending column order, compact the rows
    for (int s = threadIdx.x; s < nt; s += blockDim.x) {
      const int i = perm[s];
      const bool row_start = (s == 0) || row[perm[s - 1]] != row[i];
      if (row_start) {
        int before = 0;
        for (int q = 1; q <= s; ++q) before += (row[perm[q - 1]] != row[perm[q]]) ? 1 : 0;
        double acc = 0.0;
        for (int q = s; q < nt && row[perm[q]] == row[i]; ++q) {
          const bool gstart = (q == s) || col[perm[q - 1]] != col[perm[q]];
          if (gstart) acc = __dadd_rn(acc, prod[q]);
        }
        p.r_idx[t0 + before] = row[i];
        p.r_val[t0 + before] = acc;
      }
    }
    if (threadIdx.x == 0) {
      int cnt = nt > 0 ? 1 : 0;
      for (int q = 1; q < nt; ++q) cnt += (row[perm[q - 1]] != row[perm[q]]) ? 1 : 0;
      p.r_cnt[line] = cnt;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// (b) Forward DCT-II of the sparse r, folded with everything that does not depend on time:
//   R_k = sum_v r_v cos(pi (2v+1) k / (2N))            (exact integer angle reduction)
//   W_k = R_k / N                      if k == 0 or lambda_k == 0   (then Yhat_k = t * W_k)
//       = R_k / (N * (lambda_k * a))   otherwise                    (then Yhat_k = (e^{lambda_k a t} - 1) W_k)
// which equals the reference's q_k, the 1/(lambda a) of D_kk, the sqrt scalings of dct1d/idct1d and
// Makhoul's 1/2, all in one number per mode.  One thread per (line, mode).
// ---------------------------------------------------------------------------------------------
struct ForwardParams {
  long long num_lines;
  long long total_blocks;       // sum over lines of ceil(N / kFwdBlock)
  const int* num_nodes;
  const double* a_coeff;
  const long long* node_off;
  const long long* blk_off;     // [L+1] prefix sum of ceil(N / kFwdBlock)
  const long long* trip_off;
  const int* r_idx;
  const double* r_val;
  const int* r_cnt;
  const int* line_tab;
  const NTab* tabs;
  double* W;
};

constexpr int kFwdBlock = 32;   // consecutive modes per thread

__device__ __forceinline__ long long find_line(const long long* __restrict__ off, long long L, long long g) {
  long long lo = 0, hi = L;           // off[lo] <= g < off[hi]
  while (hi - lo > 1) {
    const long long mid = (lo + hi) >> 1;
    if (off[mid] <= g) lo = mid; else hi = mid;
  }
  return lo;
}

// One thread per (line, block of 32 consecutive modes).  For every non-zero r_v the cosine of mode k0 is taken
// from an exactly reduced angle and advanced to k0+1, k0+2, ... by the rotation e^{i pi (2v+1)/(2N)} (one complex
// product per term instead of a 64-bit modulo and a sincospi); 32 steps add < 1e-14 relative error.
__global__ void __launch_bounds__(128) em_forward_kernel(const ForwardParams p) {
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < p.total_blocks;
       g += (long long)gridDim.x * blockDim.x) {
    const long long line = find_line(p.blk_off, p.num_lines, g);
    const int N = p.num_nodes[line];
    const int k0 = (int)(g - p.blk_off[line]) * kFwdBlock;
    const int kn = min(kFwdBlock, N - k0);
    const long long t0 = p.trip_off[line];
    const int cnt = p.r_cnt[line];
    const unsigned long long den = 2ULL * (unsigned long long)N;
    double R[kFwdBlock];
#pragma unroll
    for (int i = 0; i < kFwdBlock; ++i) R[i] = 0.0;
    for (int i = 0; i < cnt; ++i) {
      const unsigned long long m = 2ULL * (unsigned long long)p.r_idx[t0 + i] + 1ULL;   // < 2*den
      const double rv = p.r_val[t0 + i];
      const cplx w = em_cis_frac(m, den);
      cplx z = em_cis_frac((m * (unsigned long long)k0) % (2ULL * den), den);
#pragma unroll
      for (int q = 0; q < kFwdBlock; ++q) {
        R[q] = fma(rv, z.x, R[q]);
        z = cmul(z, w);
      }
    }
    const double* lam = p.tabs[p.line_tab[line]].lam + k0;
    const double a = p.a_coeff[line];
    double* W = p.W + p.node_off[line] + k0;
#pragma unroll
    for (int q = 0; q < kFwdBlock; ++q) {
      if (q < kn) {
        const double l = lam[q];
        double w = R[q] / (double)N;
        if (l != 0.0) w = w / __dmul_rn(l, a);
        W[q] = w;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Constant tables per distinct N on the in-CTA FFT path (see NTab).
// ---------------------------------------------------------------------------------------------
struct TableParams {
  const NTab* tabs;       // entries handled by this launch (all share LOG2P / BLUE); pre/post/hhat writable
  const cplx* tw;
};

template <int LOG2P, bool BLUE>
__global__ void __launch_bounds__(Cfg<LOG2P, BLUE>::NT) em_tables_kernel(const TableParams p) {
  constexpr int P = 1 << LOG2P, NT = Cfg<LOG2P, BLUE>::NT;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* twb = reinterpret_cast<cplx*>(smem_raw);
  cplx* data = twb + stage_tw_total(LOG2P);
  const NTab tab = p.tabs[blockIdx.x];
  const unsigned long long N = (unsigned long long)tab.n;
  cplx* pre = const_cast<cplx*>(tab.pre);
  const int tid = threadIdx.x;
  if (!BLUE) {
    for (int k = tid; k < (int)N; k += NT) pre[k] = em_cis_frac((unsigned long long)k, 2ULL * N);
    return;
  }
  cplx* post = const_cast<cplx*>(tab.post);
  cplx* hhat = const_cast<cplx*>(tab.hhat);
  for (int i = tid; i < stage_tw_total(LOG2P); i += NT) twb[i] = p.tw[i];
  for (int k = tid; k < (int)N; k += NT) {
    const unsigned long long kk = (unsigned long long)k;
    pre[k] = em_cis_frac((kk + 2ULL * kk * kk) % (4ULL * N), 2ULL * N);
    post[k] = em_cis_frac((kk * kk) % (2ULL * N), N);
  }
  for (int i = tid; i < P; i += NT) data[pad_idx<LOG2P>(i)] = cplx{0.0, 0.0};
  __syncthreads();
  const double inv_p = 1.0 / (double)P;
  for (int j = tid; j < (int)N; j += NT) {
    const unsigned long long jj = (unsigned long long)j;
    const cplx c = em_cis_frac((jj * jj) % (2ULL * N), N);
    const cplx h = cplx{c.x * inv_p, -c.y * inv_p};
    data[pad_idx<LOG2P>(j)] = h;
    if (j) data[pad_idx<LOG2P>(P - j)] = h;
  }
  __syncthreads();
  DifStages<LOG2P, 0, plan_n(LOG2P), NT>::run(data, 1, twb, tid);
  for (int i = tid; i < P; i += NT) hhat[hh_pos_of(P, plan_radix(LOG2P, plan_n(LOG2P) - 1), i)] = data[pad_idx<LOG2P>(i)];
}

// Parameters of the short-line (dense) kernel.
struct SmallParams {
  const int* class_lines;
  int n_class_lines;
  int chunks;
  int ts_per_chunk;
  int T;
  const int* num_nodes;
  const double* a_coeff;
  const long long* node_off;
  const int* line_tab;
  const NTab* tabs;
  const double* W;
  const double* ts;
  const ulonglong2* exp_tab;
  double* stress;
  double* mx;
  int* amx;
  unsigned char* flag;
  double critical;
  int max_nodes;               // longest line of the class (sizes the dense kernel's shared memory)
};

// ---------------------------------------------------------------------------------------------
// Direct O(N^2) evaluator for any N: one CTA per (line, timestamp).  Not a production path — it is
// the device-side cross-check the tests use (opts.algorithm = DIRECT), with every cosine obtained
// from an exactly reduced angle.
// ---------------------------------------------------------------------------------------------
struct DirectParams {
  long long num_lines;         // lines of this launch
  long long line0;             // first line of this launch (a launch covers the contiguous lines of its chunks)
  int T;
  const int* num_nodes;
  const double* a_coeff;
  const long long* node_off;
  const int* line_tab;
  const NTab* tabs;
  const double* W;
  const double* ts;
  const ulonglong2* exp_tab;
  double* stress;
  double* mx;
  int* amx;
  unsigned char* flag;
  double critical;
};

__global__ void __launch_bounds__(256) em_eval_direct_kernel(const DirectParams p) {
  constexpr int KC = 512;
  __shared__ ulonglong2 exp_tab[kExpTableEntries];
  __shared__ double ych[KC];
  __shared__ double red_v[8];
  __shared__ int red_i[8];
  __shared__ double first_s;
  for (int i = threadIdx.x; i < kExpTableEntries; i += blockDim.x) exp_tab[i] = p.exp_tab[i];
  __syncthreads();
  const long long rel = blockIdx.x / p.T;
  const int j = (int)(blockIdx.x - rel * p.T);
  const long long line = p.line0 + rel;
  const int N = p.num_nodes[line];
  const double a = p.a_coeff[line];
  const long long noff = p.node_off[line];
  const double* lam = p.tabs[p.line_tab[line]].lam;
  const double t = p.ts[j];
  const double w = __dmul_rn(a, t);
  const unsigned long long den = 2ULL * (unsigned long long)N;
  double best = -__longlong_as_double(0x7ff0000000000000LL);
  int besti = 0x7fffffff;
  for (int n0 = 0; n0 < N; n0 += 256) {
    const int n = n0 + threadIdx.x;
    double acc = 0.0;
    for (int k0 = 0; k0 < N; k0 += KC) {
      __syncthreads();
      for (int k = k0 + threadIdx.x; k < min(N, k0 + KC); k += 256) {
        double wk = p.W[noff + k];
        if (k > 0) wk = wk + wk;
        ych[k - k0] = em_modal_exact(lam[k], wk, w, t, exp_tab);
      }
      __syncthreads();
      if (n < N) {
        for (int k = k0; k < min(N, k0 + KC); ++k) {
          const unsigned long long num = ((unsigned long long)k * (2ULL * n + 1ULL)) % (2ULL * den);
          acc = fma(ych[k - k0], em_cos_frac(num, den), acc);
        }
      }
    }
    if (n < N) {
      if (p.stress) p.stress[noff * (long long)p.T + (long long)j * N + n] = acc;
      if (acc > best) { best = acc; besti = n; }
      if (n == 0) first_s = acc;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
    if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
  }
  if ((threadIdx.x & 31) == 0) { red_v[threadIdx.x >> 5] = best; red_i[threadIdx.x >> 5] = besti; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int q = 1; q < 8; ++q)
      if (red_v[q] > best || (red_v[q] == best && red_i[q] < besti)) { best = red_v[q]; besti = red_i[q]; }
    if (besti == 0x7fffffff) { besti = 0; best = first_s; }
    const long long o = line * p.T + j;
    p.mx[o] = best; p.amx[o] = besti; p.flag[o] = (best > p.critical) ? 1 : 0;
  }
}

// exp() probe used by the bit-exactness test.
__global__ void em_exp_probe_kernel(const double* x, double* y, long long n, const ulonglong2* tab) {
  __shared__ ulonglong2 exp_tab[kExpTableEntries];
  for (int i = threadIdx.x; i < kExpTableEntries; i += blockDim.x) exp_tab[i] = tab[i];
  __syncthreads();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = em_exp(x[i], exp_tab);
}

// cos(pi m / 2N), m < 4N: the cosine table em_eval_dense_kernel builds in shared memory (test hook for the CPU twin).
__global__ void em_costab_probe_kernel(double* y, int N) {
  for (int m = blockIdx.x * blockDim.x + threadIdx.x; m < 4 * N; m += gridDim.x * blockDim.x)
    y[m] = em_cos_frac((unsigned long long)m, 2ULL * N);
}

// exp(x) - 1.0 as the fused kernels evaluate it (em_expm1_ref), for the parity test of the modal factor.
__global__ void em_expm1_probe_kernel(const double* x, double* y, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = em_expm1_ref(x[i]);
}

}  // namespace em

namespace em {

// ---------------------------------------------------------------------------------------------
// Short lines on the FP64 tensor-core path: sigma[N x T] = C[N x N] * Y[N x T] as DMMA m8n8k4 tiles
// (C[n][k] = cos(pi k (2n+1) / 2N), Y[k][t] = modal value of mode k at timestamp t).  One CTA per (line, chunk of
// timestamps): the cosine matrix is built once in shared memory in A-fragment order, every warp then takes tiles of 8
// timestamps; each lane evaluates the modal factor of exactly the (k, t) element the B fragment wants from it, so no
// data is exchanged between lanes before the MMAs.  Replaces calc_analytical_solution + idct1d
// (bin/EMtool_analytical@0xc68c, @0xc2a0) for N <= kDenseMaxN.
// ---------------------------------------------------------------------------------------------
constexpr int kDenseMaxN = 96;

__device__ __forceinline__ void dmma_m8n8k4(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int MTMAX>    // m-tiles of 8 nodes the instantiation can hold: lines of up to 8 * MTMAX nodes
__global__ void __launch_bounds__(128) em_eval_dense_kernel(const SmallParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* ctab = reinterpret_cast<double*>(smem_raw);              // [4 * Nmax]
  double* afrag = ctab + 4 * p.max_nodes;                          // [MT * KS * 32]
  const int item = blockIdx.x;
  const int li = item / p.chunks, chunk = item - li * p.chunks;
  const int line = p.class_lines[li];
  const int N = p.num_nodes[line];
  const double a = p.a_coeff[line];
  const long long noff = p.node_off[line];
  const int MT = (N + 7) >> 3, KS = (N + 3) >> 2;
  for (int m = threadIdx.x; m < 4 * N; m += blockDim.x) ctab[m] = em_cos_frac((unsigned long long)m, 2ULL * N);
  __syncthreads();
  // A fragment of tile (mt, ks): lane holds C[n = 8 mt + lane/4][k = 4 ks + lane%4]
  for (int e = threadIdx.x; e < MT * KS * 32; e += blockDim.x) {
    const int l = e & 31, t = e >> 5;
    const int ks = t % KS, mt = t / KS;
    const int n = 8 * mt + (l >> 2), k = 4 * ks + (l & 3);
    afrag[e] = (n < N && k < N) ? ctab[(k * (2 * n + 1)) % (4 * N)] : 0.0;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const double* lam = p.tabs[p.line_tab[line]].lam;
  const double* W = p.W + noff;
  const int j0 = chunk * p.ts_per_chunk, j1 = min(p.T, j0 + p.ts_per_chunk);
  const double ninf = -__longlong_as_double(0x7ff0000000000000LL);
  for (int tb = j0 + 8 * warp; tb < j1; tb += 8 * nwarps) {
    const int tB = tb + (lane >> 2);                       // timestamp of this lane's B elements
    const bool tvalid = tB < j1;
    const double tt = p.ts[tvalid ? tB : j0];
    const double w = __dmul_rn(a, tt);
    double d[MTMAX][2];
#pragma unroll
    for (int mt = 0; mt < MTMAX; ++mt) { d[mt][0] = 0.0; d[mt][1] = 0.0; }
    for (int ks = 0; ks < KS; ++ks) {
      const int k = 4 * ks + (lane & 3);
      double b = 0.0;
      if (k < N && tvalid) {
        double wk = W[k];
        if (k > 0) wk = wk + wk;              // W carries Makhoul's 1/2 for k >= 1; the plain cosine sum does not want it
        b = em_modal(lam[k], wk, w, tt);
      }
      const double* af = afrag + ks * 32 + lane;
#pragma unroll
      for (int mt = 0; mt < MTMAX; ++mt)
        if (mt < MT) dmma_m8n8k4(d[mt][0], d[mt][1], af[mt * KS * 32], b);
    }
    // D fragment: lane holds sigma[n = 8 mt + lane/4][t = tb + 2 (lane%4) + {0, 1}]
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int tj = tb + 2 * (lane & 3) + e;
      double best = ninf;
      int besti = 0x7fffffff;
#pragma unroll
      for (int mt = 0; mt < MTMAX; ++mt) {
        const int n = 8 * mt + (lane >> 2);
        if (mt < MT && n < N && tj < j1) {
          const double v = d[mt][e];
          if (p.stress) __stcs(p.stress + noff * (long long)p.T + (long long)tj * N + n, v);
          if (v > best) { best = v; besti = n; }
        }
      }
#pragma unroll
      for (int o = 4; o < 32; o <<= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
        if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
      }
      const double first = __shfl_sync(0xffffffffu, d[0][e], lane & 3);   // n = 0 lives in lanes 0..3 of tile 0
      if ((lane >> 2) == 0 && tj < j1) {
        if (besti == 0x7fffffff) { besti = 0; best = first; }
        const long long o = (long long)line * p.T + tj;
        p.mx[o] = best; p.amx[o] = besti; p.flag[o] = (best > p.critical) ? 1 : 0;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Current histories (proton_em_analyze_history_host): the stress of a piecewise-constant current is the sum of the
// reference solutions of the current increments, each started at its switching time.  The increments are evaluated
// by the kernels above into a second field; these two kernels add it to the running total and take the per-(line, t)
// reductions of the total.  Both are plain HBM streams (24 and 8 bytes per eval).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) em_accumulate_kernel(double* __restrict__ total, const double* __restrict__ add,
                                                            long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) total[i] += add[i];
}

struct RowMaxParams {
  const double* stress;        // indexed by global node offset (node_off[line] * T + t * N)
  const long long* node_off;
  const int* num_nodes;
  long long line0;             // first line of this launch
  int T;
  double* mx; int* amx; unsigned char* flag;
  double critical;
};

__global__ void __launch_bounds__(128) em_rowmax_kernel(const RowMaxParams p) {
  const long long row = blockIdx.x;
  const long long line = p.line0 + row / p.T;
  const int tj = (int)(row % p.T);
  const int N = p.num_nodes[line];
  const double* src = p.stress + p.node_off[line] * (long long)p.T + (long long)tj * N;
  double best = -__longlong_as_double(0x7ff0000000000000LL);
  int besti = 0x7fffffff;
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    const double v = src[n];
    if (v > best) { best = v; besti = n; }
  }
  __shared__ double sv[4];
  __shared__ int si[4];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
    if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = besti; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int q = 1; q < 4; ++q)
      if (sv[q] > best || (sv[q] == best && si[q] < besti)) { best = sv[q]; besti = si[q]; }
    if (besti == 0x7fffffff) { besti = 0; best = src[0]; }       // all NaN: report the first value like the fused kernels
    const long long o = line * p.T + tj;
    p.mx[o] = best; p.amx[o] = besti; p.flag[o] = (best > p.critical) ? 1 : 0;
  }
}

// ---------------------------------------------------------------------------------------------
// Initial stress (no counterpart in the reference, whose kernel starts from zero stress,
// bin/EMtool_analytical@0xc68c).  With sigma(0) = sigma0 the discretised equation has the solution
//     sigma(t) = sigma0 + sum_k [ (W_k + Z_k) (e^{lambda_k a t} - 1) ] cos(..)  + t W_0,   Z_k = DCT-II(sigma0)_k / N,
// i.e. the kernels above evaluate it unchanged once Z is folded into the modal weights (k >= 1), and sigma0 itself is
// added to every timestamp's row afterwards.  Two kernels:
//   em_forward_sigma0_kernel  dense forward DCT-II of a field: W_k += (1/N) sum_n sigma0[n] cos(pi k (2n+1) / 2N), k >= 1.
//       One CTA per (line, slab of 4096 modes), one thread per 32 consecutive modes; the field streams through shared
//       memory in chunks fetched by bulk asynchronous copies (TMA, cp.async.bulk + mbarrier, two buffers: the next chunk
//       lands while this one is consumed).  Cosines: exact integer angle at every 64th node, in between two rotations
//       (along n and along k), so the cost is ~5 FP64 instructions per (mode, node) pair.  O(N^2) per line: a grid's
//       worth (sum N^2 ~ 1e11) takes tens of milliseconds, once per restart.
//   em_add_sigma0_kernel      stress[t][n] += sigma0[n] for the rows of a launch, with max / argmax / flag of the sums.
// ---------------------------------------------------------------------------------------------
constexpr int kS0Chunk = 2048;            // doubles per staged chunk (16 KB)
constexpr int kS0Modes = 32;              // modes per thread
constexpr int kS0Threads = 128;


struct Sigma0Params {
  const int* num_nodes;
  const long long* node_off;
  const long long* slab_off;     // [L+1] prefix sum of ceil(N / 4096): CTA -> (line, slab)
  long long num_lines;
  const double* sigma0;          // plan-owned copy, line l at node_off[l]; two doubles of slack at the end
  double* W;
};

__global__ void __launch_bounds__(kS0Threads) em_forward_sigma0_kernel(const Sigma0Params p) {
  __shared__ __align__(16) double buf[2][kS0Chunk + 2];
  __shared__ __align__(8) unsigned long long bars[2];
  const long long line = find_line(p.slab_off, p.num_lines, (long long)blockIdx.x);
  const int slab = (int)((long long)blockIdx.x - p.slab_off[line]);
  const int N = p.num_nodes[line];
  const long long noff = p.node_off[line];
  const int tid = threadIdx.x;
  const int k0 = (slab * kS0Threads + tid) * kS0Modes;
  // chunks start on even global elements so that every bulk copy is 16-byte aligned: element `shift` of the first
  // buffer is node 0 of the line
  const int shift = (int)(noff & 1);
  const double* src0 = p.sigma0 + (noff - shift);
  const int total = N + shift;                                   // doubles from src0 to the end of the line
  const int nchunks = (total + kS0Chunk - 1) / kS0Chunk;
  const unsigned bar0 = smem_u32(&bars[0]), bar1 = smem_u32(&bars[1]);
  auto issue = [&](int c) {
    const int len = min(kS0Chunk, total - c * kS0Chunk);
    const unsigned bytes = (unsigned)(((len + 1) & ~1) * sizeof(double));      // (the slack behind the array covers the pad)
    const unsigned bar = (c & 1) ? bar1 : bar0;
    mbar_expect_tx(bar, bytes);
    bulk_g2s(smem_u32(&buf[c & 1][0]), src0 + (long long)c * kS0Chunk, bytes, bar);
  };
  if (tid == 0) {
    mbar_init(bar0, 1);
    mbar_init(bar1, 1);
    mbar_init_fence();
  }
  __syncthreads();
  if (tid == 0) { issue(0); if (nchunks > 1) issue(1); }
  double R[kS0Modes];
#pragma unroll
  for (int q = 0; q < kS0Modes; ++q) R[q] = 0.0;
  const bool live = k0 < N;
  const unsigned long long den = 2ULL * (unsigned long long)N;
  // per-thread constants: step of the k0-angle and of the base angle when n advances by one
  const cplx dz = em_cis_frac((2ULL * (unsigned long long)k0) % (2ULL * den), den);     // e^{i pi 2 k0 / 2N}
  const cplx dw = em_cis_frac(2ULL, den);                                               // e^{i pi 2 / 2N}
  cplx z = cplx{1.0, 0.0}, w = cplx{1.0, 0.0};
  for (int c = 0; c < nchunks; ++c) {
    mbar_wait((c & 1) ? bar1 : bar0, (unsigned)((c >> 1) & 1));
    const double* b = buf[c & 1];
    const int e0 = c * kS0Chunk;                                  // first element (relative to src0) of this chunk
    const int len = min(kS0Chunk, total - e0);
    if (live) {
      for (int e = (c == 0 ? shift : 0); e < len; ++e) {
        const int n = e0 + e - shift;                             // node index
        if ((n & 63) == 0) {                                      // exact angles every 64 nodes
          const unsigned long long m = 2ULL * (unsigned long long)n + 1ULL;
          w = em_cis_frac(m, den);
          z = em_cis_frac((m * (unsigned long long)k0) % (2ULL * den), den);
        }
        const double x = b[e];
        cplx zz = z;
#pragma unroll
        for (int q = 0; q < kS0Modes; ++q) {
          R[q] = fma(x, zz.x, R[q]);
          zz = cmul(zz, w);
        }
        z = cmul(z, dz);
        w = cmul(w, dw);
      }
    }
    __syncthreads();                                              // everyone is done with this buffer
    if (tid == 0 && c + 2 < nchunks) issue(c + 2);
  }
  if (live) {
    double* W = p.W + noff + k0;
#pragma unroll
    for (int q = 0; q < kS0Modes; ++q)
      if (k0 + q < N && k0 + q >= 1) W[q] = W[q] + R[q] / (double)N;
  }
}

struct AddSigma0Params {
  double* stress;              // indexed by global node offset (node_off[line] * T + t * N)
  const double* sigma0;
  const long long* node_off;
  const int* num_nodes;
  long long line0;             // first line of this launch
  int T;
  double* mx; int* amx; unsigned char* flag;
  double critical;
};

__global__ void __launch_bounds__(128) em_add_sigma0_kernel(const AddSigma0Params p) {
  const long long row = blockIdx.x;
  const long long line = p.line0 + row / p.T;
  const int tj = (int)(row % p.T);
  const int N = p.num_nodes[line];
  double* dst = p.stress + p.node_off[line] * (long long)p.T + (long long)tj * N;
  const double* s0 = p.sigma0 + p.node_off[line];
  double best = -__longlong_as_double(0x7ff0000000000000LL);
  int besti = 0x7fffffff;
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    const double v = dst[n] + s0[n];
    dst[n] = v;
    if (v > best) { best = v; besti = n; }
  }
  __shared__ double sv[4];
  __shared__ int si[4];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
    if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = besti; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int q = 1; q < 4; ++q)
      if (sv[q] > best || (sv[q] == best && si[q] < besti)) { best = sv[q]; besti = si[q]; }
    if (besti == 0x7fffffff) { besti = 0; best = dst[0]; }
    const long long o = line * p.T + tj;
    p.mx[o] = best; p.amx[o] = besti; p.flag[o] = (best > p.critical) ? 1 : 0;
  }
}

}  // namespace em
