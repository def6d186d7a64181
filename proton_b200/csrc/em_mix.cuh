// em_mix.cuh — in-CTA Bluestein with mixed-radix lengths P = ODD x 2^LR (ODD = 5 ... 16 rows of 2^LR points).
//
// em_eval_fft_kernel<p, true> pads every line to the next power of two P >= 2N-1, so a line just above a power of two
// does twice the necessary transform length.  With 5 and 6 rows (1.25 and 1.5 x 2^k) in between, the average excess over
// 2N dropped from 44 % to 19 % for line lengths spread evenly in log N (synthetic ragged batch 66.0 -> 57.7 ms per step on
// B200, T = 256); with every row count up to 16 that pays (rows of 256 and 512 points) it is 16 % overall and 7 % from
// 640 nodes up (52.3 -> 49.2 ms).
//
// Structure: the long path's three passes (em_long.cuh) inside one CTA, with shared memory in place of the L2 scratch.
// The P points are ODD rows of R = 2^LR points, element index i = j + q R (row q, column j):
//   modal phase   chirped, packed spectrum a_k written at (row k / R, column k % R)               [em_fft.cuh phase 1]
//   column pass   y_r[j] = (sum_q x[j + q R] w_ODD^{-q r}) w_P^{-j r}     (inputs with j + q R >= N are zero)
//   row pass      R-point DIF on the ODD rows at once (the rows play the part of em_fft.cuh's transforms in flight),
//                 last DIF stage x H x first DIT stage in registers (H stored row by row, lane-major inside a row: hh_pos),
//                 DIT back down to stage 0
//   column pass   u[j + q R] = sum_r (X_r[j] w_P^{+j r}) w_ODD^{+q r}, post-chirp, de-interleave (Makhoul's permutation),
//                 stress store straight to global memory, max / argmax / flag per (line, timestamp)
// One CTA owns one (line, range of timestamp pairs) item and runs its transforms one after the other.
#pragma once
#include "em_kernels.cuh"
#include "em_long.cuh"

namespace em {

// Row counts.  5 and 6 rows (P = 1.25 and 1.5 x 2^k; six half-length rows rather than three: an even row count splits
// evenly over the row groups a CTA works on at a time, measured 63 vs 42 Gevals/s at N = 3000) came first; the column
// DFT handles any row count up to 16 in registers, so rows of 256 points give every multiple of 256 up to 4096 and rows
// of 512 points every multiple of 512 up to 8192: the excess of P over 2N drops to a few per cent, and rows of 512
// points need one stage less than the 1024-point rows the 5120- and 6144-point classes started with.
template <int LR, int ODD> struct MixCfg {
  static constexpr int R = 1 << LR, P = ODD * R;
  static constexpr int NT = P > 6144 ? 512 : P > 3072 ? 256 : P >= 1024 ? 128 : 64;
  // resident threads per SM the register budget is sized for: 512 (128 registers) where the rows run radix-16 stages,
  // the column DFT holds nine or more values or the rows are long (shared memory allows no more CTAs there anyway),
  // 768 (80 registers) for the short ones (measured +3..13 % there, neutral for LR = 9)
  static constexpr int MINB = ((plan16(LR) || LR >= 9 || ODD >= 9) ? 512 : 768) / NT;
  static constexpr int RSTRIDE = padded_len<LR>();                    // padded row length (complex elements)
  static constexpr size_t tw_bytes = (size_t)stage_tw_total(LR) * sizeof(cplx);
  static constexpr size_t smem_bytes = tw_bytes + (size_t)ODD * RSTRIDE * sizeof(cplx);
  static constexpr int QN = (ODD + 1) / 2;                             // rows that can hold outputs m < N <= (P + 1) / 2
};

// code of a mixed class as the host keeps it in ClassOf::log2p / NTab::mix: (LR - 6) * 32 + ODD
__host__ __device__ constexpr int mix_code(int lr, int odd) { return (lr - 6) * 32 + odd; }

// Middle of Bluestein on ODD rows: last DIF stage, product with this row's slice of H, first DIT stage, in registers.
template <int LR, int RADIX, int NT, int ODD>
__device__ __forceinline__ void mix_middle(cplx* __restrict__ data, const cplx* __restrict__ hhat, int tid) {
  constexpr int R = 1 << LR, NB = R / RADIX, RS = padded_len<LR>();
  constexpr int BSTEP = NB < NT ? NB : NT;
  constexpr int GSTEP = NB < NT ? NT / NB : 1;
  const int g0 = NB < NT ? tid / NB : 0;
  for (int b = (NB < NT ? tid % NB : tid); b < NB; b += BSTEP) {
    const int base = b * RADIX;
    const int pbase = pad_idx<LR>(base);
    for (int g = g0; g < ODD; g += GSTEP) {
      cplx* d = data + g * RS + pbase;
      const cplx* h = hhat + g * R;
      cplx v[RADIX];
#pragma unroll
      for (int q = 0; q < RADIX; ++q) v[q] = d[q];
      Dft<RADIX, -1>::run(v);
#pragma unroll
      for (int q = 0; q < RADIX; ++q) v[q] = cmul(v[q], ldg_cplx(h + hh_pos<RADIX, NB>(b, q)));
      Dft<RADIX, +1>::run(v);
#pragma unroll
      for (int q = 0; q < RADIX; ++q) d[q] = v[q];
    }
  }
}

// Forward column pass over whatever is in the rows (valid inputs: element index < n_valid, or everything).
template <int LR, int ODD, int NT, bool PRUNED>
__device__ __forceinline__ void mix_colfwd(cplx* __restrict__ data, const cplx* __restrict__ twl, int tid, int n_valid) {
  constexpr int R = 1 << LR, RS = padded_len<LR>();
  for (int j = tid; j < R; j += NT) {
    const int pj = pad_idx<LR>(j);
    cplx v[ODD];
#pragma unroll
    for (int q = 0; q < ODD; ++q) v[q] = (!PRUNED || j + q * R < n_valid) ? data[q * RS + pj] : cplx{0.0, 0.0};
    DftAny<ODD, -1>::run(v);
    const cplx t = ldg_cplx(twl + j);
    col_twiddle<ODD>(v, cplx{t.x, -t.y});
#pragma unroll
    for (int r = 0; r < ODD; ++r) data[r * RS + pj] = v[r];
  }
}

struct MixParams {
  EvalParams e;
  const cplx* twl;         // [R]  w_P^{+j}
};

template <int LR, int ODD>
__global__ void __launch_bounds__(MixCfg<LR, ODD>::NT, MixCfg<LR, ODD>::MINB) em_eval_mix_kernel(const MixParams mp) {
  using C = MixCfg<LR, ODD>;
  constexpr int R = C::R, NT = C::NT, RS = C::RSTRIDE, NST = plan_n(LR), QN = C::QN;
  const EvalParams& prm = mp.e;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* twb = reinterpret_cast<cplx*>(smem_raw);
  cplx* data = reinterpret_cast<cplx*>(smem_raw + C::tw_bytes);
  __shared__ double red_v[2][NT / 32];
  __shared__ int red_i[2][NT / 32];
  const int tid = threadIdx.x;
  for (int i = tid; i < stage_tw_total(LR); i += NT) twb[i] = prm.tw[i];

  const int item = blockIdx.x;
  const int li = item / prm.chunks, chunk = item - li * prm.chunks;
  const int line = prm.class_lines[li];
  const int N = prm.num_nodes[line];
  const double a = prm.a_coeff[line];
  const long long noff = prm.node_off[line];
  const NTab tab = prm.tabs[prm.line_tab[line]];
  const double* __restrict__ lam = tab.lam;
  const double* __restrict__ W = prm.W + noff;
  const int T = prm.T;
  const int npairs = (T + 1) >> 1;
  const int pair_begin = chunk * prm.pairs_per_chunk;
  const int pair_end = min(npairs, pair_begin + prm.pairs_per_chunk);
  const int halfN = (N + 1) >> 1;
  const int lane = tid & 31, warp = tid >> 5;
  const double ninf = -__longlong_as_double(0x7ff0000000000000LL);
  __syncthreads();

  for (int pb = pair_begin; pb < pair_end; ++pb) {
    const int j1 = 2 * pb;
    const double t1 = prm.ts[j1];
    const double t2 = (j1 + 1 < T) ? prm.ts[j1 + 1] : t1;
    const double w1 = __dmul_rn(a, t1), w2 = __dmul_rn(a, t2);
    // ---- modal values -> chirped packed spectrum at its natural position (row k / R, column k % R) ----------------
    {
      const int KP = (N - 1) >> 1;
      // software pipeline: the loads of the next mode pair are issued before the four exponentials of this one
      // (the tables of a line exceed what is left of L1 beside the rows, so each pass over them pays L2 latency: +3 %)
      int k = tid + 1;
      double lamk = 0.0, wk = 0.0, lamkk = 0.0, wkk = 0.0;
      cplx prek = cplx{0.0, 0.0}, prekk = cplx{0.0, 0.0};
      if (k <= KP) {
        lamk = __ldg(lam + k); wk = __ldg(W + k); lamkk = __ldg(lam + N - k); wkk = __ldg(W + N - k);
        prek = ldg_cplx(tab.pre + k); prekk = ldg_cplx(tab.pre + N - k);
      }
      for (; k <= KP; k += NT) {
        const int kk = N - k, kn = k + NT;
        const double l0 = lamk, w0 = wk, l1 = lamkk, ww1 = wkk;
        const cplx p0 = prek, pp1 = prekk;
        if (kn <= KP) {
          lamk = __ldg(lam + kn); wk = __ldg(W + kn); lamkk = __ldg(lam + N - kn); wkk = __ldg(W + N - kn);
          prek = ldg_cplx(tab.pre + kn); prekk = ldg_cplx(tab.pre + N - kn);
        }
        const double y1k = em_modal(l0, w0, w1, t1), y2k = em_modal(l0, w0, w2, t2);
        const double y1kk = em_modal(l1, ww1, w1, t1), y2kk = em_modal(l1, ww1, w2, t2);
        data[(k >> LR) * RS + pad_idx<LR>(k & (R - 1))] = cmul(p0, cplx{y1k + y2kk, y2k - y1kk});
        data[(kk >> LR) * RS + pad_idx<LR>(kk & (R - 1))] = cmul(pp1, cplx{y1kk + y2k, y2kk - y1k});
      }
      if (tid < 2) {
        const int k = tid ? (N >> 1) : 0;
        if (tid == 0 || (((N & 1) == 0) && N > 1)) {
          const double y1 = em_modal(lam[k], W[k], w1, t1), y2 = em_modal(lam[k], W[k], w2, t2);
          data[(k >> LR) * RS + pad_idx<LR>(k & (R - 1))] =
              tid ? cmul(ldg_cplx(tab.pre + k), cplx{y1 + y2, y2 - y1}) : cplx{y1, y2};
        }
      }
    }
    __syncthreads();
    // ---- forward: column pass, row DIF, x H, row DIT ------------------------------------------------------------
    mix_colfwd<LR, ODD, NT, true>(data, mp.twl, tid, N);
    __syncthreads();
    DifStages<LR, 0, NST - 1, NT>::run(data, ODD, twb, tid);
    mix_middle<LR, stage_radix<LR, NST - 1>(), NT, ODD>(data, tab.hhat, tid);
    stage_barrier<LR, NST - 2>();
    DitStages<LR, NST - 2, 0, NT>::run(data, ODD, twb, tid);
    // ---- inverse column pass + post-chirp + de-interleave + stress store + max -----------------------------------
    double* d1 = prm.stress ? prm.stress + noff * (long long)T + (long long)j1 * N : nullptr;
    double* d2 = (prm.stress && j1 + 1 < T) ? d1 + N : nullptr;
    double b1 = ninf, b2 = ninf;
    int i1 = 0x7fffffff, i2 = 0x7fffffff;
    for (int j = tid; j < R; j += NT) {
      const int pj = pad_idx<LR>(j);
      cplx v[ODD];
#pragma unroll
      for (int r = 0; r < ODD; ++r) v[r] = data[r * RS + pj];
      col_twiddle<ODD>(v, ldg_cplx(mp.twl + j));
      DftAny<ODD, +1>::run(v);
#pragma unroll
      for (int q = 0; q < QN; ++q) {
        const int m = j + q * R;
        if (m < N) {
          const cplx u = cmul(v[q], ldg_cplx(tab.post + m));
          const int idx = (m < halfN) ? 2 * m : 2 * (N - 1 - m) + 1;
          if (d1) __stcs(d1 + idx, u.x);
          if (d2) __stcs(d2 + idx, u.y);
          if (u.x > b1 || (u.x == b1 && idx < i1)) { b1 = u.x; i1 = idx; }
          if (u.y > b2 || (u.y == b2 && idx < i2)) { b2 = u.y; i2 = idx; }
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ob1 = __shfl_xor_sync(0xffffffffu, b1, o); const int oi1 = __shfl_xor_sync(0xffffffffu, i1, o);
      if (ob1 > b1 || (ob1 == b1 && oi1 < i1)) { b1 = ob1; i1 = oi1; }
      const double ob2 = __shfl_xor_sync(0xffffffffu, b2, o); const int oi2 = __shfl_xor_sync(0xffffffffu, i2, o);
      if (ob2 > b2 || (ob2 == b2 && oi2 < i2)) { b2 = ob2; i2 = oi2; }
    }
    if (lane == 0) { red_v[0][warp] = b1; red_i[0][warp] = i1; red_v[1][warp] = b2; red_i[1][warp] = i2; }
    __syncthreads();                        // partials in place; every thread is done with the rows
    if (tid < 2 && j1 + tid < T) {
      double b = red_v[tid][0]; int bi = red_i[tid][0];
      for (int q = 1; q < NT / 32; ++q) {
        const double ob = red_v[tid][q]; const int oi = red_i[tid][q];
        if (ob > b || (ob == b && oi < bi)) { b = ob; bi = oi; }
      }
      if (bi == 0x7fffffff) {               // every value NaN: report node 0 like the other kernels
        bi = 0;
        b = __longlong_as_double(0x7ff8000000000000LL);
      }
      const long long o = (long long)line * T + j1 + tid;
      prm.mx[o] = b; prm.amx[o] = bi; prm.flag[o] = (b > prm.critical) ? 1 : 0;
    }
    // (the next pair's modal phase writes the rows; red_* is rewritten only after its own barriers)
  }
}

// ---- tables of a mixed class: pre / post chirps and H = forward transform of the conjugate chirp / P -----------------
template <int LR, int ODD>
__global__ void __launch_bounds__(MixCfg<LR, ODD>::NT) em_tables_mix_kernel(const TableParams p, const cplx* twl) {
  using C = MixCfg<LR, ODD>;
  constexpr int R = C::R, P = C::P, NT = C::NT, RS = C::RSTRIDE;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* twb = reinterpret_cast<cplx*>(smem_raw);
  cplx* data = reinterpret_cast<cplx*>(smem_raw + C::tw_bytes);
  const NTab tab = p.tabs[blockIdx.x];
  const unsigned long long N = (unsigned long long)tab.n;
  cplx* pre = const_cast<cplx*>(tab.pre);
  cplx* post = const_cast<cplx*>(tab.post);
  cplx* hhat = const_cast<cplx*>(tab.hhat);
  const int tid = threadIdx.x;
  for (int i = tid; i < stage_tw_total(LR); i += NT) twb[i] = p.tw[i];
  for (int k = tid; k < (int)N; k += NT) {
    const unsigned long long kk = (unsigned long long)k;
    pre[k] = em_cis_frac((kk + 2ULL * kk * kk) % (4ULL * N), 2ULL * N);
    post[k] = em_cis_frac((kk * kk) % (2ULL * N), N);
  }
  for (int i = tid; i < P; i += NT) data[(i >> LR) * RS + pad_idx<LR>(i & (R - 1))] = cplx{0.0, 0.0};
  __syncthreads();
  const double inv_p = 1.0 / (double)P;
  for (int j = tid; j < (int)N; j += NT) {
    const unsigned long long jj = (unsigned long long)j;
    const cplx c = em_cis_frac((jj * jj) % (2ULL * N), N);
    const cplx h = cplx{c.x * inv_p, -c.y * inv_p};
    data[(j >> LR) * RS + pad_idx<LR>(j & (R - 1))] = h;
    if (j) { const int i = P - j; data[(i >> LR) * RS + pad_idx<LR>(i & (R - 1))] = h; }
  }
  __syncthreads();
  mix_colfwd<LR, ODD, NT, false>(data, twl, tid, 0);
  __syncthreads();
  DifStages<LR, 0, plan_n(LR), NT>::run(data, ODD, twb, tid);
  for (int i = tid; i < P; i += NT)
    hhat[(i >> LR) * R + hh_pos_of(R, plan_radix(LR, plan_n(LR) - 1), i & (R - 1))] = data[(i >> LR) * RS + pad_idx<LR>(i & (R - 1))];
}

// w_P^{+j}, j < R, of a mixed length (one table per class, built at first use)
__global__ void __launch_bounds__(256) em_mix_twiddle_kernel(cplx* twl, int R, int P) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < R) twl[j] = em_cis_frac(2ULL * (unsigned long long)j, (unsigned long long)P);
}

}  // namespace em
