// em_long.cuh — long lines: transforms that do not fit one CTA's shared memory (P = P1 * 4096 complex
// points, P1 in {3..16, 18, 20, .., 32}; Bluestein lines with 4096 < N <= 65536, power-of-two lines up to 131072).
//
// Same mathematics as em_fft.cuh (two timestamps per complex transform, Makhoul permutation, Bluestein
// with a DIF forward / DIT inverse pair so no reordering pass exists), factored as P = P1 x P2 with
// P2 = 4096:  element index i = j + q*P2  (column j, row q).
//   forward  : column pass  y_r[j] = (sum_q x[j + q P2] w_P1^{-qr}) w_P^{-jr}      -> X[r*P2 + j]
//              row pass     P2-point DIF on every row, product with H, P2-point DIT   (one CTA per row,
//                           the in-CTA stages of em_fft.cuh)
//   inverse  : column pass  u[j + q P2] = sum_r (X[r*P2 + j] w_P^{+jr}) w_P1^{+qr}
// The passes exchange data through a scratch buffer in global memory that is sized to stay resident in
// the 126 MB L2 (a wave of transforms at a time).  Kernels:
//   em_long_colfwd_paired_kernel  modal factors + pre-chirp + forward column pass (Bluestein evaluation)
//   em_long_colfwd_kernel     forward column pass over the chirp (table build, once per distinct N)
//   em_long_scatter_kernel    power-of-two lines: modal factors, U_k written where the row DIT expects it
//   em_long_row_kernel        row pass (forward * H * inverse, forward only, or inverse only)
//   em_long_colinv_kernel     inverse column pass + post-chirp + de-interleave + stress store + partial max
//   em_long_reduce_kernel     per-(line, timestamp) max / argmax / flag from the partials
#pragma once
#include "em_fft.cuh"

namespace em {

constexpr int kLongLog2P2 = 12;
constexpr int kLongP2 = 1 << kLongLog2P2;
constexpr int kLongColsPerCta = 128;
constexpr int kLongRowThreads = plan16(kLongLog2P2) ? 256 : 512;     // one butterfly of the outermost radix per thread

// Twiddles of the small column DFTs whose length is not a power of two: kDftTw[P1][m] = e^{+2 pi i m / P1}
// (filled by the host at context creation).  Indices are compile-time constants after unrolling, so the values
// are read as constant-bank operands.
constexpr int kLongMaxP1 = 32;
__constant__ cplx kDftTw[kLongMaxP1 + 1][kLongMaxP1];

// Column DFT of any length R <= 32: radix-2 recursion for powers of two and even lengths, the symmetric O(R^2 / 4)
// form for odd lengths.
template <int R, int SIGN> struct DftAny {
  static __device__ __forceinline__ void run(cplx (&v)[R]) {
    if constexpr ((R & (R - 1)) == 0) {
      DftBig<R, SIGN>::run(v);
    } else if constexpr (R % 2 == 0) {
      // even length: one radix-2 split on top of two DFTs of length R/2
      cplx e[R / 2], o[R / 2];
#pragma unroll
      for (int i = 0; i < R / 2; ++i) { e[i] = v[2 * i]; o[i] = v[2 * i + 1]; }
      DftAny<R / 2, SIGN>::run(e);
      DftAny<R / 2, SIGN>::run(o);
#pragma unroll
      for (int r = 0; r < R / 2; ++r) {
        cplx t = o[r];
        if (r != 0) {
          const cplx w = kDftTw[R][r];
          const double wy = SIGN > 0 ? w.y : -w.y;
          t = cplx{fma(o[r].x, w.x, -(o[r].y * wy)), fma(o[r].x, wy, o[r].y * w.x)};
        }
        v[r] = cadd(e[r], t);
        v[r + R / 2] = csub(e[r], t);
      }
    } else {
      // odd length: pair q with R-q.  e^{i t} v_q + e^{-i t} v_{R-q} = cos t (v_q + v_{R-q}) + i sin t (v_q - v_{R-q}), so
      // outputs r and R-r share A_r = sum_q cos(2 pi q r / R) s_q and B_r = sum_q sin(2 pi q r / R) d_q: (R-1)^2 real
      // FMAs instead of 4 R (R-1).
      constexpr int H = (R - 1) / 2;
      cplx sm[H + 1], df[H + 1];
      cplx y0 = v[0];
#pragma unroll
      for (int q = 1; q <= H; ++q) {
        sm[q] = cadd(v[q], v[R - q]);
        df[q] = csub(v[q], v[R - q]);
        y0 = cadd(y0, sm[q]);
      }
      cplx y[R];
      y[0] = y0;
#pragma unroll
      for (int r = 1; r <= H; ++r) {
        cplx A = v[0], B = cplx{0.0, 0.0};
#pragma unroll
        for (int q = 1; q <= H; ++q) {
          const cplx w = kDftTw[R][(q * r) % R];          // (cos, sin) of 2 pi q r / R
          A.x = fma(w.x, sm[q].x, A.x); A.y = fma(w.x, sm[q].y, A.y);
          B.x = fma(w.y, df[q].x, B.x); B.y = fma(w.y, df[q].y, B.y);
        }
        constexpr double sg = (double)SIGN;
        y[r] = cplx{A.x - sg * B.y, A.y + sg * B.x};
        y[R - r] = cplx{A.x + sg * B.y, A.y - sg * B.x};
      }
#pragma unroll
      for (int r = 0; r < R; ++r) v[r] = y[r];
    }
  }
};

// Programmatic dependent launch (PDL): the three kernels of a wave run back to back on one stream.  Each lets the
// next one start its CTAs as soon as all of its own are resident (pdl_launch_dependents at entry) and blocks only
// where it first touches what its predecessor produced or still reads (pdl_wait).  Without the launch attribute both
// instructions are no-ops.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// v[r] *= w^r for r = 1..R-1 (|w| = 1): w^2 by squaring, then two interleaved chains w^{r+2} = w^r w^2, so the
// rounding error grows like r/2 ulp and no table of R twiddles per column has to come from memory.
template <int R> __device__ __forceinline__ void col_twiddle(cplx (&v)[R], cplx w) {
  if constexpr (R >= 2) v[1] = cmul(v[1], w);
  if constexpr (R >= 3) {
    const cplx w2 = cmul(w, w);
    cplx a = w2, b = w;              // a = w^{2m}, b = w^{2m-1}
    v[2] = cmul(v[2], a);
#pragma unroll
    for (int r = 3; r < R; ++r) {
      if (r & 1) { b = cmul(b, w2); v[r] = cmul(v[r], b); }
      else       { a = cmul(a, w2); v[r] = cmul(v[r], a); }
    }
  }
}

// Everything a long-path kernel needs to know about one transform slot, in one 80-byte record that the host builds
// once per (plan, T): the kernels used to walk slot -> line -> table index -> table -> pointer -> data, five dependent
// loads from L2 before the first useful one (24 % of the inverse column kernel's stall samples sat on the post-chirp
// loads at the end of that chain, profiles/r01_ncu_long_full.txt); now it is slot -> record -> data.
struct alignas(16) SlotRec {
  const double* lam;       // lambda_k of the line's length
  const double* W;         // modal weights of the line
  const cplx* pre;         // pre-chirp / pre-twiddle
  const cplx* post;        // post-chirp (Bluestein)
  const cplx* hhat;        // transformed chirp (Bluestein)
  long long stress_off;    // node_off[line] * T + 2 * pair * N: first timestamp's row in the stress field
  double a;                // A_coeff of the line
  int N;
  int line;
  int pair;
  int reserved;
};

struct LongWave {
  // transforms of this wave: slot s -> record
  const SlotRec* recs;
  int n_tr;
  int tr_base;           // index of slot 0 among all transforms of the class (partials are per class)
  int T;
  const double* ts;
  cplx* x;               // [n_tr][P]
  double* stress;
  double* part_v;        // [n_tr][2][colblocks]
  int* part_i;
  double* mx; int* amx; unsigned char* flag;
  double critical;
};

// ---- forward column pass ---------------------------------------------------------------------------------
// Parameters shared by the column kernels.  `mode` / `tabs_build` are used by the table build only
// (em_long_colfwd_kernel: x = conj chirp / P, one "transform" per entry of `tabs_build`).
struct ColParams {
  LongWave w;
  int mode;
  const NTab* tabs_build;
  int tr_per_cta;
  const cplx* twl;       // [P1][P2]  w_P^{+j r}  (em_long_twiddle_kernel); the evaluation kernels read row r = 1 only
};

__global__ void __launch_bounds__(256) em_long_twiddle_kernel(cplx* twl, int P1) {
  const long long P = (long long)P1 * kLongP2;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P1 * kLongP2) return;
  const int r = idx / kLongP2, j = idx - r * kLongP2;
  twl[idx] = em_cis_frac((unsigned long long)j * r * 2ULL, (unsigned long long)P);
}

// pre / post chirps of a long Bluestein line (the in-CTA classes get theirs from em_tables_kernel).
__global__ void __launch_bounds__(256) em_long_prepost_kernel(const NTab* tabs_build, int bluestein) {
  const NTab tab = tabs_build[blockIdx.y];
  const unsigned long long N = (unsigned long long)tab.n;
  cplx* pre = const_cast<cplx*>(tab.pre);
  cplx* post = const_cast<cplx*>(tab.post);
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < (int)N; k += gridDim.x * blockDim.x) {
    const unsigned long long kk = (unsigned long long)k;
    if (bluestein) {
      pre[k] = em_cis_frac((kk + 2ULL * kk * kk) % (4ULL * N), 2ULL * N);
      post[k] = em_cis_frac((kk * kk) % (2ULL * N), N);
    } else {
      pre[k] = em_cis_frac(kk, 2ULL * N);
    }
  }
}

// Table build: forward column pass over the conjugate chirp of each new length (run once per distinct N).
template <int P1>
__global__ void __launch_bounds__(kLongColsPerCta) em_long_colfwd_kernel(const ColParams p) {
  constexpr int P2 = kLongP2;
  constexpr long long P = (long long)P1 * P2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* twt = reinterpret_cast<cplx*>(smem_raw);          // [P1][cols]  w_P^{-j r}
  const int j = blockIdx.x * kLongColsPerCta + threadIdx.x;
  for (int r = 1; r < P1; ++r) {
    const cplx c = ldg_cplx(p.twl + (long long)r * P2 + j);
    twt[r * kLongColsPerCta + threadIdx.x] = cplx{c.x, -c.y};
  }
  const int s0 = blockIdx.y * p.tr_per_cta;
  const int s1 = min(p.w.n_tr, s0 + p.tr_per_cta);
  for (int s = s0; s < s1; ++s) {
    cplx v[P1];
    const NTab tab = p.tabs_build[s];
    const int N = tab.n;
    const double inv_p = 1.0 / (double)P;
#pragma unroll
    for (int q = 0; q < P1; ++q) {
      const long long i = j + (long long)q * P2;
      long long d = -1;
      if (i < N) d = i; else if (P - i < N) d = P - i;
      v[q] = cplx{0.0, 0.0};
      if (d >= 0) {
        const unsigned long long dd = (unsigned long long)d;
        const cplx c = em_cis_frac((dd * dd) % (2ULL * N), (unsigned long long)N);
        v[q] = cplx{c.x * inv_p, -c.y * inv_p};
      }
    }
    DftAny<P1, -1>::run(v);
    cplx* x = p.w.x + (long long)s * P;
    x[j] = v[0];
#pragma unroll
    for (int r = 1; r < P1; ++r) x[(long long)r * P2 + j] = cmul(v[r], twt[r * kLongColsPerCta + threadIdx.x]);
  }
}

// Forward column pass of the evaluation with every modal factor evaluated ONCE: entry k of the packed
// spectrum needs the modal values of k and of its mirror N-k, which lives in column (N mod 4096) - j.  A CTA therefore
// owns 64 column PAIRS {j, mirror(j)} (threads 0..63 one side, 64..127 the other), each thread evaluates only its own
// modes and the two sides swap through shared memory.  Slot a pairs columns c/2 + a and c/2 - a (c = N mod 4096; for
// even c the two self-mirrored columns c/2 and c/2 + 2048 stand alone), so the grid has 33 blocks of 64 slots.
constexpr int kLongPairBlocks = kLongP2 / 128 + 1;

template <int P1>
__global__ void __launch_bounds__(128) em_long_colfwd_paired_kernel(const ColParams p) {
  constexpr int P2 = kLongP2, QMAX = (P1 + 1) / 2 + 1;
  constexpr long long P = (long long)P1 * P2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* ys = reinterpret_cast<cplx*>(smem_raw);           // [QMAX][128] modal pairs (Yhat1, Yhat2) of this thread's modes
  cplx wbase = cplx{1.0, 0.0};                            // w_P^{-j} of this thread's column
  const int tid = threadIdx.x;
  pdl_launch_dependents();
  const int a = blockIdx.x * 64 + (tid & 63);
  const int side = tid >> 6;
  int prev_c = -1, j = 0, partner = tid;
  bool active = false;
  const int s0 = blockIdx.y * p.tr_per_cta;
  const int s1 = min(p.w.n_tr, s0 + p.tr_per_cta);
  for (int s = s0; s < s1; ++s) {
    const SlotRec rec = p.w.recs[s];
    const int line = rec.line, pr = rec.pair;
    const int N = rec.N;
    const int c = N & (P2 - 1);
    if (c != prev_c) {
      prev_c = c;
      if (c & 1) {
        active = a < P2 / 2;
        j = side == 0 ? ((c + 1) / 2 + a) & (P2 - 1) : ((c - 1) / 2 - a) & (P2 - 1);
        partner = tid ^ 64;
      } else {
        const bool self = (a == 0) || (a == P2 / 2);
        active = a <= P2 / 2 && !(self && side == 1);
        j = side == 0 ? (c / 2 + a) & (P2 - 1) : (c / 2 - a) & (P2 - 1);
        partner = self ? tid : (tid ^ 64);
      }
      if (active) {
        const cplx t = ldg_cplx(p.twl + P2 + j);          // row r = 1 of the table: w_P^{+j}
        wbase = cplx{t.x, -t.y};
      }
    }
    const double* W = rec.W;
    const double av = rec.a;
    const double t1 = p.w.ts[2 * pr];
    const double t2 = (2 * pr + 1 < p.w.T) ? p.w.ts[2 * pr + 1] : t1;
    const double w1 = __dmul_rn(av, t1), w2 = __dmul_rn(av, t2);
    cplx yk[QMAX];
    double lq[QMAX], wq[QMAX];
#pragma unroll
    for (int q = 0; q < QMAX; ++q) {              // all loads first, then the arithmetic: written as one loop, the
                                                  // compiler issued each mode's loads right before its exp (-3 % per step)
      const int k = j + q * P2;
      lq[q] = 0.0; wq[q] = 0.0;
      if (active && q < P1 && k < N) { lq[q] = __ldg(rec.lam + k); wq[q] = __ldg(W + k); }
    }
#pragma unroll
    for (int q = 0; q < QMAX; ++q) {
      const int k = j + q * P2;
      yk[q] = cplx{0.0, 0.0};
      if (active && q < P1 && k < N) yk[q] = cplx{em_modal(lq[q], wq[q], w1, t1), em_modal(lq[q], wq[q], w2, t2)};
      ys[q * 128 + tid] = yk[q];
    }
    __syncthreads();
    cplx v[P1];
#pragma unroll
    for (int q = 0; q < P1; ++q) {
      v[q] = cplx{0.0, 0.0};
      if (q < QMAX) {
        const int k = j + q * P2;
        if (active && k < N) {
          const int kk = N - k;                      // mirror mode; kk == N (k == 0) has no partner
          cplx ym = cplx{0.0, 0.0};
          if (kk < N) ym = ys[(kk >> kLongLog2P2) * 128 + partner];
          const cplx z = cplx{yk[q].x + ym.y, yk[q].y - ym.x};
          v[q] = k == 0 ? z : cmul(ldg_cplx(rec.pre + k), z);
        }
      }
    }
    __syncthreads();
    if (s == s0) pdl_wait();       // the scratch may still be read by the previous wave's inverse column pass
    if (active) {
      DftAny<P1, -1>::run(v);
      col_twiddle<P1>(v, wbase);
      cplx* x = p.w.x + (long long)s * P;
#pragma unroll
      for (int r = 0; r < P1; ++r) x[(long long)r * P2 + j] = v[r];
    }
  }
}

// Position (inside the P-point buffer) where the row DIT expects spectrum entry k: row k mod P1, digit-
// reversed place of k / P1 inside the row.
template <int LOG2P1> __device__ __forceinline__ long long long_pos(int k) {
  constexpr int P1 = 1 << LOG2P1;
  return (long long)(k & (P1 - 1)) * kLongP2 + digit_pos<kLongLog2P2>(k >> LOG2P1);
}

// ---- power-of-two long lines: write U_k where the inverse expects it --------------------------------
template <int LOG2P1>
__global__ void __launch_bounds__(256) em_long_scatter_kernel(const LongWave w) {
  constexpr long long P = (long long)1 << (LOG2P1 + kLongLog2P2);
  const int s = blockIdx.y;
  const SlotRec rec = w.recs[s];
  const int line = rec.line;
  const int N = rec.N;
  const double* W = rec.W;
  const double a = rec.a;
  const int pr = rec.pair;
  const double t1 = w.ts[2 * pr];
  const double t2 = (2 * pr + 1 < w.T) ? w.ts[2 * pr + 1] : t1;
  const double w1 = __dmul_rn(a, t1), w2 = __dmul_rn(a, t2);
  cplx* x = w.x + (long long)s * P;
  pdl_launch_dependents();
  pdl_wait();
  // one thread per mode pair (k, N-k): the four modal values serve both packed entries
  const double* __restrict__ lam = rec.lam;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; 2 * k <= N; k += gridDim.x * blockDim.x) {
    const double l = lam[k], wk = W[k];
    const double y1 = em_modal(l, wk, w1, t1), y2 = em_modal(l, wk, w2, t2);
    if (k == 0) { x[long_pos<LOG2P1>(0)] = cplx{y1, y2}; continue; }
    const int kk = N - k;
    if (kk == k) { x[long_pos<LOG2P1>(k)] = cmul(ldg_cplx(rec.pre + k), cplx{y1 + y2, y2 - y1}); continue; }
    const double lm = lam[kk], wm = W[kk];
    const double m1 = em_modal(lm, wm, w1, t1), m2 = em_modal(lm, wm, w2, t2);
    x[long_pos<LOG2P1>(k)] = cmul(ldg_cplx(rec.pre + k), cplx{y1 + m2, y2 - m1});
    x[long_pos<LOG2P1>(kk)] = cmul(ldg_cplx(rec.pre + kk), cplx{m1 + y2, m2 - y1});
  }
}

// ---- row pass ------------------------------------------------------------------------------------------
// mode 0: DIF, * H, DIT (Bluestein);  mode 1: DIF only, result written to `hhat_out` (table build);
// mode 2: DIT only (power-of-two lines).
struct RowParams {
  cplx* x;
  int p1;                  // rows per transform
  int mode;
  const SlotRec* recs;     // evaluation: the transform slots of the wave
  const NTab* tabs_build;
  const cplx* tw;          // per-stage twiddle tables of the 4096-point transform
};

// Outermost stage (stride P2/R0) of the row transform with one side in global memory: the DIF form reads its
// butterfly inputs straight from the row (lanes read consecutive columns, so every access is a coalesced 512-byte
// segment), the DIT form writes its outputs straight back.  Saves the staging pass through shared memory and one
// CTA barrier at each end of the kernel.
template <int NT>
__device__ __forceinline__ void row_dif_from_global(const cplx* __restrict__ row, cplx* __restrict__ data,
                                                    const cplx* __restrict__ twb, int tid) {
  constexpr int LOG2P = kLongLog2P2, R0 = plan_radix(LOG2P, 0), LOG2S = LOG2P - lg2(R0), S = 1 << LOG2S;
  for (int j = tid; j < S; j += NT) {
    cplx v[R0];
#pragma unroll
    for (int q = 0; q < R0; ++q) v[q] = row[j + q * S];
    Dft<R0, -1>::run(v);
    twiddle_jit<R0, -1>(v, twb[j]);
    cplx* d = data + pad_idx<LOG2P>(j);
#pragma unroll
    for (int q = 0; q < R0; ++q) d[pad_off<LOG2P, LOG2S>(q)] = v[q];
  }
}
template <int NT>
__device__ __forceinline__ void row_dit_to_global(cplx* __restrict__ row, const cplx* __restrict__ data,
                                                  const cplx* __restrict__ twb, int tid) {
  constexpr int LOG2P = kLongLog2P2, R0 = plan_radix(LOG2P, 0), LOG2S = LOG2P - lg2(R0), S = 1 << LOG2S;
  for (int j = tid; j < S; j += NT) {
    const cplx* d = data + pad_idx<LOG2P>(j);
    cplx v[R0];
#pragma unroll
    for (int q = 0; q < R0; ++q) v[q] = d[pad_off<LOG2P, LOG2S>(q)];
    twiddle_jit<R0, +1>(v, twb[j]);
    Dft<R0, +1>::run(v);
#pragma unroll
    for (int q = 0; q < R0; ++q) row[j + q * S] = v[q];
  }
}

__global__ void __launch_bounds__(kLongRowThreads, 2) em_long_row_kernel(const RowParams p) {
  constexpr int LOG2P = kLongLog2P2, P2 = kLongP2, NT = kLongRowThreads, NST = plan_n(LOG2P);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* twb = reinterpret_cast<cplx*>(smem_raw);
  cplx* data = twb + stage_tw_total(LOG2P);
  const int tid = threadIdx.x;
  const int r = blockIdx.x, s = blockIdx.y;
  pdl_launch_dependents();
  for (int i = tid; i < stage_tw_total(LOG2P); i += NT) twb[i] = p.tw[i];
  const long long P = (long long)P2 * p.p1;
  cplx* row = p.x + (long long)s * P + (long long)r * P2;
  const cplx* hh = (p.mode == 0) ? p.recs[s].hhat + (long long)r * P2 : nullptr;   // (not written by the predecessor)
  pdl_wait();                                         // the row comes from the preceding column / scatter kernel
  if (p.mode != 2) {
    __syncthreads();                                  // twiddles in place
    row_dif_from_global<NT>(row, data, twb, tid);
    __syncthreads();
    DifStages<LOG2P, 1, NST - 1, NT>::run(data, 1, twb, tid);
    if (p.mode == 1) {
      fft_stage<LOG2P, stage_radix<LOG2P, NST - 1>(), 0, true, -1, NT, 0>(data, 1, twb, tid);
      __syncthreads();
      cplx* out = const_cast<cplx*>(p.tabs_build[s].hhat) + (long long)r * P2;
      for (int i = tid; i < P2; i += NT) out[hh_pos_of(P2, stage_radix<LOG2P, NST - 1>(), i)] = data[pad_idx<LOG2P>(i)];
      return;
    }
    fft_middle<LOG2P, stage_radix<LOG2P, NST - 1>(), NT>(data, 1, hh, tid);
    stage_barrier<LOG2P, NST - 2>();
    DitStages<LOG2P, NST - 2, 1, NT>::run(data, 1, twb, tid);
  } else {
    // power-of-two lines: the innermost DIT stage (stride 1, no twiddles) takes its consecutive inputs straight
    // from the row
    constexpr int RM = stage_radix<LOG2P, NST - 1>();
    for (int b = tid; b < P2 / RM; b += NT) {
      cplx v[RM];
#pragma unroll
      for (int q = 0; q < RM; ++q) v[q] = row[b * RM + q];
      Dft<RM, +1>::run(v);
      cplx* d = data + pad_idx<LOG2P>(b * RM);
#pragma unroll
      for (int q = 0; q < RM; ++q) d[q] = v[q];
    }
    stage_barrier<LOG2P, NST - 2>();
    DitStages<LOG2P, NST - 2, 1, NT>::run(data, 1, twb, tid);
  }
  row_dit_to_global<NT>(row, data, twb, tid);
}

// ---- inverse column pass + epilogue -----------------------------------------------------------------------
template <int P1, bool BLUE>
__global__ void __launch_bounds__(kLongColsPerCta) em_long_colinv_kernel(const ColParams p) {
  constexpr int P2 = kLongP2;
  constexpr long long P = (long long)P1 * P2;
  __shared__ double red_v[2][kLongColsPerCta / 32];
  __shared__ int red_i[2][kLongColsPerCta / 32];
  const int j = blockIdx.x * kLongColsPerCta + threadIdx.x;
  pdl_launch_dependents();
  const cplx wbase = ldg_cplx(p.twl + P2 + j);            // w_P^{+j}
  const int s0 = blockIdx.y * p.tr_per_cta;
  const int s1 = min(p.w.n_tr, s0 + p.tr_per_cta);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  pdl_wait();                                         // x is the row kernel's output
  const double ninf = -__longlong_as_double(0x7ff0000000000000LL);
  for (int s = s0; s < s1; ++s) {
    const SlotRec rec = p.w.recs[s];
    const int pr = rec.pair;
    const int N = rec.N;
    const int halfN = (N + 1) >> 1;
    const cplx* x = p.w.x + (long long)s * P;
    // the post-chirp values of this thread's outputs: asked for first, needed last
    constexpr int QNE = BLUE ? ((P1 + 1) / 2 + 1 < P1 ? (P1 + 1) / 2 + 1 : P1) : 1;
    cplx postq[QNE];
    if (BLUE) {
#pragma unroll
      for (int q = 0; q < QNE; ++q) {
        postq[q] = cplx{1.0, 0.0};
        if (j + q * P2 < N) postq[q] = ldg_cplx(rec.post + j + q * P2);
      }
    }
    cplx v[P1];
#pragma unroll
    for (int r = 0; r < P1; ++r) v[r] = x[(long long)r * P2 + j];
    col_twiddle<P1>(v, wbase);
    DftAny<P1, +1>::run(v);
    const int T = p.w.T;
    const int tj1 = 2 * pr, tj2 = 2 * pr + 1;
    double* d1 = p.w.stress ? p.w.stress + rec.stress_off : nullptr;
    double* d2 = (p.w.stress && tj2 < T) ? d1 + N : nullptr;
    double b1 = ninf, b2 = ninf;
    int i1 = 0x7fffffff, i2 = 0x7fffffff;
#pragma unroll
    for (int q = 0; q < (BLUE ? (P1 + 1) / 2 + 1 : P1); ++q) {
      if (BLUE && q >= P1) break;
      const int m = j + q * P2;
      if (m < N) {
        const cplx u = BLUE ? cmul(v[q], postq[q < QNE ? q : 0]) : v[q];
        const int idx = (m < halfN) ? 2 * m : 2 * (N - 1 - m) + 1;
        if (d1) __stcs(d1 + idx, u.x);
        if (d2) __stcs(d2 + idx, u.y);
        if (u.x > b1 || (u.x == b1 && idx < i1)) { b1 = u.x; i1 = idx; }
        if (u.y > b2 || (u.y == b2 && idx < i2)) { b2 = u.y; i2 = idx; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ob1 = __shfl_xor_sync(0xffffffffu, b1, o); const int oi1 = __shfl_xor_sync(0xffffffffu, i1, o);
      if (ob1 > b1 || (ob1 == b1 && oi1 < i1)) { b1 = ob1; i1 = oi1; }
      const double ob2 = __shfl_xor_sync(0xffffffffu, b2, o); const int oi2 = __shfl_xor_sync(0xffffffffu, i2, o);
      if (ob2 > b2 || (ob2 == b2 && oi2 < i2)) { b2 = ob2; i2 = oi2; }
    }
    __syncthreads();
    if (lane == 0) { red_v[0][warp] = b1; red_i[0][warp] = i1; red_v[1][warp] = b2; red_i[1][warp] = i2; }
    __syncthreads();
    if (threadIdx.x < 2) {
      double b = red_v[threadIdx.x][0]; int bi = red_i[threadIdx.x][0];
      for (int q = 1; q < kLongColsPerCta / 32; ++q) {
        const double ob = red_v[threadIdx.x][q]; const int oi = red_i[threadIdx.x][q];
        if (ob > b || (ob == b && oi < bi)) { b = ob; bi = oi; }
      }
      const long long o = ((long long)(p.w.tr_base + s) * 2 + threadIdx.x) * gridDim.x + blockIdx.x;
      p.w.part_v[o] = b; p.w.part_i[o] = bi;
    }
  }
}

// ---- final reduction of the per-column-block partials -----------------------------------------------------
__global__ void __launch_bounds__(128) em_long_reduce_kernel(const LongWave w, int colblocks) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= 2 * w.n_tr) return;
  const int s = e >> 1, which = e & 1;
  const SlotRec rec = w.recs[s];
  const int line = rec.line;
  const int tj = 2 * rec.pair + which;
  if (tj >= w.T) return;
  double b = -__longlong_as_double(0x7ff0000000000000LL);
  int bi = 0x7fffffff;
  for (int c = 0; c < colblocks; ++c) {
    const double ob = w.part_v[(long long)e * colblocks + c];
    const int oi = w.part_i[(long long)e * colblocks + c];
    if (ob > b || (ob == b && oi < bi)) { b = ob; bi = oi; }
  }
  if (bi == 0x7fffffff) {
    bi = 0;
    b = w.stress ? w.stress[rec.stress_off + (long long)which * rec.N]
                 : __longlong_as_double(0x7ff8000000000000LL);
  }
  const long long o = (long long)line * w.T + tj;
  w.mx[o] = b; w.amx[o] = bi; w.flag[o] = (b > w.critical) ? 1 : 0;
}

}  // namespace em
