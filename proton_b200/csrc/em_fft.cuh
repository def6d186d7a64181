// em_fft.cuh — fused modal evaluation + inverse DCT (DCT-III) + stress write-back + per-line max.
//
// Replaces calc_analytical_solution + idct1d of the reference (bin/EMtool_analytical@0xc68c, @0xc2a0;
// SURVEY.md §8a rows 10-12) for every timestamp of a line in one kernel.
//
// Math.  For one line with N nodes the reference computes, per timestamp t,
//     sigma_n = sum_k Y_k cos(pi k (2n+1) / (2N)),   Y_k = D_k(t) * (forward-DCT weight of mode k).
// Two timestamps (t1, t2) are packed into ONE length-N complex inverse DFT (Makhoul's permutation):
//     U_0 = Y1_0 + i Y2_0,   U_k = e^{i pi k/2N} [ (Y1_k + Y2_{N-k}) + i (Y2_k - Y1_{N-k}) ]   (k >= 1,
//     the factor 1/2 of Makhoul's formula is folded into the stored weights)
//     u_m = sum_k U_k e^{+2 pi i m k / N}  ->  sigma1_{perm(m)} = Re u_m, sigma2_{perm(m)} = Im u_m,
//     perm(m) = 2m for m < ceil(N/2), else 2(N-1-m)+1.
// N a power of two: U is written in digit-reversed order and one in-place radix-8/4/2 DIT FFT of
// length P = N runs in shared memory.  Any other N: Bluestein's chirp-z with P = pow2 >= 2N-1:
//     a_k = U_k c_k,  c_j = e^{i pi j^2/N};  A = DIF-FFT_P(a);  A *= H (FFT of conj chirp, / P);
//     DIT-IFFT_P;  u_m = c_m * conv_m.   DIF leaves its output digit-reversed, which is exactly what
// the DIT consumes, so no reordering pass exists; the last DIF stage, the product with H and the
// first DIT stage happen in registers.
//
// One CTA owns one (line, range of timestamp pairs) item and keeps G transforms in flight in shared
// memory so that per-thread twiddles / modal constants are loaded once per G transforms.
#pragma once
#include "em_math.cuh"

namespace em {

// Per-distinct-N constant tables (built once per context, cached by N).
struct NTab {
  const double* lam;    // [N]   lambda_k = 2cos(k pi/N) - 2, computed by the HOST libm (see proton_em.cu)
  const cplx* pre;      // [N]   pow2: e^{i pi k/2N};  Bluestein: e^{i pi (k + 2k^2)/(2N)}
  const cplx* post;     // [N]   Bluestein only: c_m = e^{i pi m^2/N}
  const cplx* hhat;     // [P]   Bluestein only: DIF-FFT_P of conj chirp, digit-reversed order, scaled 1/P
  int n;
  int log2p;            // FFT length class (0 = not on the in-CTA FFT path)
  int bluestein;
  int p1;               // long classes: rows of 4096 points per transform (0 otherwise)
  int mix;              // mixed-radix in-CTA classes: mix_code(LR, ODD) (0 otherwise; em_mix.cuh)
};

struct EvalParams {
  const int* class_lines;      // line ids of this class
  int n_class_lines;
  int chunks;                  // items per line
  int pairs_per_chunk;
  int T;
  const int* num_nodes;
  const double* a_coeff;
  const long long* node_off;
  const int* line_tab;         // line -> index into tabs
  const NTab* tabs;
  const double* W;             // [sum N] folded modal weights (forward kernel)
  const double* ts;            // [T]
  const cplx* tw;              // base twiddles of this P, per stage (stage_tw_offset)
  const int* pos_tab;          // [P] padded shared-memory position of spectrum entry k for the DIT input (pow2 classes)
  double* stress;              // may be null (reduce-only)
  double* mx;                  // [L*T]
  int* amx;                    // [L*T]
  unsigned char* flag;         // [L*T]
  double critical;
};

// ---- small DFTs in registers ---------------------------------------------------------------------
template <int SIGN> __device__ __forceinline__ cplx mul_i(cplx v) {   // v * (SIGN * i)
  return SIGN > 0 ? cplx{-v.y, v.x} : cplx{v.y, -v.x};
}
template <int R, int SIGN> struct Dft;
template <int SIGN> struct Dft<2, SIGN> {
  static __device__ __forceinline__ void run(cplx (&v)[2]) {
    const cplx a = v[0];
    v[0] = cadd(a, v[1]); v[1] = csub(a, v[1]);
  }
};
template <int SIGN> struct Dft<4, SIGN> {
  static __device__ __forceinline__ void run(cplx (&v)[4]) {
    const cplx a0 = cadd(v[0], v[2]), a1 = csub(v[0], v[2]);
    const cplx b0 = cadd(v[1], v[3]), b1 = mul_i<SIGN>(csub(v[1], v[3]));
    v[0] = cadd(a0, b0); v[2] = csub(a0, b0);
    v[1] = cadd(a1, b1); v[3] = csub(a1, b1);
  }
};
template <int SIGN> struct Dft<8, SIGN> {
  static __device__ __forceinline__ void run(cplx (&v)[8]) {
    cplx e[4] = {v[0], v[2], v[4], v[6]};
    cplx o[4] = {v[1], v[3], v[5], v[7]};
    Dft<4, SIGN>::run(e);
    Dft<4, SIGN>::run(o);
    // o1 * e^{SIGN i pi/4} and o3 * e^{SIGN 3 i pi/4}: the factor 1/sqrt(2) rides on the final add as an FMA
    constexpr double h = 0.70710678118654752440;
    constexpr double s = (double)SIGN;
    const cplx t1 = cplx{o[1].x - s * o[1].y, s * o[1].x + o[1].y};
    const cplx o2 = mul_i<SIGN>(o[2]);
    const cplx t3 = cplx{-o[3].x - s * o[3].y, s * o[3].x - o[3].y};
    v[0] = cadd(e[0], o[0]); v[4] = csub(e[0], o[0]);
    v[1] = cplx{fma(h, t1.x, e[1].x), fma(h, t1.y, e[1].y)};
    v[5] = cplx{fma(-h, t1.x, e[1].x), fma(-h, t1.y, e[1].y)};
    v[2] = cadd(e[2], o2);   v[6] = csub(e[2], o2);
    v[3] = cplx{fma(h, t3.x, e[3].x), fma(h, t3.y, e[3].y)};
    v[7] = cplx{fma(-h, t3.x, e[3].x), fma(-h, t3.y, e[3].y)};
  }
};

// cos(2 pi m / 32), m in [0, 32)
__host__ __device__ constexpr double cos32(int m) {
  m &= 31;
  if (m > 16) m = 32 - m;            // cos is even around 16
  const bool neg = m > 8;
  if (neg) m = 16 - m;
  const double c = m == 0 ? 1.0 : m == 1 ? 0.98078528040323044913 : m == 2 ? 0.92387953251128675613
                   : m == 3 ? 0.83146961230254523708 : m == 4 ? 0.70710678118654752440
                   : m == 5 ? 0.55557023301960222474 : m == 6 ? 0.38268343236508977173
                   : m == 7 ? 0.19509032201612826785 : 0.0;
  return neg ? -c : c;
}
__host__ __device__ constexpr double sin32(int m) { return cos32(m - 8); }

// Radix-2 recursion on top of Dft<8>: natural order in and out.
template <int R, int SIGN> struct DftBig {
  static __device__ __forceinline__ void run(cplx (&v)[R]) {
    if constexpr (R <= 8) {
      Dft<R, SIGN>::run(v);
    } else {
      cplx e[R / 2], o[R / 2];
#pragma unroll
      for (int i = 0; i < R / 2; ++i) { e[i] = v[2 * i]; o[i] = v[2 * i + 1]; }
      DftBig<R / 2, SIGN>::run(e);
      DftBig<R / 2, SIGN>::run(o);
#pragma unroll
      for (int r = 0; r < R / 2; ++r) {
        cplx t;
        if (r == 0) t = o[r];
        else if (4 * r == R) t = mul_i<SIGN>(o[r]);
        else {
          // (explicit FMAs, like every product-sum of these kernels: the CPU twin of the tests repeats them bit for bit)
          const double c = cos32(r * (32 / R)), s = (double)SIGN * sin32(r * (32 / R));
          t = cmul(o[r], cplx{c, s});
        }
        v[r] = cadd(e[r], t);
        v[r + R / 2] = csub(e[r], t);
      }
    }
  }
};
template <int SIGN> struct Dft<16, SIGN> {
  static __device__ __forceinline__ void run(cplx (&v)[16]) { DftBig<16, SIGN>::run(v); }
};

// ---- stage plans: radices in DIF order (largest stride first); the first radix is always 8 --------
__host__ __device__ constexpr int lg2(int r) { return r == 2 ? 1 : r == 4 ? 2 : r == 8 ? 3 : 4; }
// LOG2P = 3*n8 + rem: n8 radix-8 stages followed by one radix-2 (rem 1) or radix-4 (rem 2) stage.  The 4096-point
// transform (in-CTA p = 12 classes and the rows of the long path) runs as 16 x 16 x 16 instead: 8 instead of 12 sweeps
// over shared memory, which is the busiest pipe of these kernels; a radix-16 butterfly keeps 64 data
// registers per thread, so those kernels run P/16 threads per CTA.
#ifndef EM_RADIX16
#define EM_RADIX16 1
#endif
// p = 8: 16 x 16 (the rows of the mixed-radix classes with up to 16 rows of 256 points, em_mix.cuh);  p = 12: 16 x 16 x 16;
// p = 13: 16 x 16 x 16 x 2.
#ifndef EM_PLAN16_8
#define EM_PLAN16_8 1
#endif
__host__ __device__ constexpr bool plan16(int log2p) {
  return EM_RADIX16 != 0 && ((EM_PLAN16_8 != 0 && log2p == 8) || log2p == 12 || log2p == 13);
}
__host__ __device__ constexpr int plan_n(int log2p) {
  return plan16(log2p) ? log2p / 4 + (log2p % 4 ? 1 : 0) : log2p / 3 + (log2p % 3 ? 1 : 0);
}
__host__ __device__ constexpr int plan_radix(int log2p, int i) {
  return plan16(log2p) ? (i < log2p / 4 ? 16 : (log2p % 4 == 1 ? 2 : log2p % 4 == 2 ? 4 : 8))
                       : (i < log2p / 3 ? 8 : (log2p % 3 == 1 ? 2 : 4));
}
template <int LOG2P, int I> __host__ __device__ constexpr int stage_radix() { return plan_radix(LOG2P, I); }
template <int LOG2P, int I> __host__ __device__ constexpr int stage_log2s() {
  int s = LOG2P;
  for (int j = 0; j <= I; ++j) s -= lg2(plan_radix(LOG2P, j));
  return s;
}

// Twiddles: ONE base value per (stage, j), w1 = e^{+2 pi i j / (R S)}, kept in shared memory; the powers
// w1^q (q < R) are formed in registers with R-2 complex products (the FP64 pipe has the slack, the L1/shared
// data pipe does not).  Stages with S == 1 need none.
__host__ __device__ constexpr int stage_tw_count(int log2p, int i) {
  int s = log2p;
  for (int j = 0; j <= i; ++j) s -= lg2(plan_radix(log2p, j));
  return s == 0 ? 0 : (1 << s);
}
__host__ __device__ constexpr int stage_tw_offset(int log2p, int i) {
  int off = 0;
  for (int j = 0; j < i; ++j) off += stage_tw_count(log2p, j);
  return off;
}
__host__ __device__ constexpr int stage_tw_total(int log2p) { return stage_tw_offset(log2p, plan_n(log2p)); }

// w[q] = w1^q for q = 1..R-1 (w[0] unused).  |w1| = 1, so w^{q+1} = 2 Re(w1) w^q - w^{q-1}: two FMAs per power
// (error growth is bounded by the Chebyshev polynomials U_q <= q + 1, i.e. a few ulp for q <= 7).
template <int RADIX> __device__ __forceinline__ void tw_powers(cplx w1, cplx (&w)[RADIX]) {
  w[1] = w1;
  if constexpr (RADIX >= 4) {
    const double tx = w1.x + w1.x;
    w[2] = cplx{fma(tx, w1.x, -1.0), tx * w1.y};
#pragma unroll
    for (int q = 3; q < RADIX; ++q) w[q] = cplx{fma(tx, w[q - 1].x, -w[q - 2].x), fma(tx, w[q - 1].y, -w[q - 2].y)};
  }
}

// v[q] *= w^q (q = 1..R-1), w = (tv.x, SIGN tv.y): the same recurrence, nothing kept (radix 16 has no registers to spare).
template <int RADIX, int SIGN> __device__ __forceinline__ void twiddle_jit(cplx (&v)[RADIX], cplx tv) {
  const double tx = tv.x + tv.x;
  cplx wp = cplx{1.0, 0.0}, wc = cplx{tv.x, SIGN > 0 ? tv.y : -tv.y};
  v[1] = cmul(v[1], wc);
#pragma unroll
  for (int q = 2; q < RADIX; ++q) {
    const cplx wn = cplx{fma(tx, wc.x, -wp.x), fma(tx, wc.y, -wp.y)};
    v[q] = cmul(v[q], wn);
    wp = wc; wc = wn;
  }
}

// Shared-memory index padding (units of 16-byte complex): one pad slot per 8 (16 when the innermost radix is 16)
// elements keeps the stride-R accesses of the s = 1 stages conflict-free, one more per P/R0 elements does the same
// for the digit-reversed scatter of the power-of-two path (checked by brute force for every plan).
// Shifts found by exhaustive search per plan (tests/test_fft_plans.py repeats the check): every stage, the natural-order
// accesses and the digit-reversed scatter are free of bank conflicts for each of them.
__host__ __device__ constexpr int pad_sh1(int log2p) { return plan_radix(log2p, plan_n(log2p) - 1) == 16 ? 4 : 3; }
__host__ __device__ constexpr int pad_sh2(int log2p) {
  return log2p == 6 ? 0 : log2p == 7 ? 6 : log2p == 8 ? (plan16(8) ? 0 : 6) : log2p - lg2(plan_radix(log2p, 0));
}
__host__ __device__ constexpr int pad_sh3(int log2p) { return (log2p == 13 && plan16(13)) ? 4 : 0; }
__host__ __device__ constexpr int pad_idx_rt(int log2p, int p) {
  return p + (p >> pad_sh1(log2p)) + (pad_sh2(log2p) ? (p >> pad_sh2(log2p)) : 0) + (pad_sh3(log2p) ? (p >> pad_sh3(log2p)) : 0);
}
template <int LOG2P> __host__ __device__ constexpr int pad_idx(int p) { return pad_idx_rt(LOG2P, p); }
template <int LOG2P> __host__ __device__ constexpr int padded_len() { return (pad_idx_rt(LOG2P, (1 << LOG2P) - 1) + 8) & ~7; }

// Padded offset of element q of a butterfly relative to its first element: for every stage of the plans above it
// does not depend on the butterfly, so shared-memory accesses use immediate offsets.
template <int LOG2P, int LOG2S> __host__ __device__ constexpr int pad_off(int q) {
  return pad_idx<LOG2P>(q << LOG2S);
}

// Position where DIF leaves output k (and where DIT expects input k).
template <int LOG2P> __device__ __forceinline__ int digit_pos(int k) {
  int p = 0;
  int sh = LOG2P;
#pragma unroll
  for (int i = 0; i < plan_n(LOG2P); ++i) {
    const int lr = lg2(plan_radix(LOG2P, i));
    sh -= lr;
    p |= (k & ((1 << lr) - 1)) << sh;
    k >>= lr;
  }
  return p;
}

// Class configuration: threads per CTA and transforms in flight.
template <int LOG2P, bool BLUE> struct Cfg {
  static constexpr int P = 1 << LOG2P;
#ifndef EM_NT_MID
#define EM_NT_MID 256
#endif
#ifndef EM_G10
#define EM_G10 2
#endif
  static constexpr int NT = LOG2P >= 13 ? (plan16(13) ? 512 : 1024) : LOG2P == 12 ? (plan16(12) ? 256 : 512)
                            : (LOG2P == 8 && plan16(8)) ? 128 : EM_NT_MID;      // (p = 8: 16 butterflies x 8 transforms in flight)
  // rows of the staging buffer: N <= P (pow2) or N <= P/2 + 1 (Bluestein)
  static constexpr int NSTAGE = BLUE ? (P / 2 + 8) : P;
  static constexpr int G = LOG2P <= 6 ? 32 : LOG2P == 7 ? 16 : LOG2P == 8 ? 8 : LOG2P == 9 ? 4 : LOG2P == 10 ? EM_G10 : 1;
  static constexpr bool ALIAS_STAGE = LOG2P >= 12;       // staging rows reuse the data buffer
  static constexpr size_t data_bytes = (size_t)G * padded_len<LOG2P>() * sizeof(cplx);
  static constexpr size_t stage_bytes = ALIAS_STAGE ? 0 : (size_t)G * 2 * NSTAGE * sizeof(double);
  static constexpr size_t tw_bytes = (size_t)stage_tw_total(LOG2P) * sizeof(cplx);
  static constexpr size_t head_bytes = tw_bytes;
  static constexpr size_t smem_bytes = head_bytes + data_bytes + stage_bytes;
};

// One in-place stage over the G transforms in flight.
//   DIF: y = DFT_R(x) then y_r *= w^(j r);   DIT: x_q *= w^(j q) then DFT_R.   w = e^{SIGN 2 pi i / (R S)}.
template <int LOG2P, int RADIX, int LOG2S, bool DIF, int SIGN, int NT, int TWOFF>
__device__ __forceinline__ void fft_stage(cplx* __restrict__ data, int gcount, const cplx* __restrict__ tw, int tid) {
  constexpr int P = 1 << LOG2P, S = 1 << LOG2S, LOG2R = lg2(RADIX), NB = P / RADIX;
  constexpr int BSTEP = NB < NT ? NB : NT;
  constexpr int GSTEP = NB < NT ? NT / NB : 1;
  constexpr int TSTRIDE = padded_len<LOG2P>();
  const int g0 = NB < NT ? tid / NB : 0;
  for (int b = (NB < NT ? tid % NB : tid); b < NB; b += BSTEP) {
    const int j = b & (S - 1);
    const int base = ((b >> LOG2S) << (LOG2S + LOG2R)) | j;
    const int pbase = pad_idx<LOG2P>(base);
    if constexpr (RADIX == 16) {
      // no room for a table of powers next to 16 complex values: twiddles are formed as they are applied
      cplx tv = cplx{1.0, 0.0};
      if (S > 1) tv = tw[TWOFF + j];
      for (int g = g0; g < gcount; g += GSTEP) {
        cplx* d = data + g * TSTRIDE + pbase;
        cplx v[RADIX];
#pragma unroll
        for (int q = 0; q < RADIX; ++q) v[q] = d[pad_off<LOG2P, LOG2S>(q)];
        if (DIF) {
          Dft<RADIX, SIGN>::run(v);
          if (S > 1) twiddle_jit<RADIX, SIGN>(v, tv);
        } else {
          if (S > 1) twiddle_jit<RADIX, SIGN>(v, tv);
          Dft<RADIX, SIGN>::run(v);
        }
#pragma unroll
        for (int q = 0; q < RADIX; ++q) d[pad_off<LOG2P, LOG2S>(q)] = v[q];
      }
      continue;
    }
    cplx w[RADIX];
    if (S > 1) {
      const cplx tv = tw[TWOFF + j];
      tw_powers<RADIX>(cplx{tv.x, SIGN > 0 ? tv.y : -tv.y}, w);
    }
    for (int g = g0; g < gcount; g += GSTEP) {
      cplx* d = data + g * TSTRIDE + pbase;
      cplx v[RADIX];
#pragma unroll
      for (int q = 0; q < RADIX; ++q) v[q] = d[pad_off<LOG2P, LOG2S>(q)];
      if (DIF) {
        Dft<RADIX, SIGN>::run(v);
        if (S > 1) {
#pragma unroll
          for (int q = 1; q < RADIX; ++q) v[q] = cmul(v[q], w[q]);
        }
      } else {
        if (S > 1) {
#pragma unroll
          for (int q = 1; q < RADIX; ++q) v[q] = cmul(v[q], w[q]);
        }
        Dft<RADIX, SIGN>::run(v);
      }
#pragma unroll
      for (int q = 0; q < RADIX; ++q) d[pad_off<LOG2P, LOG2S>(q)] = v[q];
    }
  }
}

// First DIF stage of Bluestein (radix R0, stride P/R0) with the zero padding never materialised: the chirped
// spectrum occupies positions < N <= P/2, so the upper half of the butterfly inputs is zero and the lower half is read
// only where it falls below N.  Saves the zero-fill pass and half of this stage's shared-memory reads.
template <int LOG2P, int NT>
__device__ __forceinline__ void fft_stage0_pruned(cplx* __restrict__ data, int gcount, const cplx* __restrict__ tw,
                                                  int tid, int n_valid) {
  constexpr int R0 = plan_radix(LOG2P, 0);
  constexpr int LOG2S = LOG2P - lg2(R0), S = 1 << LOG2S, NB = S;
  constexpr int BSTEP = NB < NT ? NB : NT;
  constexpr int GSTEP = NB < NT ? NT / NB : 1;
  constexpr int TSTRIDE = padded_len<LOG2P>();
  const int g0 = NB < NT ? tid / NB : 0;
  if constexpr (R0 == 16) {
    // inputs 8..15 are zero: X[2r] = DFT_8(x)[r], X[2r+1] = DFT_8(x_q w16^q)[r]
    for (int j = (NB < NT ? tid % NB : tid); j < NB; j += BSTEP) {
      const cplx tv = tw[j];
      const int pbase = pad_idx<LOG2P>(j);
      for (int g = g0; g < gcount; g += GSTEP) {
        cplx* d = data + g * TSTRIDE + pbase;
        cplx a[8], b[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          a[q] = (j + q * S < n_valid) ? d[pad_off<LOG2P, LOG2S>(q)] : cplx{0.0, 0.0};
          const double c = cos32(2 * q), sn = -sin32(2 * q);
          b[q] = q == 0 ? a[q] : cmul(a[q], cplx{c, sn});
        }
        Dft<8, -1>::run(a);
        Dft<8, -1>::run(b);
        cplx v[16];
#pragma unroll
        for (int r = 0; r < 8; ++r) { v[2 * r] = a[r]; v[2 * r + 1] = b[r]; }
        twiddle_jit<16, -1>(v, tv);
#pragma unroll
        for (int q = 0; q < 16; ++q) d[pad_off<LOG2P, LOG2S>(q)] = v[q];
      }
    }
    return;
  }
  for (int j = (NB < NT ? tid % NB : tid); j < NB; j += BSTEP) {
    cplx w[8];
    const cplx tv = tw[j];
    tw_powers<8>(cplx{tv.x, -tv.y}, w);
    const int pbase = pad_idx<LOG2P>(j);
    const bool in1 = j + S < n_valid, in2 = j + 2 * S < n_valid, in3 = j + 3 * S < n_valid;
    for (int g = g0; g < gcount; g += GSTEP) {
      cplx* d = data + g * TSTRIDE + pbase;
      const cplx z{0.0, 0.0};
      const cplx v0 = (j < n_valid) ? d[0] : z;
      const cplx v1 = in1 ? d[pad_off<LOG2P, LOG2S>(1)] : z;
      const cplx v2 = in2 ? d[pad_off<LOG2P, LOG2S>(2)] : z;
      const cplx v3 = in3 ? d[pad_off<LOG2P, LOG2S>(3)] : z;
      // DFT_8 (sign -1) of (v0, v1, v2, v3, 0, 0, 0, 0)
      const cplx i2 = mul_i<-1>(v2), i3 = mul_i<-1>(v3);
      const cplx e0 = cadd(v0, v2), e2 = csub(v0, v2), e1 = cadd(v0, i2), e3 = csub(v0, i2);
      const cplx o0 = cadd(v1, v3), o2 = csub(v1, v3), o1 = cadd(v1, i3), o3 = csub(v1, i3);
      constexpr double h = 0.70710678118654752440;
      const cplx r1 = cplx{o1.x + o1.y, o1.y - o1.x};                  // o1 * e^{-i pi/4} * sqrt(2)
      const cplx r2 = mul_i<-1>(o2);
      const cplx r3 = cplx{o3.y - o3.x, -o3.x - o3.y};                 // o3 * e^{-3 i pi/4} * sqrt(2)
      cplx v[8];
      v[0] = cadd(e0, o0); v[4] = csub(e0, o0);
      v[1] = cplx{fma(h, r1.x, e1.x), fma(h, r1.y, e1.y)}; v[5] = cplx{fma(-h, r1.x, e1.x), fma(-h, r1.y, e1.y)};
      v[2] = cadd(e2, r2); v[6] = csub(e2, r2);
      v[3] = cplx{fma(h, r3.x, e3.x), fma(h, r3.y, e3.y)}; v[7] = cplx{fma(-h, r3.x, e3.x), fma(-h, r3.y, e3.y)};
#pragma unroll
      for (int q = 1; q < 8; ++q) v[q] = cmul(v[q], w[q]);
#pragma unroll
      for (int q = 0; q < 8; ++q) d[pad_off<LOG2P, LOG2S>(q)] = v[q];
    }
  }
}

// Where the transformed chirp H of a row of NB * RADIX points keeps element q of the innermost butterfly b (logical,
// digit-reversed index b * RADIX + q): at q * NB + b, so that the 16-byte loads of a warp — consecutive b, one q — are
// contiguous (four 128-byte lines per request).  In butterfly order every lane sat in a line of its own: 32 lines per
// request, and the H loads alone were 43 % of the L1 data-pipe wavefronts of the long row kernel
// (profiles/r02_ncu_long_full.txt: 2.09 M of 4.9 M).
#ifndef EM_HH_LANE_MAJOR
#define EM_HH_LANE_MAJOR 1
#endif
template <int RADIX, int NB> __host__ __device__ constexpr int hh_pos(int b, int q) {
  return EM_HH_LANE_MAJOR ? q * NB + b : b * RADIX + q;
}
// the same from the logical index i = b * RADIX + q of a row of `len` points
__host__ __device__ constexpr int hh_pos_of(int len, int radix, int i) {
  return EM_HH_LANE_MAJOR ? (i % radix) * (len / radix) + i / radix : i;
}

// Middle of Bluestein: last DIF stage (s = 1), product with H, first DIT stage (s = 1) in registers.
template <int LOG2P, int RADIX, int NT>
__device__ __forceinline__ void fft_middle(cplx* __restrict__ data, int gcount, const cplx* __restrict__ hhat, int tid) {
  constexpr int P = 1 << LOG2P, NB = P / RADIX;
  constexpr int BSTEP = NB < NT ? NB : NT;
  constexpr int GSTEP = NB < NT ? NT / NB : 1;
  constexpr int TSTRIDE = padded_len<LOG2P>();
  const int g0 = NB < NT ? tid / NB : 0;
  for (int b = (NB < NT ? tid % NB : tid); b < NB; b += BSTEP) {
    const int base = b * RADIX;
    const int pbase = pad_idx<LOG2P>(base);
    if constexpr (RADIX == 16) {
      for (int g = g0; g < gcount; g += GSTEP) {
        cplx* d = data + g * TSTRIDE + pbase;
        cplx v[RADIX];
#pragma unroll
        for (int q = 0; q < RADIX; ++q) v[q] = d[q];
        Dft<RADIX, -1>::run(v);
#pragma unroll
        for (int q = 0; q < RADIX; ++q) v[q] = cmul(v[q], ldg_cplx(hhat + hh_pos<RADIX, NB>(b, q)));
        Dft<RADIX, +1>::run(v);
#pragma unroll
        for (int q = 0; q < RADIX; ++q) d[q] = v[q];
      }
      continue;
    }
    cplx h[RADIX];
#pragma unroll
    for (int q = 0; q < RADIX; ++q) h[q] = ldg_cplx(hhat + hh_pos<RADIX, NB>(b, q));
    for (int g = g0; g < gcount; g += GSTEP) {
      cplx* d = data + g * TSTRIDE + pbase;
      cplx v[RADIX];
#pragma unroll
      for (int q = 0; q < RADIX; ++q) v[q] = d[q];
      Dft<RADIX, -1>::run(v);
#pragma unroll
      for (int q = 0; q < RADIX; ++q) v[q] = cmul(v[q], h[q]);
      Dft<RADIX, +1>::run(v);
#pragma unroll
      for (int q = 0; q < RADIX; ++q) d[q] = v[q];
    }
  }
}

// The two innermost stages (strides R and 1) of a plan that ends in two equal radices touch, per thread, the same
// blocks of R*R consecutive points, and both give butterfly b of transform g to the same thread: a warp's outputs of
// one are exactly its inputs of the other, so a warp barrier replaces the CTA barrier between them.
template <int LOG2P, int I> __host__ __device__ constexpr bool warp_local_after() {   // between stage I and stage I+1
  return I + 1 == plan_n(LOG2P) - 1 && plan_radix(LOG2P, I) == plan_radix(LOG2P, I + 1);
}
template <int LOG2P, int I> __device__ __forceinline__ void stage_barrier() {
  if constexpr (warp_local_after<LOG2P, I>()) __syncwarp(); else __syncthreads();
}

// Stage recursion helpers (compile-time unrolled over the plan).
template <int LOG2P, int I, int IEND, int NT> struct DifStages {   // stages I .. IEND-1, SIGN = -1
  static __device__ __forceinline__ void run(cplx* data, int gcount, const cplx* tw, int tid) {
    if constexpr (I < IEND) {
      fft_stage<LOG2P, stage_radix<LOG2P, I>(), stage_log2s<LOG2P, I>(), true, -1, NT, stage_tw_offset(LOG2P, I)>(data, gcount, tw, tid);
      stage_barrier<LOG2P, I>();
      DifStages<LOG2P, I + 1, IEND, NT>::run(data, gcount, tw, tid);
    }
  }
};
template <int LOG2P, int I, int ISTOP, int NT> struct DitStages {   // stages I down to ISTOP, SIGN = +1
  static __device__ __forceinline__ void run(cplx* data, int gcount, const cplx* tw, int tid) {
    if constexpr (I >= ISTOP) {
      fft_stage<LOG2P, stage_radix<LOG2P, I>(), stage_log2s<LOG2P, I>(), false, +1, NT, stage_tw_offset(LOG2P, I)>(data, gcount, tw, tid);
      if constexpr (I >= 1) stage_barrier<LOG2P, I - 1>(); else __syncthreads();
      DitStages<LOG2P, I - 1, ISTOP, NT>::run(data, gcount, tw, tid);
    }
  }
};

template <int LOG2P, bool BLUE>
__global__ void __launch_bounds__(Cfg<LOG2P, BLUE>::NT, (LOG2P <= 11 ? 3 : LOG2P == 12 ? 2 : 1))
em_eval_fft_kernel(const EvalParams prm) {
  using C = Cfg<LOG2P, BLUE>;
  constexpr int P = C::P, NT = C::NT, G = C::G, NST = plan_n(LOG2P);
  constexpr int TSTRIDE = padded_len<LOG2P>();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* twb = reinterpret_cast<cplx*>(smem_raw);
  cplx* data = reinterpret_cast<cplx*>(smem_raw + C::head_bytes);
  double* stage = C::ALIAS_STAGE ? reinterpret_cast<double*>(data)
                                 : reinterpret_cast<double*>(smem_raw + C::head_bytes + C::data_bytes);
  const int tid = threadIdx.x;
  for (int i = tid; i < stage_tw_total(LOG2P); i += NT) twb[i] = prm.tw[i];

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
  __syncthreads();

  for (int pb = pair_begin; pb < pair_end; pb += G) {
    const int gcount = min(G, pair_end - pb);
    // ---- phase 1: modal values -> packed complex spectrum U in shared memory -------------------
    {
      // mode pairs (k, N-k), k = 1 .. KP; the two unpaired modes (k = 0, and k = N/2 for even N) come after
      const int KP = (N - 1) >> 1;
      int kp0, kstep, g0, gstep;
      if (KP >= NT) { kp0 = tid + 1; kstep = NT; g0 = 0; gstep = 1; }
      else if (KP > 0) {
        const int groups = NT / KP;
        kp0 = tid % KP + 1; kstep = KP; g0 = tid / KP; gstep = groups;
        if (g0 >= groups) kp0 = KP + 1;
      } else { kp0 = 1; kstep = 1; g0 = 0; gstep = 1; }
      for (int kp = kp0; kp <= KP; kp += kstep) {
        const int k = kp, kk = N - kp;
        const double lamk = lam[k], wk = W[k], lamkk = lam[kk], wkk = W[kk];
        const cplx prek = ldg_cplx(tab.pre + k), prekk = ldg_cplx(tab.pre + kk);
        const int posk = BLUE ? pad_idx<LOG2P>(k) : __ldg(prm.pos_tab + k);
        const int poskk = BLUE ? pad_idx<LOG2P>(kk) : __ldg(prm.pos_tab + kk);
        for (int g = g0; g < gcount; g += gstep) {
          const int j1 = 2 * (pb + g);
          const double t1 = prm.ts[j1];
          const double t2 = (j1 + 1 < T) ? prm.ts[j1 + 1] : t1;
          const double w1 = __dmul_rn(a, t1), w2 = __dmul_rn(a, t2);
          cplx* d = data + g * TSTRIDE;
          const double y1k = em_modal(lamk, wk, w1, t1);
          const double y2k = em_modal(lamk, wk, w2, t2);
          const double y1kk = em_modal(lamkk, wkk, w1, t1);
          const double y2kk = em_modal(lamkk, wkk, w2, t2);
          d[posk] = cmul(prek, cplx{y1k + y2kk, y2k - y1kk});
          d[poskk] = cmul(prekk, cplx{y1kk + y2k, y2kk - y1k});
        }
      }
      if (tid < 2 * gcount) {
        const int g = tid >> 1, which = tid & 1;
        const int k = which ? (N >> 1) : 0;
        if (which == 0 || (((N & 1) == 0) && N > 1)) {
          const int j1 = 2 * (pb + g);
          const double t1 = prm.ts[j1];
          const double t2 = (j1 + 1 < T) ? prm.ts[j1 + 1] : t1;
          const double w1 = __dmul_rn(a, t1), w2 = __dmul_rn(a, t2);
          const double y1 = em_modal(lam[k], W[k], w1, t1);
          const double y2 = em_modal(lam[k], W[k], w2, t2);
          const int posk = pad_idx<LOG2P>(BLUE ? k : digit_pos<LOG2P>(k));
          data[g * TSTRIDE + posk] = which ? cmul(ldg_cplx(tab.pre + k), cplx{y1 + y2, y2 - y1}) : cplx{y1, y2};
        }
      }
    }
    __syncthreads();
    // ---- phase 2: transforms ------------------------------------------------------------------
    if (BLUE) {
      fft_stage0_pruned<LOG2P, NT>(data, gcount, twb, tid, N);   // N <= P/2: the zero padding is implicit
      __syncthreads();
      DifStages<LOG2P, 1, NST - 1, NT>::run(data, gcount, twb, tid);
      fft_middle<LOG2P, stage_radix<LOG2P, NST - 1>(), NT>(data, gcount, tab.hhat, tid);
      stage_barrier<LOG2P, NST - 2>();
      DitStages<LOG2P, NST - 2, 1, NT>::run(data, gcount, twb, tid);
    } else {
      DitStages<LOG2P, NST - 1, 1, NT>::run(data, gcount, twb, tid);
    }
    // ---- last DIT stage (radix R0, stride P/R0) fused with chirp post-multiply + de-interleave ---
    {
      constexpr int R0 = plan_radix(LOG2P, 0), LOG2S0 = LOG2P - lg2(R0);
      constexpr int S = P / R0, NB = P / R0;
      constexpr int QN = BLUE ? R0 / 2 : R0;    // Bluestein: only outputs m < N <= P/2 are wanted
      if constexpr (!C::ALIAS_STAGE) {
        constexpr int BSTEP = NB < NT ? NB : NT;
        constexpr int GSTEP = NB < NT ? NT / NB : 1;
        const int g0 = NB < NT ? tid / NB : 0;
        for (int b = (NB < NT ? tid % NB : tid); b < NB; b += BSTEP) {
          const int j = b;
          cplx w[8];
          if constexpr (R0 == 8) tw_powers<8>(twb[j], w);
          cplx post[QN];
          if (BLUE) {
#pragma unroll
            for (int q = 0; q < QN; ++q) {
              const int m = j + q * S;
              post[q] = cplx{1.0, 0.0};
              if (m < N) post[q] = ldg_cplx(tab.post + m);
            }
          }
          for (int g = g0; g < gcount; g += GSTEP) {
            const cplx* d = data + g * TSTRIDE + pad_idx<LOG2P>(j);
            cplx v[R0];
#pragma unroll
            for (int q = 0; q < R0; ++q) v[q] = d[pad_off<LOG2P, LOG2S0>(q)];
            if constexpr (R0 == 8) {
#pragma unroll
              for (int q = 1; q < 8; ++q) v[q] = cmul(v[q], w[q]);
            } else {
              twiddle_jit<R0, +1>(v, twb[j]);
            }
            Dft<R0, +1>::run(v);
            double* s1 = stage + (size_t)g * 2 * C::NSTAGE;
            double* s2 = s1 + C::NSTAGE;
#pragma unroll
            for (int q = 0; q < QN; ++q) {
              const int m = j + q * S;
              if (m < N) {
                const cplx u = BLUE ? cmul(v[q], post[q]) : v[q];
                const int idx = (m < halfN) ? 2 * m : 2 * (N - 1 - m) + 1;
                s1[idx] = u.x; s2[idx] = u.y;
              }
            }
          }
        }
      } else {
        // G == 1 and the staging rows live inside the data buffer: keep results in registers across a
        // barrier, then overwrite the buffer.
        static_assert(!C::ALIAS_STAGE || (G == 1 && NB % NT == 0), "alias staging needs G == 1");
        constexpr int ITERS = NB / NT;
        cplx vv[ITERS][QN];
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
          const int j = tid + it * NT;
          cplx v[R0];
#pragma unroll
          for (int q = 0; q < R0; ++q) v[q] = data[pad_idx<LOG2P>(j) + pad_off<LOG2P, LOG2S0>(q)];
          twiddle_jit<R0, +1>(v, twb[j]);
          Dft<R0, +1>::run(v);
#pragma unroll
          for (int q = 0; q < QN; ++q) {
            const int m = j + q * S;
            vv[it][q] = v[q];
            if (BLUE && m < N) vv[it][q] = cmul(v[q], ldg_cplx(tab.post + m));
          }
        }
        __syncthreads();
        double* s1 = stage;
        double* s2 = s1 + C::NSTAGE;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
#pragma unroll
          for (int q = 0; q < QN; ++q) {
            const int m = tid + it * NT + q * S;
            if (m < N) {
              const int idx = (m < halfN) ? 2 * m : 2 * (N - 1 - m) + 1;
              s1[idx] = vv[it][q].x; s2[idx] = vv[it][q].y;
            }
          }
        }
      }
    }
    __syncthreads();
    // ---- phase 3: coalesced store + max / argmax; every row is split over WPR warps so that all warps of the
    // CTA take part (with one transform in flight there are only two rows) -------------------------------------
    {
      constexpr int NWARPS = NT >> 5;
      constexpr int WPR = (NWARPS / (2 * G)) > 1 ? (NWARPS / (2 * G)) : 1;     // warps per row
      constexpr int RSTEP = NWARPS / WPR;                                       // rows handled per sweep
      __shared__ double red_v[NWARPS];
      __shared__ int red_i[NWARPS];
      const int warp = tid >> 5, lane = tid & 31;
      const int seg = warp % WPR;
      const int seglen = (((N + WPR - 1) / WPR) + 31) & ~31;
      const int n_begin = seg * seglen, n_end = min(N, n_begin + seglen);
      const double ninf = -__longlong_as_double(0x7ff0000000000000LL);
      for (int rr = warp / WPR; rr < 2 * gcount; rr += RSTEP) {
        const int g = rr >> 1, which = rr & 1;
        const int tj = 2 * (pb + g) + which;
        const double* src = stage + (size_t)g * 2 * C::NSTAGE + (size_t)which * C::NSTAGE;
        double* dst = (prm.stress && tj < T) ? prm.stress + noff * (long long)T + (long long)tj * N : nullptr;
        double best = ninf;
        int besti = 0x7fffffff;
        if (tj < T) {
          for (int n = n_begin + lane; n < n_end; n += 32) {
            const double v = src[n];
            if (dst) __stcs(dst + n, v);
            if (v > best) { best = v; besti = n; }
          }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const double ob = __shfl_xor_sync(0xffffffffu, best, o);
          const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
          if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
        }
        if (WPR == 1) {
          if (lane == 0 && tj < T) {
            if (besti == 0x7fffffff) { besti = 0; best = src[0]; }
            const long long o = (long long)line * T + tj;
            prm.mx[o] = best; prm.amx[o] = besti;
            prm.flag[o] = (best > prm.critical) ? 1 : 0;
          }
        } else if (lane == 0) {
          red_v[warp] = best; red_i[warp] = besti;
        }
      }
      if (WPR > 1) {
        // WPR > 1 means RSTEP >= 2*G: every warp had at most one row; combine its segments through smem
        __syncthreads();
        const int rr = warp / WPR;
        const int g = rr >> 1, which = rr & 1;
        const int tj = 2 * (pb + g) + which;
        if (seg == 0 && lane == 0 && rr < 2 * gcount && tj < T) {
          double best = red_v[warp];
          int besti = red_i[warp];
          for (int q = 1; q < WPR; ++q) {
            const double ob = red_v[warp + q]; const int oi = red_i[warp + q];
            if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
          }
          if (besti == 0x7fffffff) { besti = 0; best = stage[(size_t)g * 2 * C::NSTAGE + (size_t)which * C::NSTAGE]; }
          const long long o = (long long)line * T + tj;
          prm.mx[o] = best; prm.amx[o] = besti;
          prm.flag[o] = (best > prm.critical) ? 1 : 0;
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace em
