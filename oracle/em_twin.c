/* oracle/em_twin.c — TEST INFRASTRUCTURE ONLY: a deterministic CPU twin of the GPU ARITHMETIC.
 *
 * em_oracle.c restates the reference's algorithm (FFTW-style DCTs); two different DCT algorithms cannot agree in
 * the last bit, so the GPU path is compared with it norm-wise (1e-9).  This file restates instead what the CUDA
 * kernels themselves do — the same butterflies, in the same order, with the same fused multiply-adds — so that the
 * stress field, and with it max / argmax / flag, can be required to be BIT-IDENTICAL (SURVEY.md §7 hard part 2).
 * Twins (tests/test_twin.py):
 *   em_twin_dense   em_eval_dense_kernel         proton_b200/csrc/em_kernels.cuh   (N <= 88: DMMA m8n8k4 chains)
 *   em_twin_pow2    em_eval_fft_kernel<p,false>  proton_b200/csrc/em_fft.cuh       (N = 2^p, 64..8192: in-CTA DIT FFT)
 *   em_twin_blue    em_eval_fft_kernel<p,true>   proton_b200/csrc/em_fft.cuh       (other N <= 4096: Bluestein, P = 2^p)
 *   em_twin_rows    em_eval_mix_kernel<LR,ROWS>  proton_b200/csrc/em_mix.cuh       (Bluestein, P = 5 or 6 rows of 2^LR)
 *                   em_long_colfwd_paired / em_long_row / em_long_colinv            (4096 < N <= 65536: P = P1 x 4096)
 *   em_twin_long_pow2  em_long_scatter / em_long_row (inverse only) / em_long_colinv (N = 2^p, 16384..131072)
 * Inputs that come from the DEVICE's sincospi (pre-twiddles, cosine table, modal weights W) are read back through
 * proton_em_debug_line_table and handed in (also the transformed chirp H, which the device builds with the same
 * kernels); lambda is the host-libm table both sides share; the stage twiddles and the small-DFT twiddles are
 * recomputed here exactly as proton_em_create does (long double cosl / sinl).
 * Build: -O2 -ffp-contract=off (every FMA below is explicit; the device code is built with -fmad=false).
 * Only tests/ loads this; nothing in the product links it.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { double x, y; } cplx;

static inline cplx cmul(cplx a, cplx b) { cplx r = {fma(a.x, b.x, -(a.y * b.y)), fma(a.x, b.y, a.y * b.x)}; return r; }
static inline cplx cadd(cplx a, cplx b) { cplx r = {a.x + b.x, a.y + b.y}; return r; }
static inline cplx csub(cplx a, cplx b) { cplx r = {a.x - b.x, a.y - b.y}; return r; }
static inline cplx mul_i(cplx v, int sign) { cplx r; if (sign > 0) { r.x = -v.y; r.y = v.x; } else { r.x = v.y; r.y = -v.x; } return r; }

/* ---- modal factor: em_math.cuh em_expm1_ref / em_modal --------------------------------------------------------- */
static const double kEmC[8] = {0x1.71547652b82fep0 * 128.0, 0x1.8p52, -0x1.62e42fefa0000p-8, -0x1.cf79abc9e3b3ap-47,
                               0x1.ffffffffffdbdp-2, 0x1.555555555543cp-3, 0x1.55555cf172b91p-5, 0x1.1111167a4d017p-7};
static const double kEmP[16] = {1.4426950408889634074, 6755399441055744.0, -6.93147180369123816490e-01,
                                -1.90821492927058770002e-10, 1.0 / 2, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040,
                                1.0 / 40320, 1.0 / 362880, 1.0 / 3628800, 1.0 / 39916800, 1.0 / 479001600,
                                1.0 / 6227020800.0};

static double expm1_ref(double x) {
  uint64_t bits;
  memcpy(&bits, &x, 8);
  const uint32_t hi = (uint32_t)(bits >> 32) & 0x7fffffffu;
  if (hi < 0x3f662e00u) {
    const double r = x, r2 = r * r;
    const double p23 = fma(r, kEmC[5], kEmC[4]);
    const double p45 = fma(r, kEmC[7], kEmC[6]);
    double tmp = fma(p23, r2, r);
    const double r4 = r2 * r2;
    tmp = fma(r4, p45, tmp);
    return (1.0 + tmp) + -1.0;
  }
  if (hi >= 0x40862800u) return x < 0.0 ? -1.0 : x * 0x1p1000;
  const double zn = fma(x, kEmP[0], kEmP[1]);
  uint64_t zb;
  memcpy(&zb, &zn, 8);
  const int n = (int)(uint32_t)zb;                       /* __double2loint */
  const double nd = zn + -kEmP[1];
  double r = fma(nd, kEmP[2], x);
  r = fma(nd, kEmP[3], r);
  double q = kEmP[15];
  for (int i = 14; i >= 4; --i) q = fma(q, r, kEmP[i]);
  const double p = fma(r * r, q, r);
  const uint64_t sb = (uint64_t)(int64_t)(1023 + n) << 52;
  double s;
  memcpy(&s, &sb, 8);
  return fma(s, p, s + -1.0);
}

static double modal(double lam, double wk, double w, double t) {
  const double x = lam * w;
  const double f = (lam == 0.0) ? t : expm1_ref(x);
  return f * wk;
}

/* ---- dense path: em_eval_dense_kernel ----------------------------------------------------------------------------
 * sigma[n][t] = the DMMA accumulator after KS = ceil(N/4) instructions; one m8n8k4 instruction adds its four products
 * to the accumulator one after the other (k ascending), each as one fused multiply-add.
 * ctab[m] = cos(pi m / 2N), m < 4N, read back from the device. */
int em_twin_dense(int N, double a, const double* lam, const double* W, const double* ctab, int T, const double* ts,
                  double* stress) {
  const int KS = (N + 3) >> 2;
  for (int j = 0; j < T; ++j) {
    const double tt = ts[j], w = a * tt;
    for (int n = 0; n < N; ++n) {
      double d = 0.0;
      for (int ks = 0; ks < KS; ++ks)
        for (int kk = 0; kk < 4; ++kk) {
          const int k = 4 * ks + kk;
          double b = 0.0, av = 0.0;
          if (k < N) {
            double wk = W[k];
            if (k > 0) wk = wk + wk;
            b = modal(lam[k], wk, w, tt);
            av = ctab[(int)(((long long)k * (2 * n + 1)) % (4LL * N))];
          }
          d = fma(av, b, d);
        }
      stress[(size_t)j * N + n] = d;
    }
  }
  return 0;
}

/* ---- stage plans: em_fft.cuh plan16 / plan_n / plan_radix / stage_tw_* / digit_pos ------------------------------- */
static int lg2r(int r) { return r == 2 ? 1 : r == 4 ? 2 : r == 8 ? 3 : 4; }
static int plan16(int p) { return p == 8 || p == 12 || p == 13; }
static int plan_n(int p) { return plan16(p) ? p / 4 + (p % 4 ? 1 : 0) : p / 3 + (p % 3 ? 1 : 0); }
static int plan_radix(int p, int i) {
  return plan16(p) ? (i < p / 4 ? 16 : (p % 4 == 1 ? 2 : p % 4 == 2 ? 4 : 8)) : (i < p / 3 ? 8 : (p % 3 == 1 ? 2 : 4));
}
static int stage_log2s(int p, int i) { int s = p; for (int j = 0; j <= i; ++j) s -= lg2r(plan_radix(p, j)); return s; }
static int stage_tw_count(int p, int i) { const int s = stage_log2s(p, i); return s == 0 ? 0 : (1 << s); }
static int stage_tw_offset(int p, int i) { int o = 0; for (int j = 0; j < i; ++j) o += stage_tw_count(p, j); return o; }
static int digit_pos(int p, int k) {
  int q = 0, sh = p;
  for (int i = 0; i < plan_n(p); ++i) {
    const int lr = lg2r(plan_radix(p, i));
    sh -= lr;
    q |= (k & ((1 << lr) - 1)) << sh;
    k >>= lr;
  }
  return q;
}

/* ---- small DFTs: em_fft.cuh Dft<2|4|8>, DftBig<16> ------------------------------------------------------------ */
static void dft2(cplx* v) { const cplx a = v[0]; v[0] = cadd(a, v[1]); v[1] = csub(a, v[1]); }
static void dft4(cplx* v, int sign) {
  const cplx a0 = cadd(v[0], v[2]), a1 = csub(v[0], v[2]);
  const cplx b0 = cadd(v[1], v[3]), b1 = mul_i(csub(v[1], v[3]), sign);
  v[0] = cadd(a0, b0); v[2] = csub(a0, b0);
  v[1] = cadd(a1, b1); v[3] = csub(a1, b1);
}
static void dft8(cplx* v, int sign) {
  cplx e[4] = {v[0], v[2], v[4], v[6]}, o[4] = {v[1], v[3], v[5], v[7]};
  dft4(e, sign);
  dft4(o, sign);
  const double h = 0.70710678118654752440, s = (double)sign;
  const cplx t1 = {o[1].x - s * o[1].y, s * o[1].x + o[1].y};
  const cplx o2 = mul_i(o[2], sign);
  const cplx t3 = {-o[3].x - s * o[3].y, s * o[3].x - o[3].y};
  v[0] = cadd(e[0], o[0]); v[4] = csub(e[0], o[0]);
  v[1].x = fma(h, t1.x, e[1].x); v[1].y = fma(h, t1.y, e[1].y);
  v[5].x = fma(-h, t1.x, e[1].x); v[5].y = fma(-h, t1.y, e[1].y);
  v[2] = cadd(e[2], o2); v[6] = csub(e[2], o2);
  v[3].x = fma(h, t3.x, e[3].x); v[3].y = fma(h, t3.y, e[3].y);
  v[7].x = fma(-h, t3.x, e[3].x); v[7].y = fma(-h, t3.y, e[3].y);
}
static double cos32(int m) {
  static const double c[9] = {1.0, 0.98078528040323044913, 0.92387953251128675613, 0.83146961230254523708,
                              0.70710678118654752440, 0.55557023301960222474, 0.38268343236508977173,
                              0.19509032201612826785, 0.0};
  m &= 31;
  if (m > 16) m = 32 - m;
  const int neg = m > 8;
  if (neg) m = 16 - m;
  return neg ? -c[m] : c[m];
}
static double sin32(int m) { return cos32(m - 8); }
static void dft16(cplx* v, int sign) {
  cplx e[8], o[8];
  for (int i = 0; i < 8; ++i) { e[i] = v[2 * i]; o[i] = v[2 * i + 1]; }
  dft8(e, sign);
  dft8(o, sign);
  for (int r = 0; r < 8; ++r) {
    cplx t;
    if (r == 0) t = o[r];
    else if (r == 4) t = mul_i(o[r], sign);
    else {
      const cplx w = {cos32(2 * r), (double)sign * sin32(2 * r)};
      t = cmul(o[r], w);
    }
    v[r] = cadd(e[r], t);
    v[r + 8] = csub(e[r], t);
  }
}
static void dft(cplx* v, int R, int sign) {
  if (R == 2) dft2(v); else if (R == 4) dft4(v, sign); else if (R == 8) dft8(v, sign); else dft16(v, sign);
}
/* tw_powers: w[q] = w1^q by the Chebyshev recurrence */
static void tw_powers(cplx w1, int R, cplx* w) {
  w[1] = w1;
  if (R >= 4) {
    const double tx = w1.x + w1.x;
    w[2].x = fma(tx, w1.x, -1.0); w[2].y = tx * w1.y;
    for (int q = 3; q < R; ++q) { w[q].x = fma(tx, w[q - 1].x, -w[q - 2].x); w[q].y = fma(tx, w[q - 1].y, -w[q - 2].y); }
  }
}
/* twiddle_jit: v[q] *= w^q with the same recurrence, nothing kept */
static void twiddle_jit(cplx* v, int R, cplx tv, int sign) {
  const double tx = tv.x + tv.x;
  cplx wp = {1.0, 0.0}, wc = {tv.x, sign > 0 ? tv.y : -tv.y};
  v[1] = cmul(v[1], wc);
  for (int q = 2; q < R; ++q) {
    const cplx wn = {fma(tx, wc.x, -wp.x), fma(tx, wc.y, -wp.y)};
    v[q] = cmul(v[q], wn);
    wp = wc; wc = wn;
  }
}

/* stage twiddles exactly as proton_em_create builds them: e^{+2 pi i j / (R S)}, long double cosl / sinl rounded once */
static cplx* build_stage_tw(int p) {
  int total = stage_tw_offset(p, plan_n(p));
  cplx* tw = (cplx*)calloc((size_t)(total > 0 ? total : 1), sizeof(cplx));
  const long double two_pi = 6.283185307179586476925286766559005768L;
  int log2s = p;
  for (int i = 0; i < plan_n(p); ++i) {
    const int R = plan_radix(p, i);
    log2s -= lg2r(R);
    const int S = 1 << log2s;
    if (S == 1) continue;
    const int off = stage_tw_offset(p, i);
    for (int j = 0; j < S; ++j) {
      const long double ang = two_pi * (long double)j / ((long double)R * (long double)S);
      tw[off + j].x = (double)cosl(ang);
      tw[off + j].y = (double)sinl(ang);
    }
  }
  return tw;
}

/* One DIT stage (em_fft.cuh fft_stage<.., DIF = false, SIGN = +1>) over the whole transform. */
static void dit_stage(cplx* data, int p, int i, const cplx* tw) {
  const int R = plan_radix(p, i), log2s = stage_log2s(p, i), S = 1 << log2s, lr = lg2r(R), NB = (1 << p) / R;
  const int twoff = stage_tw_offset(p, i);
  for (int b = 0; b < NB; ++b) {
    const int j = b & (S - 1);
    const int base = ((b >> log2s) << (log2s + lr)) | j;
    cplx v[16], w[16];
    for (int q = 0; q < R; ++q) v[q] = data[base + q * S];
    if (S > 1) {
      const cplx tv = tw[twoff + j];
      if (R == 16) twiddle_jit(v, R, tv, +1);
      else {
        tw_powers(tv, R, w);
        for (int q = 1; q < R; ++q) v[q] = cmul(v[q], w[q]);
      }
    }
    dft(v, R, +1);
    for (int q = 0; q < R; ++q) data[base + q * S] = v[q];
  }
}

/* ---- power-of-two path: em_eval_fft_kernel<p, false> -------------------------------------------------------------
 * pre[k] = e^{i pi k / 2N} read back from the device.  stress is [T][N]. */
int em_twin_pow2(int N, double a, const double* lam, const double* W, const double* pre_ri, int T, const double* ts,
                 double* stress) {
  int p = 0;
  while ((1 << p) < N) ++p;
  if ((1 << p) != N || p < 6 || p > 13) return -1;
  const cplx* pre = (const cplx*)pre_ri;
  cplx* tw = build_stage_tw(p);
  cplx* data = (cplx*)malloc((size_t)N * sizeof(cplx));
  const int NST = plan_n(p), halfN = (N + 1) >> 1, KP = (N - 1) >> 1;
  const int npairs = (T + 1) >> 1;
  for (int pr = 0; pr < npairs; ++pr) {
    const int j1 = 2 * pr;
    const double t1 = ts[j1], t2 = (j1 + 1 < T) ? ts[j1 + 1] : t1;
    const double w1 = a * t1, w2 = a * t2;
    /* phase 1: packed spectrum at the positions the DIT expects */
    for (int k = 1; k <= KP; ++k) {
      const int kk = N - k;
      const double y1k = modal(lam[k], W[k], w1, t1), y2k = modal(lam[k], W[k], w2, t2);
      const double y1kk = modal(lam[kk], W[kk], w1, t1), y2kk = modal(lam[kk], W[kk], w2, t2);
      const cplx zk = {y1k + y2kk, y2k - y1kk}, zkk = {y1kk + y2k, y2kk - y1k};
      data[digit_pos(p, k)] = cmul(pre[k], zk);
      data[digit_pos(p, kk)] = cmul(pre[kk], zkk);
    }
    {
      const double y1 = modal(lam[0], W[0], w1, t1), y2 = modal(lam[0], W[0], w2, t2);
      data[digit_pos(p, 0)].x = y1; data[digit_pos(p, 0)].y = y2;
      const int k = N >> 1;
      const double m1 = modal(lam[k], W[k], w1, t1), m2 = modal(lam[k], W[k], w2, t2);
      const cplx z = {m1 + m2, m2 - m1};
      data[digit_pos(p, k)] = cmul(pre[k], z);
    }
    /* phase 2: stages NST-1 .. 1 in place */
    for (int i = NST - 1; i >= 1; --i) dit_stage(data, p, i, tw);
    /* last stage (stride P/R0) fused with the de-interleave */
    const int R0 = plan_radix(p, 0), S = N / R0;
    double* s1 = stress + (size_t)j1 * N;
    double* s2 = (j1 + 1 < T) ? s1 + N : NULL;
    for (int j = 0; j < S; ++j) {
      cplx v[16], w[16];
      for (int q = 0; q < R0; ++q) v[q] = data[j + q * S];
      if (R0 == 16) twiddle_jit(v, 16, tw[j], +1);
      else {
        tw_powers(tw[j], 8, w);
        for (int q = 1; q < 8; ++q) v[q] = cmul(v[q], w[q]);
      }
      dft(v, R0, +1);
      for (int q = 0; q < R0; ++q) {
        const int m = j + q * S;
        const int idx = (m < halfN) ? 2 * m : 2 * (N - 1 - m) + 1;
        s1[idx] = v[q].x;
        if (s2) s2[idx] = v[q].y;
      }
    }
  }
  free(data);
  free(tw);
  return 0;
}


/* ---- any-length small DFT: em_long.cuh DftAny (powers of two by radix-2 recursion on Dft<8>, even lengths by one
 * radix-2 split, odd lengths in the symmetric form); kDftTw[R][m] = e^{+2 pi i m / R} from long double cosl / sinl ------ */
static cplx dft_tw(int R, int m) {
  const long double two_pi = 6.283185307179586476925286766559005768L;
  const long double ang = two_pi * (long double)m / (long double)R;
  cplx w = {(double)cosl(ang), (double)sinl(ang)};
  return w;
}
static void dft_big(cplx* v, int R, int sign) {            /* DftBig<R>: R a power of two <= 32 */
  if (R <= 8) { dft(v, R, sign); return; }
  cplx e[16], o[16];
  for (int i = 0; i < R / 2; ++i) { e[i] = v[2 * i]; o[i] = v[2 * i + 1]; }
  dft_big(e, R / 2, sign);
  dft_big(o, R / 2, sign);
  for (int r = 0; r < R / 2; ++r) {
    cplx t;
    if (r == 0) t = o[r];
    else if (4 * r == R) t = mul_i(o[r], sign);
    else {
      const cplx w = {cos32(r * (32 / R)), (double)sign * sin32(r * (32 / R))};
      t = cmul(o[r], w);
    }
    v[r] = cadd(e[r], t);
    v[r + R / 2] = csub(e[r], t);
  }
}
static void dft_any(cplx* v, int R, int sign) {
  if ((R & (R - 1)) == 0) { if (R > 1) dft_big(v, R, sign); return; }
  if (R % 2 == 0) {
    cplx e[16], o[16];
    for (int i = 0; i < R / 2; ++i) { e[i] = v[2 * i]; o[i] = v[2 * i + 1]; }
    dft_any(e, R / 2, sign);
    dft_any(o, R / 2, sign);
    for (int r = 0; r < R / 2; ++r) {
      cplx t = o[r];
      if (r != 0) {
        const cplx w = dft_tw(R, r);
        const double wy = sign > 0 ? w.y : -w.y;
        t.x = fma(o[r].x, w.x, -(o[r].y * wy));
        t.y = fma(o[r].x, wy, o[r].y * w.x);
      }
      v[r] = cadd(e[r], t);
      v[r + R / 2] = csub(e[r], t);
    }
    return;
  }
  const int H = (R - 1) / 2;
  cplx sm[17], df[17], y[33];
  cplx y0 = v[0];
  for (int q = 1; q <= H; ++q) { sm[q] = cadd(v[q], v[R - q]); df[q] = csub(v[q], v[R - q]); y0 = cadd(y0, sm[q]); }
  y[0] = y0;
  const double sg = (double)sign;
  for (int r = 1; r <= H; ++r) {
    cplx A = v[0], B = {0.0, 0.0};
    for (int q = 1; q <= H; ++q) {
      const cplx w = dft_tw(R, (q * r) % R);
      A.x = fma(w.x, sm[q].x, A.x); A.y = fma(w.x, sm[q].y, A.y);
      B.x = fma(w.y, df[q].x, B.x); B.y = fma(w.y, df[q].y, B.y);
    }
    y[r].x = A.x - sg * B.y; y[r].y = A.y + sg * B.x;
    y[R - r].x = A.x + sg * B.y; y[R - r].y = A.y - sg * B.x;
  }
  for (int r = 0; r < R; ++r) v[r] = y[r];
}
/* col_twiddle: v[r] *= w^r, two interleaved power chains */
static void col_twiddle(cplx* v, int R, cplx w) {
  if (R >= 2) v[1] = cmul(v[1], w);
  if (R >= 3) {
    const cplx w2 = cmul(w, w);
    cplx a = w2, b = w;
    v[2] = cmul(v[2], a);
    for (int r = 3; r < R; ++r) {
      if (r & 1) { b = cmul(b, w2); v[r] = cmul(v[r], b); }
      else { a = cmul(a, w2); v[r] = cmul(v[r], a); }
    }
  }
}

/* One DIF stage (fft_stage<.., DIF = true, SIGN = -1>) over one transform of 2^p points. */
static void dif_stage(cplx* data, int p, int i, const cplx* tw) {
  const int R = plan_radix(p, i), log2s = stage_log2s(p, i), S = 1 << log2s, lr = lg2r(R), NB = (1 << p) / R;
  const int twoff = stage_tw_offset(p, i);
  for (int b = 0; b < NB; ++b) {
    const int j = b & (S - 1);
    const int base = ((b >> log2s) << (log2s + lr)) | j;
    cplx v[16], w[16];
    for (int q = 0; q < R; ++q) v[q] = data[base + q * S];
    dft(v, R, -1);
    if (S > 1) {
      const cplx tv = tw[twoff + j];
      if (R == 16) twiddle_jit(v, R, tv, -1);
      else {
        const cplx tc = {tv.x, -tv.y};
        tw_powers(tc, R, w);
        for (int q = 1; q < R; ++q) v[q] = cmul(v[q], w[q]);
      }
    }
    for (int q = 0; q < R; ++q) data[base + q * S] = v[q];
  }
}
/* fft_middle: last DIF stage, x H, first DIT stage on blocks of `R` consecutive points */
static void middle(cplx* data, int p, const cplx* hhat) {
  const int R = plan_radix(p, plan_n(p) - 1), NB = (1 << p) / R;
  for (int b = 0; b < NB; ++b) {
    cplx v[16];
    for (int q = 0; q < R; ++q) v[q] = data[b * R + q];
    dft(v, R, -1);
    for (int q = 0; q < R; ++q) v[q] = cmul(v[q], hhat[b * R + q]);
    dft(v, R, +1);
    for (int q = 0; q < R; ++q) data[b * R + q] = v[q];
  }
}
/* modal phase of the Bluestein paths: chirped packed spectrum a_k at its natural position k < N (the rest stays unread) */
static void blue_modal(cplx* data, int N, double a, const double* lam, const double* W, const cplx* pre, double t1, double t2,
                       int long_path) {
  const double w1 = a * t1, w2 = a * t2;
  const int KP = (N - 1) >> 1;
  for (int k = 1; k <= KP; ++k) {
    const int kk = N - k;
    const double y1k = modal(lam[k], W[k], w1, t1), y2k = modal(lam[k], W[k], w2, t2);
    const double y1kk = modal(lam[kk], W[kk], w1, t1), y2kk = modal(lam[kk], W[kk], w2, t2);
    const cplx zk = {y1k + y2kk, y2k - y1kk}, zkk = {y1kk + y2k, y2kk - y1k};
    data[k] = cmul(pre[k], zk);
    data[kk] = cmul(pre[kk], zkk);
  }
  const double y1 = modal(lam[0], W[0], w1, t1), y2 = modal(lam[0], W[0], w2, t2);
  if (long_path) { data[0].x = y1 + 0.0; data[0].y = y2 - 0.0; }      /* the paired column kernel adds an absent mirror */
  else { data[0].x = y1; data[0].y = y2; }
  if ((N & 1) == 0 && N > 1) {
    const int k = N >> 1;
    const double m1 = modal(lam[k], W[k], w1, t1), m2 = modal(lam[k], W[k], w2, t2);
    const cplx z = {m1 + m2, m2 - m1};
    data[k] = cmul(pre[k], z);
  }
}

/* ---- in-CTA Bluestein, P = 2^p: em_eval_fft_kernel<p, true> ------------------------------------------------------ */
int em_twin_blue(int N, double a, const double* lam, const double* W, const double* pre_ri, const double* post_ri,
                 const double* hhat_ri, int T, const double* ts, double* stress) {
  int p = 0;
  while ((1LL << p) < 2LL * N - 1) ++p;
  if (p < 6 || p > 13) return -1;
  const int P = 1 << p, NST = plan_n(p), halfN = (N + 1) >> 1;
  const cplx *pre = (const cplx*)pre_ri, *post = (const cplx*)post_ri, *hhat = (const cplx*)hhat_ri;
  cplx* tw = build_stage_tw(p);
  cplx* data = (cplx*)calloc((size_t)P, sizeof(cplx));
  for (int pr = 0; pr < (T + 1) / 2; ++pr) {
    const int j1 = 2 * pr;
    const double t1 = ts[j1], t2 = (j1 + 1 < T) ? ts[j1 + 1] : t1;
    blue_modal(data, N, a, lam, W, pre, t1, t2, 0);
    /* first DIF stage with the zero padding implicit (fft_stage0_pruned) */
    const int R0 = plan_radix(p, 0), S = P / R0;
    for (int j = 0; j < S; ++j) {
      const cplx tv = tw[j];
      cplx v[16];
      if (R0 == 16) {
        cplx av[8], bv[8];
        for (int q = 0; q < 8; ++q) {
          if (j + q * S < N) av[q] = data[j + q * S]; else { av[q].x = 0.0; av[q].y = 0.0; }
          const cplx w = {cos32(2 * q), -sin32(2 * q)};
          bv[q] = q == 0 ? av[q] : cmul(av[q], w);
        }
        dft8(av, -1);
        dft8(bv, -1);
        for (int r = 0; r < 8; ++r) { v[2 * r] = av[r]; v[2 * r + 1] = bv[r]; }
        twiddle_jit(v, 16, tv, -1);
        for (int q = 0; q < 16; ++q) data[j + q * S] = v[q];
      } else {
        cplx w[8];
        const cplx tc = {tv.x, -tv.y};
        tw_powers(tc, 8, w);
        const cplx z = {0.0, 0.0};
        const cplx v0 = (j < N) ? data[j] : z, v1 = (j + S < N) ? data[j + S] : z;
        const cplx v2 = (j + 2 * S < N) ? data[j + 2 * S] : z, v3 = (j + 3 * S < N) ? data[j + 3 * S] : z;
        const cplx i2 = mul_i(v2, -1), i3 = mul_i(v3, -1);
        const cplx e0 = cadd(v0, v2), e2 = csub(v0, v2), e1 = cadd(v0, i2), e3 = csub(v0, i2);
        const cplx o0 = cadd(v1, v3), o2 = csub(v1, v3), o1 = cadd(v1, i3), o3 = csub(v1, i3);
        const double h = 0.70710678118654752440;
        const cplx r1 = {o1.x + o1.y, o1.y - o1.x};
        const cplx r2 = mul_i(o2, -1);
        const cplx r3 = {o3.y - o3.x, -o3.x - o3.y};
        v[0] = cadd(e0, o0); v[4] = csub(e0, o0);
        v[1].x = fma(h, r1.x, e1.x); v[1].y = fma(h, r1.y, e1.y); v[5].x = fma(-h, r1.x, e1.x); v[5].y = fma(-h, r1.y, e1.y);
        v[2] = cadd(e2, r2); v[6] = csub(e2, r2);
        v[3].x = fma(h, r3.x, e3.x); v[3].y = fma(h, r3.y, e3.y); v[7].x = fma(-h, r3.x, e3.x); v[7].y = fma(-h, r3.y, e3.y);
        for (int q = 1; q < 8; ++q) v[q] = cmul(v[q], w[q]);
        for (int q = 0; q < 8; ++q) data[j + q * S] = v[q];
      }
    }
    for (int i = 1; i <= NST - 2; ++i) dif_stage(data, p, i, tw);
    middle(data, p, hhat);
    for (int i = NST - 2; i >= 1; --i) dit_stage(data, p, i, tw);
    /* last DIT stage fused with post-chirp and de-interleave; only outputs m < N <= P/2 are formed */
    double* s1 = stress + (size_t)j1 * N;
    double* s2 = (j1 + 1 < T) ? s1 + N : NULL;
    for (int j = 0; j < S; ++j) {
      cplx v[16], w[16];
      for (int q = 0; q < R0; ++q) v[q] = data[j + q * S];
      if (R0 == 16) twiddle_jit(v, 16, tw[j], +1);
      else {
        tw_powers(tw[j], 8, w);
        for (int q = 1; q < 8; ++q) v[q] = cmul(v[q], w[q]);
      }
      dft(v, R0, +1);
      for (int q = 0; q < R0 / 2; ++q) {
        const int m = j + q * S;
        if (m < N) {
          const cplx u = cmul(v[q], post[m]);
          const int idx = (m < halfN) ? 2 * m : 2 * (N - 1 - m) + 1;
          s1[idx] = u.x;
          if (s2) s2[idx] = u.y;
        }
      }
    }
  }
  free(data);
  free(tw);
  return 0;
}

/* ---- Bluestein as `rows` rows of 2^lr points: em_eval_mix_kernel<LR, ROWS> (long_path = 0) and the long path
 * em_long_colfwd_paired_kernel -> em_long_row_kernel -> em_long_colinv_kernel (lr = 12, long_path = 1).
 * twl[j] = w_P^{+j}, j < 2^lr, and hhat (row by row, digit-reversed inside a row) are read back from the device. */
int em_twin_rows(int N, double a, const double* lam, const double* W, const double* pre_ri, const double* post_ri,
                 const double* hhat_ri, const double* twl_ri, int lr, int rows, int long_path, int T, const double* ts,
                 double* stress) {
  if (rows < 2 || rows > 32 || lr < 6 || lr > 13) return -1;
  const int R = 1 << lr, NST = plan_n(lr), halfN = (N + 1) >> 1;
  const long long P = (long long)rows * R;
  if (2LL * N - 1 > P) return -2;
  const cplx *pre = (const cplx*)pre_ri, *post = (const cplx*)post_ri, *hhat = (const cplx*)hhat_ri, *twl = (const cplx*)twl_ri;
  cplx* tw = build_stage_tw(lr);
  cplx* data = (cplx*)calloc((size_t)P, sizeof(cplx));
  const int qn = long_path ? ((rows + 1) / 2 + 1 < rows ? (rows + 1) / 2 + 1 : rows) : (rows + 1) / 2;
  for (int pr = 0; pr < (T + 1) / 2; ++pr) {
    const int j1 = 2 * pr;
    const double t1 = ts[j1], t2 = (j1 + 1 < T) ? ts[j1 + 1] : t1;
    blue_modal(data, N, a, lam, W, pre, t1, t2, long_path);
    /* forward column pass (inputs at or beyond N are zero) */
    for (int j = 0; j < R; ++j) {
      cplx v[32];
      for (int q = 0; q < rows; ++q) {
        if ((long long)j + (long long)q * R < N) v[q] = data[(size_t)q * R + j]; else { v[q].x = 0.0; v[q].y = 0.0; }
      }
      dft_any(v, rows, -1);
      const cplx wc = {twl[j].x, -twl[j].y};
      col_twiddle(v, rows, wc);
      for (int r = 0; r < rows; ++r) data[(size_t)r * R + j] = v[r];
    }
    /* rows: DIF, x H, DIT */
    for (int r = 0; r < rows; ++r) {
      cplx* row = data + (size_t)r * R;
      for (int i = 0; i <= NST - 2; ++i) dif_stage(row, lr, i, tw);
      middle(row, lr, hhat + (size_t)r * R);
      for (int i = NST - 2; i >= 0; --i) dit_stage(row, lr, i, tw);
    }
    /* inverse column pass, post-chirp, de-interleave */
    double* s1 = stress + (size_t)j1 * N;
    double* s2 = (j1 + 1 < T) ? s1 + N : NULL;
    for (int j = 0; j < R; ++j) {
      cplx v[32];
      for (int r = 0; r < rows; ++r) v[r] = data[(size_t)r * R + j];
      col_twiddle(v, rows, twl[j]);
      dft_any(v, rows, +1);
      for (int q = 0; q < qn; ++q) {
        const long long m = (long long)j + (long long)q * R;
        if (m < N) {
          const cplx u = cmul(v[q], post[m]);
          const int idx = (m < halfN) ? (int)(2 * m) : (int)(2 * (N - 1 - m) + 1);
          s1[idx] = u.x;
          if (s2) s2[idx] = u.y;
        }
      }
    }
  }
  free(data);
  free(tw);
  return 0;
}


/* ---- power-of-two long lines: em_long_scatter_kernel -> em_long_row_kernel (mode 2) -> em_long_colinv_kernel<P1, false>.
 * N = P1 x 4096; U_k goes to row k mod P1, digit-reversed place of k / P1 inside the row; rows are DIT-transformed,
 * then the column pass.  twl[j] = w_N^{+j}, j < 4096, read back from the device. */
int em_twin_long_pow2(int N, double a, const double* lam, const double* W, const double* pre_ri, const double* twl_ri,
                      int T, const double* ts, double* stress) {
  const int R = 4096, lr = 12, P1 = N / R;
  if (P1 * R != N || (P1 & (P1 - 1)) != 0 || P1 < 2 || P1 > 32) return -1;
  int log2p1 = 0;
  while ((1 << log2p1) < P1) ++log2p1;
  const cplx *pre = (const cplx*)pre_ri, *twl = (const cplx*)twl_ri;
  cplx* tw = build_stage_tw(lr);
  cplx* nat = (cplx*)calloc((size_t)N, sizeof(cplx));
  cplx* data = (cplx*)calloc((size_t)N, sizeof(cplx));
  const int NST = plan_n(lr), halfN = (N + 1) >> 1;
  for (int pr = 0; pr < (T + 1) / 2; ++pr) {
    const int j1 = 2 * pr;
    const double t1 = ts[j1], t2 = (j1 + 1 < T) ? ts[j1 + 1] : t1;
    blue_modal(nat, N, a, lam, W, pre, t1, t2, 0);
    for (int k = 0; k < N; ++k) data[(size_t)(k & (P1 - 1)) * R + digit_pos(lr, k >> log2p1)] = nat[k];
    for (int r = 0; r < P1; ++r)
      for (int i = NST - 1; i >= 0; --i) dit_stage(data + (size_t)r * R, lr, i, tw);
    double* s1 = stress + (size_t)j1 * N;
    double* s2 = (j1 + 1 < T) ? s1 + N : NULL;
    for (int j = 0; j < R; ++j) {
      cplx v[32];
      for (int r = 0; r < P1; ++r) v[r] = data[(size_t)r * R + j];
      col_twiddle(v, P1, twl[j]);
      dft_any(v, P1, +1);
      for (int q = 0; q < P1; ++q) {
        const int m = j + q * R;
        const int idx = (m < halfN) ? 2 * m : 2 * (N - 1 - m) + 1;
        s1[idx] = v[q].x;
        if (s2) s2[idx] = v[q].y;
      }
    }
  }
  free(nat);
  free(data);
  free(tw);
  return 0;
}
