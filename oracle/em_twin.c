/* oracle/em_twin.c — TEST INFRASTRUCTURE ONLY: a deterministic CPU twin of the GPU ARITHMETIC.
 *
 * em_oracle.c restates the reference's algorithm (FFTW-style DCTs); two different DCT algorithms cannot agree in
 * the last bit, so the GPU path is compared with it norm-wise (1e-9).  This file restates instead what the CUDA
 * kernels themselves do — the same butterflies, in the same order, with the same fused multiply-adds — so that the
 * stress field, and with it max / argmax / flag, can be required to be BIT-IDENTICAL (SURVEY.md §7 hard part 2).
 * Twins (tests/test_twin.py):
 *   em_twin_dense   em_eval_dense_kernel        proton_b200/csrc/em_kernels.cuh   (N <= 88: DMMA m8n8k4 chains)
 *   em_twin_pow2    em_eval_fft_kernel<p,false>  proton_b200/csrc/em_fft.cuh       (N = 2^p, 64..8192: in-CTA DIT FFT)
 * Inputs that come from the DEVICE's sincospi (pre-twiddles, cosine table, modal weights W) are read back through
 * proton_em_debug_line_table and handed in; lambda is the host-libm table both sides share; the stage twiddles are
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
static int plan16(int p) { return p == 12 || p == 13; }
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
