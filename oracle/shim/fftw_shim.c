/* oracle/shim/fftw_shim.c — TEST INFRASTRUCTURE ONLY.
 * Stand-in for the 4 FFTW3 entry points the reference binary imports from MKL's FFTW3 wrapper
 * (`nm -D bin/EMtool_analytical`: fftw_plan_r2r_1d, fftw_execute, fftw_destroy_plan,
 * fftw_cleanup; call sites @0xc1ca, @0xc3f7). Transforms are delegated to em_oracle_r2r
 * (oracle/em_oracle.c). EM_SHIM_FFT=double selects the double-precision O(N log N) transform
 * (used when the binary is TIMED as the CPU reference); default is the long-double checker. */
#include <stdlib.h>
#include <string.h>

int em_oracle_r2r(int n, const double* in, double* out, int kind, int use_double);

typedef struct shim_plan { int n; double* in; double* out; int kind; } shim_plan;

static int use_double_fft(void) {
  static int cached = -1;
  if (cached < 0) { const char* e = getenv("EM_SHIM_FFT"); cached = (e && strcmp(e, "double") == 0) ? 1 : 0; }
  return cached;
}

void* fftw_plan_r2r_1d(int n, double* in, double* out, int kind, unsigned flags) {
  (void)flags;
  shim_plan* p = (shim_plan*)malloc(sizeof(shim_plan));
  p->n = n; p->in = in; p->out = out; p->kind = kind;
  return p;
}
void fftw_execute(const void* plan) {
  const shim_plan* p = (const shim_plan*)plan;
  if (em_oracle_r2r(p->n, p->in, p->out, p->kind, use_double_fft()) != 0) abort();
}
void fftw_destroy_plan(void* plan) { free(plan); }
void fftw_cleanup(void) {}
