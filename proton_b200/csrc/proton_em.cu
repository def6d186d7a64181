// proton_em.cu — host side of libproton_em.so: the C ABI declared in include/proton_em.h.
//
// Owns the context (streams, constant tables, per-length table cache), the plan (a packed ragged batch
// resident in HBM) and the launch logic that routes each line to a kernel by its length:
//   N <= 88 (not 64)             em_eval_dense_kernel<MT>       (DMMA contraction with the N x N cosine matrix)
//   N = 2^p, 64..8192            em_eval_fft_kernel<p, false>   (in-CTA DIT FFT, radix 16/8/4/2)
//   other N <= 4096              em_eval_fft_kernel<p, true>    (Bluestein, P = 2^p >= 2N-1, P <= 4096), or
//                                em_eval_mix_kernel<LR, ROWS>   (Bluestein, P = ROWS x 2^LR, 5..16 rows of 128 / 256 / 512
//                                                                points: the admissible class of least measured cost)
//   other N <= 65536, 2^p <= 131072   em_long_{colfwd_paired | scatter} -> em_long_row -> em_long_colinv
//                                (P = P1 x 4096 through an L2-resident scratch, waves dealt to `lanes` streams)
// Tuning knobs, all optional environment variables read once: PROTON_EM_LANES (2), PROTON_EM_LONG_WAVE_MB (32),
// PROTON_EM_PDL (1), PROTON_EM_COL_DIV (32), PROTON_EM_DENSE_MAX (88), PROTON_EM_CHUNK_MB (256), PROTON_EM_MIX (1),
// PROTON_EM_TRACE.
// There is no CPU fallback: every entry point fails with PROTON_EM_ECUDA when no device is usable.
#include "../../include/proton_em.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <sstream>
#include <string>
#include <thread>
#include <map>
#include <new>
#include <unordered_map>
#include <vector>
#include <sys/stat.h>
#include <cerrno>

#include "em_kernels.cuh"
#include "em_long.cuh"
#include "em_mix.cuh"

namespace {

thread_local std::string g_last_error;

struct Trace {
  bool on;
  std::chrono::steady_clock::time_point t0;
  const char* what;
  explicit Trace(const char* w) : on(std::getenv("PROTON_EM_TRACE") != nullptr), t0(std::chrono::steady_clock::now()), what(w) {}
  void lap(const char* label) {
    if (!on) return;
    const auto t1 = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[proton_em] %s: %s %.3f ms\n", what, label, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};

int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}

// No exception leaves the C ABI (include/proton_em.h: "nothing here throws"): the entry points that build host-side
// containers are function-try-blocks that end in this.
#define EM_ABI_CATCH(name)                                                                          \
  catch (const std::bad_alloc&) { return fail(PROTON_EM_ENOMEM, name ": out of host memory"); }    \
  catch (const std::exception& e) { return fail(PROTON_EM_EINVAL, name ": %s", e.what()); }

#define CU_TRY(expr)                                                                              \
  do {                                                                                            \
    cudaError_t e_ = (expr);                                                                      \
    if (e_ != cudaSuccess)                                                                        \
      return fail(e_ == cudaErrorMemoryAllocation ? PROTON_EM_ENOMEM : PROTON_EM_ECUDA,           \
                  "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__);    \
  } while (0)

const unsigned long long kExpTable[2 * em::kExpTableEntries] = {
#include "exp_table.inc"
};

constexpr int kMinLog2P = 6, kMaxLog2P = 13;
constexpr int kMaxPow2N = 1 << kMaxLog2P;          // 8192
constexpr int kMaxBlueN = (1 << (kMaxLog2P - 1));  // 4096: 2N-1 <= 8192

// Long lines: P = P1 * 4096 with P1 from this list (any P1 works for Bluestein; powers of two for 2^p lines).
constexpr int kLongP1List[] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 26, 28, 30, 32};
// Measured Gevals/s of a full transform per row count (B200, T = 100, tools/profile_step.py --uniform 2048*P1-10):
// odd and large row counts cost more in the column passes, so e.g. 16 rows beat 15 and 32 beat 28 and 30
// (PROTON_EM_P1_SMALLEST=1 forces the smallest fitting count, for re-measuring this table).
constexpr double kLongP1Gevals[] = {52.5, 53.0, 52.8, 52.2, 51.7, 52.0, 50.2, 51.5, 47.1, 50.0, 45.7, 45.8, 40.5, 48.4,
                                    43.8, 40.8, 35.5, 35.2, 34.1, 33.9, 32.1, 37.7};
static_assert(sizeof(kLongP1Gevals) / sizeof(double) == sizeof(kLongP1List) / sizeof(int), "one figure per row count");
constexpr int kMaxLongPow2N = em::kLongP2 * em::kLongMaxP1;                 // 131072
constexpr int kMaxLongBlueN = (em::kLongP2 * em::kLongMaxP1) / 2;           // 65536: 2N-1 <= 131072

enum LineClass { CLS_SMALL = 0, CLS_POW2 = 1, CLS_BLUE = 2, CLS_LPOW2 = 3, CLS_LBLUE = 4, CLS_MIX = 5, CLS_NONE = 6 };

// log2p = log2 of the FFT length for the in-CTA classes; for the long classes it holds P1 (rows of 4096 points).
struct ClassOf { LineClass cls; int log2p; };

// Lines up to this length take the dense (DMMA) path; the split against the FFT path is set from measured
// throughput, PROTON_EM_DENSE_MAX overrides it for experiments.
int small_max_n() {
  static const int v = [] {
    const char* e = std::getenv("PROTON_EM_DENSE_MAX");
    const int x = e ? std::atoi(e) : 0;
    return x > 0 ? std::min(x, (int)em::kDenseMaxN) : 88;
  }();
  return v;
}

// The mixed-radix in-CTA classes that are built (rows of 2^lr points x odd rows, em_mix.cuh), each with its measured
// cost: length of the longest line it takes / Gevals/s there (B200, T = 256, profiles/r02_class_rates.txt) — a time per
// transform in arbitrary units.  A line goes to the admissible class (P >= 2N - 1) of least cost, power-of-two lengths
// (kBlueCost) included.  Row counts that lost to the next longer class everywhere were taken out again:
// 13 and 15 rows of 256 or 512 points (odd column DFTs cost O(rows^2) FP64 work; 14 x 256, 4096 and 7168, 8192 points
// are cheaper).
#define FOR_EACH_MIX(M)                                                                                               \
  M(7, 5, 3.79) M(7, 6, 4.19) M(8, 5, 7.33) M(8, 6, 8.14) M(8, 7, 9.48) M(8, 9, 13.58) M(8, 11, 17.26)                 \
  M(8, 14, 22.04) M(9, 5, 14.58) M(9, 6, 18.00) M(9, 9, 29.51) M(9, 10, 31.20) M(9, 11, 36.34) M(9, 12, 38.67)         \
  M(9, 14, 49.46) M(9, 16, 52.63)
struct MixClass { int lr, odd; double cost; };
const MixClass kMixClasses[] = {
#define M(lr, odd, cost) {lr, odd, cost},
    FOR_EACH_MIX(M)
#undef M
};
// the same for em_eval_fft_kernel<p, true>, P = 2^p (p <= 9 has no competitor: nominal)
inline double blue_cost(int p) {
  switch (p) {
    case 10: return 5.63;
    case 11: return 11.83;
    case 12: return 22.52;
    case 13: return 58.2;
    default: return 0.0055 * (double)(1 << p);
  }
}

ClassOf classify(int N) {
  if (N <= small_max_n() && !(N >= 64 && (N & (N - 1)) == 0)) return {CLS_SMALL, 0};
  if ((N & (N - 1)) == 0) {
    if (N > kMaxLongPow2N) return {CLS_NONE, 0};
    if (N > kMaxPow2N) return {CLS_LPOW2, N / em::kLongP2};
    int p = 0;
    while ((1 << p) < N) ++p;
    return {CLS_POW2, p};
  }
  if (N > kMaxLongBlueN) return {CLS_NONE, 0};
  if (N > kMaxBlueN) {
    // any row count with P1 * 4096 >= 2N-1 works; take the one with the least time ~ P1 / throughput(P1)
    const long long need = 2LL * N - 1;
    int best = 0;
    double best_cost = 0.0;
    for (size_t i = 0; i < sizeof(kLongP1List) / sizeof(kLongP1List[0]); ++i) {
      const int p1 = kLongP1List[i];
      if ((long long)p1 * em::kLongP2 < need) continue;
      static const bool smallest = [] { const char* e = std::getenv("PROTON_EM_P1_SMALLEST"); return e && std::atoi(e) != 0; }();
      const double cost = smallest ? (double)p1 : (double)p1 / kLongP1Gevals[i];   // (override: for re-measuring the table)
      if (!best || cost < best_cost) { best = p1; best_cost = cost; }
    }
    return best ? ClassOf{CLS_LBLUE, best} : ClassOf{CLS_NONE, 0};
  }
  int p = 0;
  while ((1LL << p) < 2LL * N - 1) ++p;
  // mixed-radix lengths (em_mix.cuh) when one is shorter than the power of two; PROTON_EM_MIX=0 keeps the
  // power-of-two lengths (for A/B runs)
  static const bool use_mix = [] { const char* e = std::getenv("PROTON_EM_MIX"); return !(e && std::atoi(e) == 0); }();
  if (use_mix) {
    const long long need = 2LL * N - 1;
    double best = blue_cost(p);
    int code = 0;
    for (const MixClass& m : kMixClasses) {
      const long long P = (long long)m.odd << m.lr;
      if (P >= need && m.cost < best) { best = m.cost; code = em::mix_code(m.lr, m.odd); }
    }
    if (code) return {CLS_MIX, code};
  }
  return {CLS_BLUE, p};
}
inline int mix_lr(int code) { return code / 32 + 6; }
inline int mix_odd(int code) { return code % 32; }

// Bump allocator for the per-length tables (lives as long as the context).
struct Arena {
  std::vector<void*> blocks;
  char* cur = nullptr;
  size_t left = 0;
  size_t total = 0;
  cudaError_t alloc(void** out, size_t bytes) {
    bytes = (bytes + 255) & ~size_t(255);
    if (bytes > left) {
      const size_t blk = std::max(bytes, size_t(32) << 20);
      void* p = nullptr;
      cudaError_t e = cudaMalloc(&p, blk);
      if (e != cudaSuccess) return e;
      blocks.push_back(p);
      cur = static_cast<char*>(p);
      left = blk;
      total += blk;
    }
    *out = cur;
    cur += bytes;
    left -= bytes;
    return cudaSuccess;
  }
  // One block for everything a batch of new tables needs, instead of a cudaMalloc per 32 MB.
  cudaError_t reserve(size_t bytes) {
    if (bytes <= left) return cudaSuccess;
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return e;
    blocks.push_back(p);
    cur = static_cast<char*>(p);
    left = bytes;
    total += bytes;
    return cudaSuccess;
  }
  void release() {
    for (void* p : blocks) cudaFree(p);
    blocks.clear();
    cur = nullptr;
    left = 0;
  }
};

// Device buffers of plans and of the host-buffer paths come from a per-context CUDA memory pool, stream-ordered on
// the context's main stream (every other stream the library uses forks from and joins into it): repeated
// proton_em_analyze_host calls then cost microseconds instead of ~4 ms of cudaMalloc / cudaFree per call.  The entry
// points name the pool and stream of their context here (set_device); PROTON_EM_ASYNC_ALLOC=0 falls back to cudaMalloc.
thread_local cudaMemPool_t tl_pool = nullptr;
thread_local cudaStream_t tl_stream = nullptr;

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  bool pooled = false;
  cudaError_t alloc(size_t count) {
    free();
    n = count;
    if (count == 0) return cudaSuccess;
    if (tl_pool) {
      pooled = true;
      return cudaMallocFromPoolAsync(reinterpret_cast<void**>(&p), count * sizeof(T), tl_pool, tl_stream);
    }
    pooled = false;
    return cudaMalloc(reinterpret_cast<void**>(&p), count * sizeof(T));
  }
  void free() {
    if (p) {
      if (pooled && tl_pool) cudaFreeAsync(p, tl_stream);
      else cudaFree(p);
    }
    p = nullptr;
    n = 0;
  }
  size_t bytes() const { return n * sizeof(T); }
};

}  // namespace

struct proton_em_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaMemPool_t pool = nullptr;             // stream-ordered pool behind every DevBuf of this context (may be null)
  ulonglong2* d_exp = nullptr;
  em::cplx* d_tw[kMaxLog2P + 1] = {};
  int* d_pos[kMaxLog2P + 1] = {};           // padded DIT input positions per FFT length (em_fft.cuh: pos_tab)
  Arena arena;
  std::unordered_map<int, em::NTab> tabs;   // per distinct N, device pointers into the arena
  int64_t launches = 0;
  // long-line path: column twiddles per P1 and the wave scratch (grown on demand, reused by every call)
  em::cplx* d_twl[em::kLongMaxP1 + 1] = {};   // column twiddles per P1
  em::cplx* d_twm[192] = {};                   // column twiddles per mixed class (em_mix.cuh), by mix_code
  static constexpr int kLanes = 8;          // up to 8 streams; `lanes` of them are used (PROTON_EM_LANES, default 2:
                                            // two 32 MB waves in flight measured best, the scratch stays well inside L2)
  int lanes = 2;
  cudaStream_t lane[kLanes] = {};
  cudaEvent_t ev_fork = nullptr, ev_join[kLanes] = {};
  // host-buffer path (proton_em_analyze_host): cached device staging, a copy stream, per-buffer events
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_done[2] = {}, ev_copied[2] = {};
  DevBuf<double> io_stress[2], io_mx, io_ts, io_tmp;
  DevBuf<int> io_amx;
  DevBuf<unsigned char> io_flag;
  DevBuf<em::cplx> long_x, long_y;
  DevBuf<double> long_part_v;
  DevBuf<int> long_part_i;
  // per-kernel timing (proton_em_profile): an event pair around every launch of the evaluation path
  bool prof_on = false;
  struct ProfRec { int kind; cudaEvent_t a, b; };
  std::vector<ProfRec> prof_recs;
  std::vector<cudaEvent_t> prof_pool;
  double prof_ms[16] = {};
  long long prof_n[16] = {};
};

namespace {
enum ProfKind { PK_DENSE = 0, PK_FFT_POW2, PK_FFT_BLUE, PK_FFT_MIX, PK_LONG_COLFWD, PK_LONG_ROW, PK_LONG_COLINV,
                PK_LONG_SCATTER, PK_LONG_REDUCE, PK_COUNT };
const char* const kProfNames[PK_COUNT] = {"em_eval_dense_kernel", "em_eval_fft_kernel<pow2>",
                                          "em_eval_fft_kernel<bluestein>", "em_eval_mix_kernel", "em_long_colfwd_paired_kernel",
                                          "em_long_row_kernel", "em_long_colinv_kernel", "em_long_scatter_kernel",
                                          "em_long_reduce_kernel"};
// Records an event before and after one launch on the stream it is launched on (only while profiling is on).
struct ProfScope {
  proton_em_ctx* ctx;
  cudaStream_t st;
  cudaEvent_t b = nullptr;
  ProfScope(proton_em_ctx* c, int kind, cudaStream_t s) : ctx(c), st(s) {
    if (!c->prof_on) return;
    cudaEvent_t ev[2] = {nullptr, nullptr};
    for (cudaEvent_t& e : ev) {
      if (!c->prof_pool.empty()) { e = c->prof_pool.back(); c->prof_pool.pop_back(); }
      else if (cudaEventCreate(&e) != cudaSuccess) { e = nullptr; }
    }
    if (!ev[0] || !ev[1]) return;
    cudaEventRecord(ev[0], s);
    b = ev[1];
    c->prof_recs.push_back({kind, ev[0], ev[1]});
  }
  ~ProfScope() { if (b) cudaEventRecord(b, st); }
};
}  // namespace

struct proton_em_plan {
  proton_em_ctx* ctx = nullptr;
  int64_t L = 0, total_nodes = 0, ntrip = 0, nu = 0;
  int max_nodes = 0;
  bool owns_inputs = false;
  // batch arrays on device (owned when the batch came from the host)
  const int* d_num_nodes = nullptr;
  const double* d_a = nullptr;
  const long long* d_trip_off = nullptr;
  DevBuf<int> own_num_nodes;
  DevBuf<double> own_a;
  DevBuf<long long> own_trip_off;
  // derived
  DevBuf<long long> d_node_off, d_blk_off;
  DevBuf<int> d_r_idx, d_r_cnt, d_line_tab;
  DevBuf<double> d_r_val, d_W;
  DevBuf<em::NTab> d_tabs;
  DevBuf<int> d_class_lines;     // concatenated per-class line lists
  DevBuf<double> d_sigma0;       // initial stress per node (+2 doubles of slack), empty when the plan starts from zero stress
  struct ClassList { LineClass cls; int log2p; int offset; int count; };
  // Lines are evaluated in chunks of contiguous lines (one chunk unless the host path asked for more, so
  // that the device-to-host copy of one chunk overlaps the evaluation of the next).
  struct Chunk { int64_t l0, l1; std::vector<ClassList> classes; };
  std::vector<Chunk> chunks;
  std::vector<int> h_class_lines;    // host copy of d_class_lines
  // long-line transform lists (slot -> line, timestamp pair): they depend on (T, chunk range) only, so they are
  // built and uploaded once and reused by every later evaluation of the plan
  struct LongList { int T, c0, c1; long long n; DevBuf<em::SlotRec> recs; std::vector<em::SlotRec> host; };
  std::vector<LongList*> long_lists;
  std::vector<long long> node_off;   // host copy
  std::vector<int> num_nodes;        // host copy
  std::vector<double> a_host;        // host copy of a_coeff
  std::vector<std::string> names;    // optional line names (proton_em_plan_set_line_names / proton_em_pack_from_project)
};

namespace {

int set_device(proton_em_ctx* ctx) {
  CU_TRY(cudaSetDevice(ctx->device));
  tl_pool = ctx->pool;
  tl_stream = ctx->stream;
  return PROTON_EM_OK;
}

template <int LOG2P, bool BLUE>
int set_smem_attr() {
  CU_TRY(cudaFuncSetAttribute(em::em_eval_fft_kernel<LOG2P, BLUE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)em::Cfg<LOG2P, BLUE>::smem_bytes));
  if (BLUE)
    CU_TRY(cudaFuncSetAttribute(em::em_tables_kernel<LOG2P, BLUE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)((em::padded_len<LOG2P>() + em::stage_tw_total(LOG2P)) * sizeof(em::cplx))));
  return PROTON_EM_OK;
}

#define FOR_EACH_LOG2P(M) M(6) M(7) M(8) M(9) M(10) M(11) M(12) M(13)

int init_kernel_attrs() {
#define M(p)                                                          \
  { int rc = set_smem_attr<p, false>(); if (rc) return rc; }          \
  { int rc = set_smem_attr<p, true>(); if (rc) return rc; }
  FOR_EACH_LOG2P(M)
#undef M
  return PROTON_EM_OK;
}

template <bool BLUE>
int launch_tables(proton_em_ctx* ctx, int log2p, const em::NTab* d_entries, int count) {
  em::TableParams tp{d_entries, ctx->d_tw[log2p]};
  switch (log2p) {
#define M(p)                                                                                              \
  case p:                                                                                                 \
    em::em_tables_kernel<p, BLUE><<<count, em::Cfg<p, BLUE>::NT,                                          \
                                    BLUE ? (em::padded_len<p>() + em::stage_tw_total(p)) * sizeof(em::cplx) : 0, ctx->stream>>>(tp); \
    break;
    FOR_EACH_LOG2P(M)
#undef M
    default: return fail(PROTON_EM_EINVAL, "bad log2p %d", log2p);
  }
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

int launch_eval_fft(proton_em_ctx* ctx, cudaStream_t st, bool blue, int log2p, const em::EvalParams& ep, int grid) {
  ProfScope prof(ctx, blue ? PK_FFT_BLUE : PK_FFT_POW2, st);
  switch (log2p) {
#define M(p)                                                                                                       \
  case p:                                                                                                          \
    if (blue) em::em_eval_fft_kernel<p, true><<<grid, em::Cfg<p, true>::NT, em::Cfg<p, true>::smem_bytes, st>>>(ep); \
    else em::em_eval_fft_kernel<p, false><<<grid, em::Cfg<p, false>::NT, em::Cfg<p, false>::smem_bytes, st>>>(ep);  \
    break;
    FOR_EACH_LOG2P(M)
#undef M
    default: return fail(PROTON_EM_EINVAL, "bad log2p %d", log2p);
  }
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

int init_mix_attrs() {
#define M(lr, odd, cost)                                                                                                    \
  CU_TRY(cudaFuncSetAttribute(em::em_eval_mix_kernel<lr, odd>, cudaFuncAttributeMaxDynamicSharedMemorySize,           \
                              (int)em::MixCfg<lr, odd>::smem_bytes));                                                 \
  CU_TRY(cudaFuncSetAttribute(em::em_tables_mix_kernel<lr, odd>, cudaFuncAttributeMaxDynamicSharedMemorySize,         \
                              (int)em::MixCfg<lr, odd>::smem_bytes));
  FOR_EACH_MIX(M)
#undef M
  return PROTON_EM_OK;
}

// column twiddles w_P^{+j}, j < R, of a mixed class (built at first use, on the context stream)
int ensure_mix(proton_em_ctx* ctx, int code) {
  if (ctx->d_twm[code]) return PROTON_EM_OK;
  const int lr = mix_lr(code), odd = mix_odd(code), R = 1 << lr;
  CU_TRY(cudaMalloc(reinterpret_cast<void**>(&ctx->d_twm[code]), (size_t)R * sizeof(em::cplx)));
  em::em_mix_twiddle_kernel<<<(R + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_twm[code], R, odd * R);
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

int launch_tables_mix(proton_em_ctx* ctx, int code, const em::NTab* d_entries, int count) {
  { int rc = ensure_mix(ctx, code); if (rc) return rc; }
  const int lr = mix_lr(code), odd = mix_odd(code);
  em::TableParams tp{d_entries, ctx->d_tw[lr]};
  bool done = false;
#define M(l, o, c)                                                                                                      \
  if (lr == l && odd == o) {                                                                                          \
    em::em_tables_mix_kernel<l, o><<<count, em::MixCfg<l, o>::NT, em::MixCfg<l, o>::smem_bytes, ctx->stream>>>(tp, ctx->d_twm[code]); \
    done = true;                                                                                                      \
  }
  FOR_EACH_MIX(M)
#undef M
  if (!done) return fail(PROTON_EM_EINVAL, "bad mixed class %d", code);
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

int launch_eval_mix(proton_em_ctx* ctx, cudaStream_t st, int code, const em::EvalParams& ep, int grid) {
  ProfScope prof(ctx, PK_FFT_MIX, st);
  const int lr = mix_lr(code), odd = mix_odd(code);
  em::MixParams mp{ep, ctx->d_twm[code]};
  bool done = false;
#define M(l, o, c)                                                                                                      \
  if (lr == l && odd == o) {                                                                                          \
    em::em_eval_mix_kernel<l, o><<<grid, em::MixCfg<l, o>::NT, em::MixCfg<l, o>::smem_bytes, st>>>(mp);               \
    done = true;                                                                                                      \
  }
  FOR_EACH_MIX(M)
#undef M
  if (!done) return fail(PROTON_EM_EINVAL, "bad mixed class %d", code);
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

int fft_group(bool blue, int log2p) {
  switch (log2p) {
#define M(p) case p: return blue ? em::Cfg<p, true>::G : em::Cfg<p, false>::G;
    FOR_EACH_LOG2P(M)
#undef M
  }
  return 1;
}

// ===================================== long-line path ==========================================
#define FOR_EACH_P1(M) M(3) M(4) M(5) M(6) M(7) M(8) M(9) M(10) M(11) M(12) M(13) M(14) M(15) M(16) M(18) M(20) M(22) M(24) M(26) M(28) M(30) M(32)
#define FOR_EACH_LOG2P1(M) M(2) M(3) M(4) M(5)

constexpr size_t kLongRowSmem = (em::padded_len<em::kLongLog2P2>() + em::stage_tw_total(em::kLongLog2P2)) * sizeof(em::cplx);
inline size_t long_col_smem(int p1) { return (size_t)p1 * em::kLongColsPerCta * sizeof(em::cplx); }
constexpr int kLongColBlocks = em::kLongP2 / em::kLongColsPerCta;
inline size_t long_pair_smem(int p1) { return ((size_t)(p1 + 1) / 2 + 1) * 128 * sizeof(em::cplx); }

int init_long_attrs() {
  CU_TRY(cudaFuncSetAttribute(em::em_long_row_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kLongRowSmem));
#define M(q)                                                                                                              \
  CU_TRY(cudaFuncSetAttribute(em::em_long_colfwd_kernel<q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)long_col_smem(q)));        \
  CU_TRY(cudaFuncSetAttribute(em::em_long_colfwd_paired_kernel<q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)long_pair_smem(q)));
  FOR_EACH_P1(M)
#undef M
  // twiddles of the non-power-of-two column DFTs
  std::vector<em::cplx> tw((size_t)(em::kLongMaxP1 + 1) * em::kLongMaxP1, em::cplx{0.0, 0.0});
  const long double two_pi = 6.283185307179586476925286766559005768L;
  for (int r = 1; r <= em::kLongMaxP1; ++r)
    for (int m = 0; m < r; ++m) {
      const long double ang = two_pi * (long double)m / (long double)r;
      tw[(size_t)r * em::kLongMaxP1 + m] = em::cplx{(double)cosl(ang), (double)sinl(ang)};
    }
  CU_TRY(cudaMemcpyToSymbol(em::kDftTw, tw.data(), tw.size() * sizeof(em::cplx)));
  return PROTON_EM_OK;
}

// Transforms per wave (PROTON_EM_LONG_WAVE_MB megabytes of work buffer per stream, default 32).
int long_wave_capacity(int p1) {
  static long long budget_mb = [] {
    const char* e = std::getenv("PROTON_EM_LONG_WAVE_MB");
    const long long v = e ? std::atoll(e) : 0;
    return v > 0 ? v : 32LL;
  }();
  const long long bytes = (long long)p1 * em::kLongP2 * (long long)sizeof(em::cplx);
  return (int)std::max<long long>(4, std::min<long long>(1024, (budget_mb << 20) / bytes));
}

int ensure_long(proton_em_ctx* ctx, int p1, int cap, long long n_tr_total) {
  if (p1 < 3 || p1 > em::kLongMaxP1) return fail(PROTON_EM_EINVAL, "bad long row count %d", p1);
  if (!ctx->d_twl[p1]) {
    const size_t n = (size_t)p1 * em::kLongP2;
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&ctx->d_twl[p1]), n * sizeof(em::cplx)));
    em::em_long_twiddle_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_twl[p1], p1);
    ctx->launches++;
    CU_TRY(cudaGetLastError());        // (the lanes fork from the context stream at every evaluation, so they see it)
  }
  const size_t lanes = (size_t)ctx->lanes;
  const size_t need_x = lanes * (size_t)cap * (size_t)p1 * em::kLongP2;
  const size_t need_p = (size_t)std::max<long long>(n_tr_total, 1) * 2 * kLongColBlocks;
  if (ctx->long_x.n < need_x || ctx->long_part_v.n < need_p) {
    // Earlier waves may still use the old buffers: with the pool, release and allocation are ordered on the context
    // stream, which every evaluation's lanes have joined; without it cudaFree synchronises by itself.
    if (ctx->long_x.n < need_x) CU_TRY(ctx->long_x.alloc(need_x));
    if (ctx->long_part_v.n < need_p) { CU_TRY(ctx->long_part_v.alloc(need_p)); CU_TRY(ctx->long_part_i.alloc(need_p)); }
  }
  return PROTON_EM_OK;
}

// Launch with the programmatic-stream-serialization attribute (PDL, see em_long.cuh); PROTON_EM_PDL=0 turns it off.
bool pdl_enabled() {
  static const bool v = [] { const char* e = std::getenv("PROTON_EM_PDL"); return !(e && std::atoi(e) == 0); }();
  return v;
}
template <typename Param>
cudaError_t launch_pdl(void (*kern)(const Param), dim3 grid, unsigned block, size_t smem, cudaStream_t st, const Param& prm) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, prm);
}

int launch_long_colfwd(proton_em_ctx* ctx, cudaStream_t st, int p1, const em::ColParams& cp, int groups) {
  const dim3 grid(kLongColBlocks, groups);
  switch (p1) {
#define M(q) case q: em::em_long_colfwd_kernel<q><<<grid, em::kLongColsPerCta, long_col_smem(q), st>>>(cp); break;
    FOR_EACH_P1(M)
#undef M
    default: return fail(PROTON_EM_EINVAL, "unsupported row count %d", p1);
  }
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

int launch_long_colfwd_paired(proton_em_ctx* ctx, cudaStream_t st, int p1, const em::ColParams& cp, int groups) {
  const dim3 grid(em::kLongPairBlocks, groups);
  ProfScope prof(ctx, PK_LONG_COLFWD, st);
  switch (p1) {
#define M(q) case q: CU_TRY(launch_pdl(em::em_long_colfwd_paired_kernel<q>, grid, 128, long_pair_smem(q), st, cp)); break;
    FOR_EACH_P1(M)
#undef M
    default: return fail(PROTON_EM_EINVAL, "unsupported row count %d", p1);
  }
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

int launch_long_colinv(proton_em_ctx* ctx, cudaStream_t st, int p1, bool blue, const em::ColParams& cp, int groups) {
  const dim3 grid(kLongColBlocks, groups);
  ProfScope prof(ctx, PK_LONG_COLINV, st);
  switch (p1) {
#define M(q)                                                                                              \
  case q:                                                                                                 \
    if (blue) CU_TRY(launch_pdl(em::em_long_colinv_kernel<q, true>, grid, em::kLongColsPerCta, 0, st, cp)); \
    else CU_TRY(launch_pdl(em::em_long_colinv_kernel<q, false>, grid, em::kLongColsPerCta, 0, st, cp));     \
    break;
    FOR_EACH_P1(M)
#undef M
    default: return fail(PROTON_EM_EINVAL, "unsupported row count %d", p1);
  }
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

int launch_long_scatter(proton_em_ctx* ctx, cudaStream_t st, int p1, const em::LongWave& w, int maxN) {
  const dim3 grid(std::min(256, (maxN / 2 + 256) / 256), w.n_tr);     // one thread per mode pair (k, N-k)
  ProfScope prof(ctx, PK_LONG_SCATTER, st);
  switch (p1) {
#define M(q) case (1 << q): CU_TRY(launch_pdl(em::em_long_scatter_kernel<q>, grid, 256, 0, st, w)); break;
    FOR_EACH_LOG2P1(M)
#undef M
    default: return fail(PROTON_EM_EINVAL, "unsupported row count %d", p1);
  }
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

int launch_long_row(proton_em_ctx* ctx, cudaStream_t st, const em::RowParams& rp, int n_tr) {
  const dim3 grid(rp.p1, n_tr);
  ProfScope prof(ctx, PK_LONG_ROW, st);
  CU_TRY(launch_pdl(em::em_long_row_kernel, grid, em::kLongRowThreads, kLongRowSmem, st, rp));
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

// pre / post chirps and the transformed chirp H of new long lengths (`d_e`: device array of `count` entries that share
// the row count and the kind).  Queued on the context stream, no host synchronisation.
int long_build_tables(proton_em_ctx* ctx, const em::NTab* d_e, int count, int lp1, bool blue) {
  const int cap = long_wave_capacity(lp1);
  { int rc = ensure_long(ctx, lp1, cap, 1); if (rc) return rc; }
  em::em_long_prepost_kernel<<<dim3(64, (unsigned)count), 256, 0, ctx->stream>>>(d_e, blue ? 1 : 0);
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  if (blue) {
    for (int off = 0; off < count; off += cap) {
      const int cnt = std::min(cap, count - off);
      em::ColParams cp{};
      cp.w.n_tr = cnt; cp.w.x = ctx->long_x.p;
      cp.mode = 1; cp.tabs_build = d_e + off; cp.tr_per_cta = 1; cp.twl = ctx->d_twl[lp1];
      { int rc = launch_long_colfwd(ctx, ctx->stream, lp1, cp, cnt); if (rc) return rc; }
      em::RowParams rp{ctx->long_x.p, lp1, 1, nullptr, d_e + off, ctx->d_tw[em::kLongLog2P2]};
      { int rc = launch_long_row(ctx, ctx->stream, rp, cnt); if (rc) return rc; }
    }
  }
  return PROTON_EM_OK;
}


// lambda_k exactly as the reference computes it (form_analytical_solution, bin/EMtool_analytical@0xc4e8):
// cos((double)k*pi/(double)N) from the HOST libm — the same libm the reference binary links — doubled,
// minus 2.0.  A device cos would differ in the last bit often enough to break 1e-9 parity where
// exp(lambda a t) - 1 cancels.
void host_lambda(int N, double* out) {
  const double pi = 3.141592653589793115997963468544185161590576171875;
  for (int k = 0; k < N; ++k) {                   // (host code is built with -ffp-contract=off: three roundings, as written)
    const double c = std::cos((double)k * pi / (double)N);
    const double c2 = c + c;
    out[k] = c2 - 2.0;
  }
}

// Make sure every distinct N of `lens` has its tables in the context cache.
int ensure_tables(proton_em_ctx* ctx, const std::vector<int>& distinct) {
  std::vector<int> fresh;
  for (int N : distinct)
    if (ctx->tabs.find(N) == ctx->tabs.end()) fresh.push_back(N);
  if (fresh.empty()) return PROTON_EM_OK;
  Trace tr("ensure_tables");
  size_t total = 0;
  std::vector<size_t> off(fresh.size());
  for (size_t i = 0; i < fresh.size(); ++i) { off[i] = total; total += (size_t)fresh[i]; }
  {
    // everything the new lengths need, in one allocation
    size_t need = ((total * sizeof(double) + 255) & ~size_t(255)) + ((fresh.size() * sizeof(em::NTab) + 255) & ~size_t(255));
    for (int N : fresh) {
      const ClassOf c = classify(N);
      const size_t nb = ((size_t)N * sizeof(em::cplx) + 255) & ~size_t(255);
      if (c.cls != CLS_SMALL) need += nb;
      if (c.cls == CLS_BLUE || c.cls == CLS_LBLUE || c.cls == CLS_MIX) {
        const size_t plen = c.cls == CLS_LBLUE ? (size_t)c.log2p * em::kLongP2
                          : c.cls == CLS_MIX ? ((size_t)mix_odd(c.log2p) << mix_lr(c.log2p)) : ((size_t)1 << c.log2p);
        need += nb + ((plen * sizeof(em::cplx) + 255) & ~size_t(255));
      }
    }
    CU_TRY(ctx->arena.reserve(need));
  }
  void* d_lam = nullptr;
  CU_TRY(ctx->arena.alloc(&d_lam, total * sizeof(double)));
  // group the new entries by (class, log2p) for the table kernels
  std::vector<em::NTab> entries(fresh.size());
  for (size_t i = 0; i < fresh.size(); ++i) {
    const int N = fresh[i];
    const ClassOf c = classify(N);
    em::NTab t{};
    t.lam = static_cast<const double*>(d_lam) + off[i];
    t.n = N;
    t.log2p = (c.cls == CLS_POW2 || c.cls == CLS_BLUE) ? c.log2p : 0;
    t.bluestein = c.cls == CLS_BLUE;
    t.p1 = 0;
    if (c.cls == CLS_LPOW2 || c.cls == CLS_LBLUE) { t.p1 = c.log2p; t.bluestein = c.cls == CLS_LBLUE; }
    t.mix = c.cls == CLS_MIX ? c.log2p : 0;
    if (c.cls == CLS_MIX) t.bluestein = 1;
    if (c.cls != CLS_SMALL) {
      void* p = nullptr;
      CU_TRY(ctx->arena.alloc(&p, (size_t)N * sizeof(em::cplx)));
      t.pre = static_cast<const em::cplx*>(p);
    }
    if (c.cls == CLS_BLUE || c.cls == CLS_LBLUE || c.cls == CLS_MIX) {
      void* p = nullptr;
      CU_TRY(ctx->arena.alloc(&p, (size_t)N * sizeof(em::cplx)));
      t.post = static_cast<const em::cplx*>(p);
      const size_t plen = c.cls == CLS_LBLUE ? (size_t)c.log2p * em::kLongP2
                        : c.cls == CLS_MIX ? ((size_t)mix_odd(c.log2p) << mix_lr(c.log2p)) : ((size_t)1 << c.log2p);
      CU_TRY(ctx->arena.alloc(&p, plen * sizeof(em::cplx)));
      t.hhat = static_cast<const em::cplx*>(p);
    }
    entries[i] = t;
  }
  tr.lap("arena + entries");
  // One upload of all new entries (grouped by kernel), then one table launch per group, all on the context stream
  // and without host synchronisation in between.
  std::vector<em::NTab> sorted;
  sorted.reserve(entries.size());
  struct Group { int blue, key; size_t off, cnt; bool is_long; bool is_mix = false; };
  std::vector<Group> groups;
  for (int code = 1; code < 192; ++code) {
    const size_t off = sorted.size();
    for (const em::NTab& t : entries)
      if (t.mix == code) sorted.push_back(t);
    if (sorted.size() > off) { Group g{1, code, off, sorted.size() - off, false}; g.is_mix = true; groups.push_back(g); }
  }
  for (int blue = 0; blue < 2; ++blue)
    for (int p = kMinLog2P; p <= kMaxLog2P; ++p) {
      const size_t off = sorted.size();
      for (const em::NTab& t : entries)
        if (t.log2p == p && t.bluestein == blue && t.mix == 0) sorted.push_back(t);
      if (sorted.size() > off) groups.push_back({blue, p, off, sorted.size() - off, false});
    }
  for (int blue = 0; blue < 2; ++blue)
    for (int p1 = 3; p1 <= em::kLongMaxP1; ++p1) {
      const size_t off = sorted.size();
      for (const em::NTab& t : entries)
        if (t.p1 == p1 && t.bluestein == blue) sorted.push_back(t);
      if (sorted.size() > off) groups.push_back({blue, p1, off, sorted.size() - off, true});
    }
  if (!sorted.empty()) {
    void* d_sorted_v = nullptr;
    CU_TRY(ctx->arena.alloc(&d_sorted_v, sorted.size() * sizeof(em::NTab)));
    em::NTab* d_sorted = static_cast<em::NTab*>(d_sorted_v);
    CU_TRY(cudaMemcpyAsync(d_sorted, sorted.data(), sorted.size() * sizeof(em::NTab), cudaMemcpyHostToDevice, ctx->stream));
    for (const Group& g : groups) {
      int rc;
      if (g.is_mix) rc = launch_tables_mix(ctx, g.key, d_sorted + g.off, (int)g.cnt);
      else if (g.is_long) rc = long_build_tables(ctx, d_sorted + g.off, (int)g.cnt, g.key, g.blue != 0);
      else rc = g.blue ? launch_tables<true>(ctx, g.key, d_sorted + g.off, (int)g.cnt)
                       : launch_tables<false>(ctx, g.key, d_sorted + g.off, (int)g.cnt);
      if (rc) { cudaStreamSynchronize(ctx->stream); return rc; }
    }
  }
  tr.lap("table kernels queued");
  // lambda_k on the host cores while the GPU builds the chirp tables (which do not depend on it)
  std::vector<double> lam(total);
  {
    unsigned nth = std::thread::hardware_concurrency();
    if (nth == 0) nth = 1;
    if (total < (size_t)1 << 16) nth = 1;
    nth = std::min<unsigned>(nth, 64);
    std::vector<std::thread> pool;
    std::vector<size_t> order(fresh.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = i;
    auto work = [&](unsigned w) {
      for (size_t i = w; i < fresh.size(); i += nth) host_lambda(fresh[i], lam.data() + off[i]);
    };
    for (unsigned w = 1; w < nth; ++w) pool.emplace_back(work, w);
    work(0);
    for (auto& t : pool) t.join();
  }
  tr.lap("host lambda");
  CU_TRY(cudaMemcpyAsync(d_lam, lam.data(), total * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  CU_TRY(cudaStreamSynchronize(ctx->stream));   // lam upload must finish before `lam` goes away
  tr.lap("sync");
  for (const em::NTab& t : entries) ctx->tabs[t.n] = t;
  return PROTON_EM_OK;
}


template <typename T>
cudaError_t grow(DevBuf<T>& buf, size_t n) { return buf.n >= n ? cudaSuccess : buf.alloc(n); }

void destroy_plan(proton_em_plan* p) {
  if (!p) return;
  p->own_num_nodes.free(); p->own_a.free(); p->own_trip_off.free();
  p->d_node_off.free(); p->d_blk_off.free(); p->d_r_idx.free(); p->d_r_cnt.free(); p->d_line_tab.free();
  p->d_r_val.free(); p->d_W.free(); p->d_tabs.free(); p->d_class_lines.free(); p->d_sigma0.free();
  for (auto* ll : p->long_lists) { ll->recs.free(); delete ll; }
  delete p;
}

}  // namespace

extern "C" {

const char* proton_em_last_error(void) { return g_last_error.c_str(); }
const char* proton_em_version(void) { return "proton_em 0.1 (sm_100a)"; }

int proton_em_create(proton_em_ctx** out, int device) try {
  if (!out) return fail(PROTON_EM_EINVAL, "proton_em_create: out is NULL");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(PROTON_EM_ECUDA, "no CUDA device available (%s); this library has no CPU path",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  if (device < 0 || device >= ndev) return fail(PROTON_EM_EINVAL, "device %d out of range [0, %d)", device, ndev);
  CU_TRY(cudaSetDevice(device));
  proton_em_ctx* ctx = new proton_em_ctx();
  ctx->device = device;
  cudaDeviceProp prop{};
  CU_TRY(cudaGetDeviceProperties(&prop, device));
  ctx->sm_count = prop.multiProcessorCount;
  CU_TRY(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  {
    const char* e = std::getenv("PROTON_EM_ASYNC_ALLOC");
    if (!(e && std::atoi(e) == 0)) {
      cudaMemPoolProps props = {};
      props.allocType = cudaMemAllocationTypePinned;
      props.location.type = cudaMemLocationTypeDevice;
      props.location.id = device;
      if (cudaMemPoolCreate(&ctx->pool, &props) == cudaSuccess) {
        unsigned long long keep = ~0ULL;                    // freed blocks stay in the pool until the context goes
        cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep);
      } else {
        ctx->pool = nullptr;                                // no pool support: plain cudaMalloc / cudaFree
        cudaGetLastError();
      }
    }
    tl_pool = ctx->pool;
    tl_stream = ctx->stream;
  }
  for (int i = 0; i < proton_em_ctx::kLanes; ++i) {
    CU_TRY(cudaStreamCreateWithFlags(&ctx->lane[i], cudaStreamNonBlocking));
    CU_TRY(cudaEventCreateWithFlags(&ctx->ev_join[i], cudaEventDisableTiming));
  }
  CU_TRY(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
  if (const char* e = std::getenv("PROTON_EM_LANES")) ctx->lanes = std::max(1, std::min((int)proton_em_ctx::kLanes, std::atoi(e)));
  CU_TRY(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  for (int i = 0; i < 2; ++i) {
    CU_TRY(cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&ctx->ev_copied[i], cudaEventDisableTiming));
  }
  CU_TRY(cudaMalloc(reinterpret_cast<void**>(&ctx->d_exp), sizeof(kExpTable)));
  CU_TRY(cudaMemcpy(ctx->d_exp, kExpTable, sizeof(kExpTable), cudaMemcpyHostToDevice));
  for (int p = kMinLog2P; p <= kMaxLog2P; ++p) {
    // base twiddles per stage (em_fft.cuh: stage_tw_offset): e^{+2 pi i j / (R S)}, j < S
    std::vector<em::cplx> tw((size_t)em::stage_tw_total(p));
    const long double two_pi = 6.283185307179586476925286766559005768L;
    int log2s = p;
    for (int i = 0; i < em::plan_n(p); ++i) {
      const int R = em::plan_radix(p, i);
      log2s -= em::lg2(R);
      const int S = 1 << log2s;
      if (S == 1) continue;
      const int off = em::stage_tw_offset(p, i);
      for (int j = 0; j < S; ++j) {
        const long double ang = two_pi * (long double)j / ((long double)R * (long double)S);
        tw[(size_t)off + j] = em::cplx{(double)cosl(ang), (double)sinl(ang)};
      }
    }
    {
      const int P = 1 << p;
      std::vector<int> pos(P);
      for (int k = 0; k < P; ++k) {
        int q = 0, sh = p, kk = k;
        for (int i = 0; i < em::plan_n(p); ++i) {
          const int lr = em::lg2(em::plan_radix(p, i));
          sh -= lr;
          q |= (kk & ((1 << lr) - 1)) << sh;
          kk >>= lr;
        }
        pos[k] = em::pad_idx_rt(p, q);
      }
      CU_TRY(cudaMalloc(reinterpret_cast<void**>(&ctx->d_pos[p]), P * sizeof(int)));
      CU_TRY(cudaMemcpy(ctx->d_pos[p], pos.data(), P * sizeof(int), cudaMemcpyHostToDevice));
    }
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&ctx->d_tw[p]), tw.size() * sizeof(em::cplx)));
    CU_TRY(cudaMemcpy(ctx->d_tw[p], tw.data(), tw.size() * sizeof(em::cplx), cudaMemcpyHostToDevice));
  }
  int rc = init_kernel_attrs();
  if (rc == PROTON_EM_OK) rc = init_long_attrs();
  if (rc == PROTON_EM_OK) rc = init_mix_attrs();
  if (rc == PROTON_EM_OK) {
    const int n = em::kDenseMaxN;
    const cudaError_t e = cudaFuncSetAttribute(em::em_eval_dense_kernel<12>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)((4 * n + (n / 8) * (n / 4) * 32) * sizeof(double)));
    if (e != cudaSuccess) rc = fail(PROTON_EM_ECUDA, "cudaFuncSetAttribute(dense): %s", cudaGetErrorString(e));
  }
  if (rc) { proton_em_destroy(ctx); return rc; }
  *out = ctx;
  return PROTON_EM_OK;
} EM_ABI_CATCH("proton_em_create")

int proton_em_destroy(proton_em_ctx* ctx) {
  if (!ctx) return PROTON_EM_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  tl_pool = nullptr;                                        // the device is idle: plain cudaFree for what is left
  tl_stream = nullptr;
  for (int i = 0; i < proton_em_ctx::kLanes; ++i) {
    if (ctx->lane[i]) cudaStreamDestroy(ctx->lane[i]);
    if (ctx->ev_join[i]) cudaEventDestroy(ctx->ev_join[i]);
  }
  if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
  for (const proton_em_ctx::ProfRec& r : ctx->prof_recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  for (cudaEvent_t e : ctx->prof_pool) cudaEventDestroy(e);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  for (int i = 0; i < 2; ++i) {
    if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
    if (ctx->ev_copied[i]) cudaEventDestroy(ctx->ev_copied[i]);
    ctx->io_stress[i].free();
  }
  ctx->io_mx.free(); ctx->io_ts.free(); ctx->io_amx.free(); ctx->io_flag.free(); ctx->io_tmp.free();
  if (ctx->stream) { cudaStreamSynchronize(ctx->stream); cudaStreamDestroy(ctx->stream); }
  if (ctx->d_exp) cudaFree(ctx->d_exp);
  for (auto* p : ctx->d_tw) if (p) cudaFree(p);
  for (auto* p : ctx->d_pos) if (p) cudaFree(p);
  for (auto* p : ctx->d_twl) if (p) cudaFree(p);
  for (auto* p : ctx->d_twm) if (p) cudaFree(p);
  ctx->long_x.free(); ctx->long_y.free(); ctx->long_part_v.free(); ctx->long_part_i.free();
  ctx->arena.release();
  if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
  delete ctx;
  return PROTON_EM_OK;
}

int proton_em_synchronize(proton_em_ctx* ctx) {
  if (!ctx) return fail(PROTON_EM_EINVAL, "ctx is NULL");
  CU_TRY(cudaStreamSynchronize(ctx->stream));
  return PROTON_EM_OK;
}

void* proton_em_stream(proton_em_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int64_t proton_em_launch_count(const proton_em_ctx* ctx) { return ctx ? ctx->launches : 0; }

int proton_em_profile(proton_em_ctx* ctx, int enable) {
  if (!ctx) return fail(PROTON_EM_EINVAL, "proton_em_profile: ctx is NULL");
  { int rc = set_device(ctx); if (rc) return rc; }
  CU_TRY(cudaDeviceSynchronize());
  // fold the finished records into the totals and recycle their events
  for (const proton_em_ctx::ProfRec& r : ctx->prof_recs) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) { ctx->prof_ms[r.kind] += ms; ctx->prof_n[r.kind]++; }
    ctx->prof_pool.push_back(r.a);
    ctx->prof_pool.push_back(r.b);
  }
  ctx->prof_recs.clear();
  if (enable && !ctx->prof_on)
    for (int k = 0; k < PK_COUNT; ++k) { ctx->prof_ms[k] = 0.0; ctx->prof_n[k] = 0; }
  ctx->prof_on = enable != 0;
  return PROTON_EM_OK;
}

int proton_em_profile_read(const proton_em_ctx* ctx, int kind, const char** name, double* total_ms, int64_t* launches) {
  if (!ctx) return fail(PROTON_EM_EINVAL, "proton_em_profile_read: ctx is NULL");
  if (kind < 0 || kind >= PK_COUNT) return fail(PROTON_EM_EINVAL, "kernel kind %d out of range [0, %d)", kind, (int)PK_COUNT);
  if (name) *name = kProfNames[kind];
  if (total_ms) *total_ms = ctx->prof_ms[kind];
  if (launches) *launches = ctx->prof_n[kind];
  return PROTON_EM_OK;
}

int proton_em_plan_destroy(proton_em_plan* plan) {
  if (plan && plan->ctx) { set_device(plan->ctx); cudaStreamSynchronize(plan->ctx->stream); }
  destroy_plan(plan);
  return PROTON_EM_OK;
}
int64_t proton_em_plan_total_nodes(const proton_em_plan* plan) { return plan ? plan->total_nodes : 0; }
int64_t proton_em_plan_num_lines(const proton_em_plan* plan) { return plan ? plan->L : 0; }
int32_t proton_em_plan_line_nodes(const proton_em_plan* plan, int64_t line) {
  return (plan && line >= 0 && line < plan->L) ? plan->num_nodes[(size_t)line] : 0;
}
int proton_em_plan_set_line_names(proton_em_plan* plan, const char* const* names, int64_t n) try {
  if (!plan) return fail(PROTON_EM_EINVAL, "proton_em_plan_set_line_names: null plan");
  if (n != plan->L || (n > 0 && !names)) return fail(PROTON_EM_EINVAL, "proton_em_plan_set_line_names: the plan has %lld lines, %lld names given", (long long)plan->L, (long long)n);
  std::vector<std::string> v((size_t)n);
  for (int64_t i = 0; i < n; ++i) v[(size_t)i] = names[i] ? names[i] : "";
  plan->names.swap(v);
  return PROTON_EM_OK;
} EM_ABI_CATCH("proton_em_plan_set_line_names")
const char* proton_em_plan_line_name(const proton_em_plan* plan, int64_t line) {
  if (!plan || line < 0 || (size_t)line >= plan->names.size()) return nullptr;
  return plan->names[(size_t)line].c_str();
}
int64_t proton_em_plan_scratch_bytes(const proton_em_plan* p) {
  if (!p) return 0;
  return (int64_t)(p->d_node_off.bytes() + p->d_r_idx.bytes() + p->d_r_cnt.bytes() + p->d_line_tab.bytes() +
                   p->d_r_val.bytes() + p->d_W.bytes() + p->d_tabs.bytes() + p->d_class_lines.bytes());
}

static int pack_impl(proton_em_ctx* ctx, const proton_em_batch* b, int location, int64_t chunk_nodes, proton_em_plan** out,
                     const double* sigma0 = nullptr);

int proton_em_pack(proton_em_ctx* ctx, const proton_em_batch* b, int location, proton_em_plan** out) try {
  return pack_impl(ctx, b, location, 0, out);
} EM_ABI_CATCH("proton_em_pack")

int proton_em_pack_sigma0(proton_em_ctx* ctx, const proton_em_batch* b, const double* sigma0, int location,
                          proton_em_plan** out) try {
  if (!sigma0 && b && b->num_lines > 0) return fail(PROTON_EM_EINVAL, "proton_em_pack_sigma0: sigma0 is NULL");
  return pack_impl(ctx, b, location, 0, out, sigma0);
} EM_ABI_CATCH("proton_em_pack_sigma0")

static int pack_impl(proton_em_ctx* ctx, const proton_em_batch* b, int location, int64_t chunk_nodes, proton_em_plan** out,
                     const double* sigma0) {
  if (!ctx || !b || !out) return fail(PROTON_EM_EINVAL, "proton_em_pack: NULL argument");
  *out = nullptr;
  if (location != PROTON_EM_HOST && location != PROTON_EM_DEVICE) return fail(PROTON_EM_EINVAL, "bad location %d", location);
  const int64_t L = b->num_lines;
  if (L < 0) return fail(PROTON_EM_EINVAL, "num_lines < 0");
  if (L > 0 && (!b->num_nodes || !b->a_coeff || !b->trip_off || !b->u_off))
    return fail(PROTON_EM_EINVAL, "proton_em_pack: NULL batch array");
  { int rc = set_device(ctx); if (rc) return rc; }
  Trace tr("pack");
  proton_em_plan* p = new proton_em_plan();
  p->ctx = ctx;
  p->L = L;
  struct Guard { proton_em_plan* p; ~Guard() { if (p) destroy_plan(p); } } guard{p};
  if (L == 0) { guard.p = nullptr; *out = p; return PROTON_EM_OK; }

  // ---- host view of the per-line sizes -----------------------------------------------------
  p->num_nodes.resize(L);
  std::vector<long long> trip_off(L + 1), u_off(L + 1);
  p->a_host.resize(L);
  if (location == PROTON_EM_HOST) {
    std::memcpy(p->a_host.data(), b->a_coeff, L * sizeof(double));
    std::memcpy(p->num_nodes.data(), b->num_nodes, L * sizeof(int));
    for (int64_t i = 0; i <= L; ++i) { trip_off[i] = b->trip_off[i]; u_off[i] = b->u_off[i]; }
  } else {
    CU_TRY(cudaMemcpy(p->num_nodes.data(), b->num_nodes, L * sizeof(int), cudaMemcpyDeviceToHost));
    CU_TRY(cudaMemcpy(p->a_host.data(), b->a_coeff, L * sizeof(double), cudaMemcpyDeviceToHost));
    CU_TRY(cudaMemcpy(trip_off.data(), b->trip_off, (L + 1) * sizeof(long long), cudaMemcpyDeviceToHost));
    CU_TRY(cudaMemcpy(u_off.data(), b->u_off, (L + 1) * sizeof(long long), cudaMemcpyDeviceToHost));
  }
  for (int64_t i = 0; i < L; ++i) {
    if (p->num_nodes[i] <= 0) return fail(PROTON_EM_EINVAL, "line %lld: num_nodes = %d", (long long)i, p->num_nodes[i]);
    if (trip_off[i + 1] < trip_off[i] || u_off[i + 1] < u_off[i])
      return fail(PROTON_EM_EINVAL, "line %lld: offsets are not non-decreasing", (long long)i);
    if (trip_off[i + 1] - trip_off[i] > (1 << 20))
      return fail(PROTON_EM_EINVAL, "line %lld: more than 2^20 triplets", (long long)i);
  }
  const long long tbase = trip_off[0], ubase = u_off[0];
  p->ntrip = trip_off[L] - tbase;
  p->nu = u_off[L] - ubase;
  if (p->ntrip > 0 && (!b->trip_row || !b->trip_col || !b->trip_val)) return fail(PROTON_EM_EINVAL, "NULL triplet array");
  if (p->nu > 0 && !b->curden) return fail(PROTON_EM_EINVAL, "NULL curden array");
  p->node_off.resize(L + 1);
  p->node_off[0] = 0;
  int maxN = 0;
  for (int64_t i = 0; i < L; ++i) { p->node_off[i + 1] = p->node_off[i] + p->num_nodes[i]; maxN = std::max(maxN, p->num_nodes[i]); }
  p->total_nodes = p->node_off[L];
  p->max_nodes = maxN;

  // ---- classes and per-length tables -------------------------------------------------------
  std::vector<int> distinct(p->num_nodes);
  std::sort(distinct.begin(), distinct.end());
  distinct.erase(std::unique(distinct.begin(), distinct.end()), distinct.end());
  for (int N : distinct)
    if (classify(N).cls == CLS_NONE)
      return fail(PROTON_EM_EINVAL, "a line has %d nodes; this build handles N <= %d (any N) and powers of two up to %d",
                  N, kMaxLongBlueN, kMaxLongPow2N);
  tr.lap("sizes");
  { int rc = ensure_tables(ctx, distinct); if (rc) return rc; }
  tr.lap("ensure_tables");
  std::unordered_map<int, int> tab_index;
  std::vector<em::NTab> tabs(distinct.size());
  for (size_t i = 0; i < distinct.size(); ++i) { tabs[i] = ctx->tabs[distinct[i]]; tab_index[distinct[i]] = (int)i; }
  std::vector<int> line_tab(L);
  // chunks of contiguous lines, and per chunk the class lists keyed by (class, log2 P)
  {
    int64_t l0 = 0;
    while (l0 < L) {
      int64_t l1 = l0, nodes = 0;
      while (l1 < L && (l1 == l0 || chunk_nodes <= 0 || nodes + p->num_nodes[l1] <= chunk_nodes)) { nodes += p->num_nodes[l1]; ++l1; }
      proton_em_plan::Chunk ch;
      ch.l0 = l0; ch.l1 = l1;
      p->chunks.push_back(ch);
      l0 = l1;
    }
  }
  std::vector<int> class_lines;
  class_lines.reserve(L);
  for (auto& ch : p->chunks) {
    std::map<int, std::vector<int>> lists;
    for (int64_t i = ch.l0; i < ch.l1; ++i) {
      const int N = p->num_nodes[i];
      line_tab[i] = tab_index[N];
      const ClassOf c = classify(N);
      lists[(int)c.cls * 256 + c.log2p].push_back((int)i);
    }
    for (auto& kv : lists) {
      proton_em_plan::ClassList cl;
      cl.cls = (LineClass)(kv.first / 256);
      cl.log2p = kv.first % 256;
      // longest lines first inside a class (they are the most expensive items)
      std::stable_sort(kv.second.begin(), kv.second.end(), [&](int x, int y) { return p->num_nodes[x] > p->num_nodes[y]; });
      cl.offset = (int)class_lines.size();
      cl.count = (int)kv.second.size();
      class_lines.insert(class_lines.end(), kv.second.begin(), kv.second.end());
      ch.classes.push_back(cl);
    }
  }

  tr.lap("class lists");
  // ---- device copies -----------------------------------------------------------------------
  cudaStream_t st = ctx->stream;
  DevBuf<int> tmp_row, tmp_col, tmp_perm;
  DevBuf<double> tmp_val, tmp_u, tmp_prod;
  DevBuf<long long> tmp_u_off;
  DevBuf<int> d_err;
  struct TmpGuard { DevBuf<int>&a, &b, &c, &e; DevBuf<double>&v, &u, &pr; DevBuf<long long>& uo;
    ~TmpGuard() { a.free(); b.free(); c.free(); e.free(); v.free(); u.free(); pr.free(); uo.free(); } }
      tguard{tmp_row, tmp_col, tmp_perm, d_err, tmp_val, tmp_u, tmp_prod, tmp_u_off};
  const int* d_row; const int* d_col; const double* d_val; const double* d_u; const long long* d_u_off;
  std::vector<long long> trip_rel, u_rel;
  if (location == PROTON_EM_HOST) {
    p->owns_inputs = true;
    trip_rel.resize(L + 1); u_rel.resize(L + 1);
    for (int64_t i = 0; i <= L; ++i) { trip_rel[i] = trip_off[i] - tbase; u_rel[i] = u_off[i] - ubase; }
    CU_TRY(p->own_num_nodes.alloc(L)); CU_TRY(p->own_a.alloc(L)); CU_TRY(p->own_trip_off.alloc(L + 1));
    CU_TRY(tmp_u_off.alloc(L + 1));
    CU_TRY(tmp_row.alloc(p->ntrip)); CU_TRY(tmp_col.alloc(p->ntrip)); CU_TRY(tmp_val.alloc(p->ntrip)); CU_TRY(tmp_u.alloc(p->nu));
    CU_TRY(cudaMemcpyAsync(p->own_num_nodes.p, b->num_nodes, L * sizeof(int), cudaMemcpyHostToDevice, st));
    CU_TRY(cudaMemcpyAsync(p->own_a.p, b->a_coeff, L * sizeof(double), cudaMemcpyHostToDevice, st));
    CU_TRY(cudaMemcpyAsync(p->own_trip_off.p, trip_rel.data(), (L + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
    CU_TRY(cudaMemcpyAsync(tmp_u_off.p, u_rel.data(), (L + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
    if (p->ntrip) {
      CU_TRY(cudaMemcpyAsync(tmp_row.p, b->trip_row + tbase, p->ntrip * sizeof(int), cudaMemcpyHostToDevice, st));
      CU_TRY(cudaMemcpyAsync(tmp_col.p, b->trip_col + tbase, p->ntrip * sizeof(int), cudaMemcpyHostToDevice, st));
      CU_TRY(cudaMemcpyAsync(tmp_val.p, b->trip_val + tbase, p->ntrip * sizeof(double), cudaMemcpyHostToDevice, st));
    }
    if (p->nu) CU_TRY(cudaMemcpyAsync(tmp_u.p, b->curden + ubase, p->nu * sizeof(double), cudaMemcpyHostToDevice, st));
    p->d_num_nodes = p->own_num_nodes.p; p->d_a = p->own_a.p; p->d_trip_off = p->own_trip_off.p;
    d_row = tmp_row.p; d_col = tmp_col.p; d_val = tmp_val.p; d_u = tmp_u.p; d_u_off = tmp_u_off.p;
  } else {
    // device-resident batch: num_nodes / a_coeff / trip_off stay referenced by the plan (caller keeps them alive)
    p->d_num_nodes = b->num_nodes; p->d_a = b->a_coeff;
    if (tbase == 0) p->d_trip_off = reinterpret_cast<const long long*>(b->trip_off);
    else {
      trip_rel.resize(L + 1);
      for (int64_t i = 0; i <= L; ++i) trip_rel[i] = trip_off[i] - tbase;
      CU_TRY(p->own_trip_off.alloc(L + 1));
      CU_TRY(cudaMemcpyAsync(p->own_trip_off.p, trip_rel.data(), (L + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
      p->d_trip_off = p->own_trip_off.p;
    }
    d_row = b->trip_row + tbase; d_col = b->trip_col + tbase; d_val = b->trip_val + tbase;
    if (ubase == 0) { d_u_off = reinterpret_cast<const long long*>(b->u_off); d_u = b->curden; }
    else {
      u_rel.resize(L + 1);
      for (int64_t i = 0; i <= L; ++i) u_rel[i] = u_off[i] - ubase;
      CU_TRY(tmp_u_off.alloc(L + 1));
      CU_TRY(cudaMemcpyAsync(tmp_u_off.p, u_rel.data(), (L + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
      d_u_off = tmp_u_off.p; d_u = b->curden + ubase;
    }
  }
  CU_TRY(p->d_node_off.alloc(L + 1)); CU_TRY(p->d_line_tab.alloc(L)); CU_TRY(p->d_tabs.alloc(tabs.size()));
  CU_TRY(p->d_class_lines.alloc(class_lines.size()));
  CU_TRY(p->d_r_idx.alloc(std::max<int64_t>(p->ntrip, 1))); CU_TRY(p->d_r_val.alloc(std::max<int64_t>(p->ntrip, 1)));
  CU_TRY(p->d_r_cnt.alloc(L)); CU_TRY(p->d_W.alloc(p->total_nodes));
  CU_TRY(tmp_perm.alloc(std::max<int64_t>(p->ntrip, 1))); CU_TRY(tmp_prod.alloc(std::max<int64_t>(p->ntrip, 1)));
  CU_TRY(d_err.alloc(1));
  CU_TRY(cudaMemcpyAsync(p->d_node_off.p, p->node_off.data(), (L + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(p->d_line_tab.p, line_tab.data(), L * sizeof(int), cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(p->d_tabs.p, tabs.data(), tabs.size() * sizeof(em::NTab), cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(p->d_class_lines.p, class_lines.data(), class_lines.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  p->h_class_lines = class_lines;
  CU_TRY(cudaMemsetAsync(d_err.p, 0, sizeof(int), st));

  tr.lap("allocs + uploads queued");
  // ---- (a) packing kernel, (b) forward DCT kernel ------------------------------------------
  em::PackParams pp{L, p->d_num_nodes, p->d_trip_off, d_row, d_col, d_val, d_u_off, d_u,
                    tmp_perm.p, tmp_prod.p, p->d_r_idx.p, p->d_r_val.p, p->d_r_cnt.p, d_err.p};
  const int pack_grid = (int)std::min<int64_t>(L, (int64_t)ctx->sm_count * 16);
  em::em_pack_kernel<<<pack_grid, 128, 0, st>>>(pp);
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  std::vector<long long> blk_off(L + 1);
  blk_off[0] = 0;
  for (int64_t i = 0; i < L; ++i) blk_off[i + 1] = blk_off[i] + (p->num_nodes[i] + em::kFwdBlock - 1) / em::kFwdBlock;
  CU_TRY(p->d_blk_off.alloc(L + 1));
  CU_TRY(cudaMemcpyAsync(p->d_blk_off.p, blk_off.data(), (L + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
  em::ForwardParams fp{L, blk_off[L], p->d_num_nodes, p->d_a, p->d_node_off.p, p->d_blk_off.p, p->d_trip_off,
                       p->d_r_idx.p, p->d_r_val.p, p->d_r_cnt.p, p->d_line_tab.p, p->d_tabs.p, p->d_W.p};
  const int fwd_grid = (int)std::min<int64_t>((blk_off[L] + 127) / 128, (int64_t)ctx->sm_count * 64);
  em::em_forward_kernel<<<fwd_grid, 128, 0, st>>>(fp);
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  DevBuf<long long> d_slab_off;
  struct SlabGuard { DevBuf<long long>& b; ~SlabGuard() { b.free(); } } sguard{d_slab_off};
  std::vector<long long> slab_off;
  if (sigma0) {
    // initial stress: a plan-owned copy of the field, its dense forward DCT-II folded into the modal weights
    CU_TRY(p->d_sigma0.alloc((size_t)p->total_nodes + 2));
    CU_TRY(cudaMemsetAsync(p->d_sigma0.p + p->total_nodes, 0, 2 * sizeof(double), st));
    CU_TRY(cudaMemcpyAsync(p->d_sigma0.p, sigma0, (size_t)p->total_nodes * sizeof(double),
                           location == PROTON_EM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, st));
    const int slab = em::kS0Threads * em::kS0Modes;
    slab_off.resize(L + 1);
    slab_off[0] = 0;
    for (int64_t i = 0; i < L; ++i) slab_off[i + 1] = slab_off[i] + (p->num_nodes[i] + slab - 1) / slab;
    if (slab_off[L] > 0x7fffffffLL) return fail(PROTON_EM_EINVAL, "too many lines for the initial-stress transform");
    CU_TRY(d_slab_off.alloc(L + 1));
    CU_TRY(cudaMemcpyAsync(d_slab_off.p, slab_off.data(), (L + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
    em::Sigma0Params sp{p->d_num_nodes, p->d_node_off.p, d_slab_off.p, L, p->d_sigma0.p, p->d_W.p};
    em::em_forward_sigma0_kernel<<<(unsigned)slab_off[L], em::kS0Threads, 0, st>>>(sp);
    ctx->launches++;
    CU_TRY(cudaGetLastError());
  }
  int err = 0;
  CU_TRY(cudaMemcpyAsync(&err, d_err.p, sizeof(int), cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  tr.lap("kernels + sync");
  if (err != 0)
    return fail(PROTON_EM_EINVAL, "line %d: a B triplet has a row outside [0, num_nodes) or a column outside [0, #curden)", err - 1);
  guard.p = nullptr;
  *out = p;
  return PROTON_EM_OK;
}

// All long classes of a chunk: their transforms are listed once, cut into waves per class, and the waves are dealt
// round-robin to the lanes (streams) so that small waves of different classes overlap; one join, one reduction.
static int eval_long_all(proton_em_plan* p, const std::vector<const proton_em_plan::ClassList*>& longs,
                         const double* d_ts, int T, double* d_stress, double* d_mx, int* d_amx, unsigned char* d_flag,
                         double critical, long long* n_tr_out, int chunk_begin, int chunk_end,
                         const proton_em_plan::LongList** list_out) {
  proton_em_ctx* ctx = p->ctx;
  const int npairs = (T + 1) / 2;
  long long n_tr_total = 0;
  for (const auto* cl : longs) n_tr_total += (long long)cl->count * npairs;
  if (n_tr_total > 0x7fffffffLL) return fail(PROTON_EM_EINVAL, "too many long transforms");
  int max_cap_p1 = 0;
  for (const auto* cl : longs) {
    const int cap = (int)std::min<long long>(long_wave_capacity(cl->log2p), (long long)cl->count * npairs);
    int rc = ensure_long(ctx, cl->log2p, cap, n_tr_total);
    if (rc) return rc;
    max_cap_p1 = std::max(max_cap_p1, cap * cl->log2p);
  }
  proton_em_plan::LongList* ll = nullptr;
  for (auto* c : p->long_lists)
    if (c->T == T && c->c0 == chunk_begin && c->c1 == chunk_end) { ll = c; break; }
  if (!ll) {
    // one record per transform slot (line-major, timestamp pairs of a line consecutive); the host copy lives as long
    // as the list, so the upload needs no synchronisation
    if (p->long_lists.size() >= 8) {            // a plan evaluated at many different T: drop the oldest list
      CU_TRY(cudaStreamSynchronize(ctx->stream));
      proton_em_plan::LongList* old = p->long_lists.front();
      p->long_lists.erase(p->long_lists.begin());
      old->recs.free();
      delete old;
    }
    ll = new proton_em_plan::LongList();
    ll->T = T; ll->c0 = chunk_begin; ll->c1 = chunk_end; ll->n = n_tr_total;
    p->long_lists.push_back(ll);
    std::vector<em::SlotRec>& recs = ll->host;
    recs.resize((size_t)n_tr_total);
    size_t o = 0;
    for (const auto* cl : longs)
      for (int i = 0; i < cl->count; ++i) {
        const int line = p->h_class_lines[cl->offset + i];
        const int N = p->num_nodes[line];
        const em::NTab& t = ctx->tabs[N];
        for (int pr = 0; pr < npairs; ++pr, ++o) {
          em::SlotRec& r = recs[o];
          r.lam = t.lam; r.W = p->d_W.p + p->node_off[line]; r.pre = t.pre; r.post = t.post; r.hhat = t.hhat;
          r.stress_off = p->node_off[line] * (long long)T + 2LL * pr * N;
          r.a = p->a_host[line];
          r.N = N; r.line = line; r.pair = pr; r.reserved = 0;
        }
      }
    CU_TRY(ll->recs.alloc((size_t)n_tr_total));
    CU_TRY(cudaMemcpyAsync(ll->recs.p, recs.data(), n_tr_total * sizeof(em::SlotRec), cudaMemcpyHostToDevice, ctx->stream));
  }
  *list_out = ll;
  // fork: the lanes start after everything queued so far on the main stream (uploads above included)
  CU_TRY(cudaEventRecord(ctx->ev_fork, ctx->stream));
  for (int i = 0; i < ctx->lanes; ++i) CU_TRY(cudaStreamWaitEvent(ctx->lane[i], ctx->ev_fork, 0));
  const size_t lane_stride = ctx->long_x.n / (size_t)ctx->lanes;     // complex elements per lane
  int wave = 0;
  long long base = 0;
  for (const auto* cl : longs) {
    const bool blue = cl->cls == CLS_LBLUE;
    const int lp1 = cl->log2p;          // P1 (rows of 4096 points) for the long classes
    const long long n_tr = (long long)cl->count * npairs;
    const int cap = (int)std::min<long long>(long_wave_capacity(lp1), n_tr);
    int maxN = 0;
    for (int i = 0; i < cl->count; ++i) maxN = std::max(maxN, p->num_nodes[p->h_class_lines[cl->offset + i]]);
    for (long long s0 = 0; s0 < n_tr; s0 += cap, ++wave) {
      const int cnt = (int)std::min<long long>(cap, n_tr - s0);
      const int li = wave % ctx->lanes;
      cudaStream_t st = ctx->lane[li];
      em::cplx* xbuf = ctx->long_x.p + (size_t)li * lane_stride;
      em::LongWave w{};
      w.recs = ll->recs.p + base + s0; w.n_tr = cnt; w.T = T;
      w.ts = d_ts;
      w.x = xbuf; w.stress = d_stress; w.tr_base = (int)(base + s0);
      w.part_v = ctx->long_part_v.p; w.part_i = ctx->long_part_i.p;
      w.mx = d_mx; w.amx = d_amx; w.flag = d_flag;
      w.critical = critical;
      static const int col_div = [] { const char* e = std::getenv("PROTON_EM_COL_DIV"); const int v = e ? std::atoi(e) : 0; return v > 0 ? v : 32; }();
      const int tpc = std::max(1, cnt / col_div);
      const int groups = (cnt + tpc - 1) / tpc;
      em::ColParams cp{w, 0, nullptr, tpc, ctx->d_twl[lp1]};
      em::RowParams rp{xbuf, lp1, blue ? 0 : 2, w.recs, nullptr,
                       ctx->d_tw[em::kLongLog2P2]};
      int rc;
      if (blue) rc = launch_long_colfwd_paired(ctx, st, lp1, cp, groups);
      else rc = launch_long_scatter(ctx, st, lp1, w, maxN);
      if (rc) return rc;
      rc = launch_long_row(ctx, st, rp, cnt);
      if (rc) return rc;
      rc = launch_long_colinv(ctx, st, lp1, blue, cp, groups);
      if (rc) return rc;
    }
    base += n_tr;
  }
  *n_tr_out = n_tr_total;
  return PROTON_EM_OK;
}

// Join of the lanes and the one reduction of the column-block partials (queued after the other classes so that
// their kernels overlap the long waves).
static int eval_long_finish(proton_em_plan* p, const proton_em_plan::LongList* ll, long long n_tr_total, int T,
                            double* d_stress, double* d_mx, int* d_amx, unsigned char* d_flag, double critical) {
  proton_em_ctx* ctx = p->ctx;
  for (int i = 0; i < ctx->lanes; ++i) {
    CU_TRY(cudaEventRecord(ctx->ev_join[i], ctx->lane[i]));
    CU_TRY(cudaStreamWaitEvent(ctx->stream, ctx->ev_join[i], 0));
  }
  if (n_tr_total <= 0) return PROTON_EM_OK;
  em::LongWave w{};
  w.recs = ll->recs.p; w.n_tr = (int)n_tr_total; w.T = T;
  w.stress = d_stress;
  w.part_v = ctx->long_part_v.p; w.part_i = ctx->long_part_i.p; w.mx = d_mx; w.amx = d_amx; w.flag = d_flag;
  w.critical = critical;
  {
    ProfScope prof(ctx, PK_LONG_REDUCE, ctx->stream);
    em::em_long_reduce_kernel<<<(unsigned)((2 * n_tr_total + 127) / 128), 128, 0, ctx->stream>>>(w, kLongColBlocks);
  }
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

static int eval_device_core(proton_em_plan* p, const double* d_ts, int T, const proton_em_opts* opts,
                            double* d_stress, double* d_mx, int* d_amx, unsigned char* d_flag,
                            int chunk_begin, int chunk_end);

// Evaluation of the chunks [chunk_begin, chunk_end).  A plan packed with an initial stress (proton_em_pack_sigma0) has
// its forward DCT folded into the modal weights already; what is left is to add sigma0 to every timestamp's row and to
// take the reductions of the sums (em_add_sigma0_kernel; needs a field, so a reduce-only call gets a temporary one).
static int eval_device(proton_em_plan* p, const double* d_ts, int T, const proton_em_opts* opts,
                       double* d_stress, double* d_mx, int* d_amx, unsigned char* d_flag,
                       int chunk_begin = 0, int chunk_end = -1) {
  if (p->d_sigma0.n == 0) return eval_device_core(p, d_ts, T, opts, d_stress, d_mx, d_amx, d_flag, chunk_begin, chunk_end);
  proton_em_ctx* ctx = p->ctx;
  const int c_end = chunk_end < 0 ? (int)p->chunks.size() : chunk_end;
  if (chunk_begin >= c_end) return PROTON_EM_OK;
  const long long l0 = p->chunks[chunk_begin].l0, l1 = p->chunks[c_end - 1].l1;
  const long long n0 = p->node_off[l0], n1 = p->node_off[l1];
  const long long rows = (l1 - l0) * T;
  if (rows > 0x7fffffffLL) return fail(PROTON_EM_EINVAL, "too many (line, timestamp) rows in one evaluation");
  DevBuf<double> tmp;
  struct G { DevBuf<double>& b; ~G() { b.free(); } } g{tmp};
  if (!d_stress) {
    CU_TRY(tmp.alloc((size_t)(n1 - n0) * T));
    d_stress = tmp.p - n0 * (long long)T;                       // kernels index by global node offset
  }
  { int rc = eval_device_core(p, d_ts, T, opts, d_stress, d_mx, d_amx, d_flag, chunk_begin, chunk_end); if (rc) return rc; }
  em::AddSigma0Params ap{d_stress, p->d_sigma0.p, p->d_node_off.p, p->d_num_nodes, l0, T, d_mx, d_amx, d_flag,
                         opts ? opts->critical_stress : NAN};
  em::em_add_sigma0_kernel<<<(unsigned)rows, 128, 0, ctx->stream>>>(ap);
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  return PROTON_EM_OK;
}

static int eval_device_core(proton_em_plan* p, const double* d_ts, int T, const proton_em_opts* opts,
                            double* d_stress, double* d_mx, int* d_amx, unsigned char* d_flag,
                            int chunk_begin, int chunk_end) {
  proton_em_ctx* ctx = p->ctx;
  const double critical = opts ? opts->critical_stress : NAN;
  const int alg = opts ? opts->algorithm : PROTON_EM_ALG_AUTO;
  if (alg == PROTON_EM_ALG_DIRECT) {
    // only the lines of the chunks asked for: the host paths hand in a stress pointer that is valid for one chunk
    const int c_end = chunk_end < 0 ? (int)p->chunks.size() : chunk_end;
    if (chunk_begin >= c_end) return PROTON_EM_OK;
    const long long l0 = p->chunks[chunk_begin].l0, l1 = p->chunks[c_end - 1].l1;
    if ((l1 - l0) * T > 0x7fffffffLL) return fail(PROTON_EM_EINVAL, "direct algorithm: L*T too large");
    em::DirectParams dp{l1 - l0, l0, T, p->d_num_nodes, p->d_a, p->d_node_off.p, p->d_line_tab.p, p->d_tabs.p, p->d_W.p,
                        d_ts, ctx->d_exp, d_stress, d_mx, d_amx, d_flag, critical};
    em::em_eval_direct_kernel<<<(unsigned)((l1 - l0) * T), 256, 0, ctx->stream>>>(dp);
    ctx->launches++;
    CU_TRY(cudaGetLastError());
    return PROTON_EM_OK;
  }
  if (alg != PROTON_EM_ALG_AUTO) return fail(PROTON_EM_EINVAL, "unknown algorithm %d", alg);
  const long long target_items = (long long)ctx->sm_count * 12;
  if (chunk_end < 0) chunk_end = (int)p->chunks.size();
  std::vector<const proton_em_plan::ClassList*> longs;
  for (int ci = chunk_begin; ci < chunk_end; ++ci)
    for (const auto& cl : p->chunks[ci].classes)
      if (cl.cls == CLS_LPOW2 || cl.cls == CLS_LBLUE) longs.push_back(&cl);
  // Every class kernel goes to one of the lanes (streams forked off the main stream) so that the tails of the
  // per-class launches overlap; the long lines' waves use the same lanes.  One join at the end.
  long long n_long_tr = 0;
  const proton_em_plan::LongList* long_list = nullptr;
  int next_lane = 0;
  if (longs.empty()) {
    CU_TRY(cudaEventRecord(ctx->ev_fork, ctx->stream));
    for (int i = 0; i < ctx->lanes; ++i) CU_TRY(cudaStreamWaitEvent(ctx->lane[i], ctx->ev_fork, 0));
  } else {
    int rc = eval_long_all(p, longs, d_ts, T, d_stress, d_mx, d_amx, d_flag, critical, &n_long_tr, chunk_begin, chunk_end,
                           &long_list);
    if (rc) return rc;
  }
  for (int ci = chunk_begin; ci < chunk_end; ++ci)
  for (const auto& cl : p->chunks[ci].classes) {
    if (cl.cls == CLS_LPOW2 || cl.cls == CLS_LBLUE) continue;
    cudaStream_t cst = ctx->lane[next_lane++ % ctx->lanes];
    if (cl.cls == CLS_SMALL) {
      // dense contraction on the FP64 tensor cores (DMMA)
      {
        // the class list is sorted longest first: cut it where the number of 8-node tiles drops to 8 and to 4
        int begin = 0;
        const int limits[3] = {64, 32, 0};
        for (int part = 0; part < 3 && begin < cl.count; ++part) {
          int end = begin;
          while (end < cl.count && p->num_nodes[p->h_class_lines[cl.offset + end]] > limits[part]) ++end;
          if (end == begin) continue;
          const int cnt = end - begin;
          const int maxN = p->num_nodes[p->h_class_lines[cl.offset + begin]];
          int chunks = (int)std::min<long long>((T + 31) / 32, std::max<long long>(1, target_items / cnt));
          int tpc = (((T + chunks - 1) / chunks) + 7) / 8 * 8;
          chunks = (T + tpc - 1) / tpc;
          em::SmallParams sp{p->d_class_lines.p + cl.offset + begin, cnt, chunks, tpc, T, p->d_num_nodes, p->d_a,
                             p->d_node_off.p, p->d_line_tab.p, p->d_tabs.p, p->d_W.p, d_ts, ctx->d_exp,
                             d_stress, d_mx, d_amx, d_flag, critical, maxN};
          const size_t smem = (size_t)(4 * maxN + ((maxN + 7) / 8) * ((maxN + 3) / 4) * 32) * sizeof(double);
          const unsigned items = (unsigned)((long long)cnt * chunks);
          {
            ProfScope prof(ctx, PK_DENSE, cst);
            if (maxN > 64) em::em_eval_dense_kernel<12><<<items, 128, smem, cst>>>(sp);
            else if (maxN > 32) em::em_eval_dense_kernel<8><<<items, 128, smem, cst>>>(sp);
            else em::em_eval_dense_kernel<4><<<items, 128, smem, cst>>>(sp);
          }
          ctx->launches++;
          CU_TRY(cudaGetLastError());
          begin = end;
        }
        continue;
      }
    }
    if (cl.cls == CLS_MIX) {
      // one transform at a time per CTA; the timestamp pairs of a line are cut into chunks until the grid is large enough
      const int npairs = (T + 1) / 2;
      int chunks = (int)std::min<long long>(npairs, std::max<long long>(1, (target_items + cl.count - 1) / cl.count));
      const int ppc = (npairs + chunks - 1) / chunks;
      chunks = (npairs + ppc - 1) / ppc;
      em::EvalParams ep{p->d_class_lines.p + cl.offset, cl.count, chunks, ppc, T, p->d_num_nodes, p->d_a,
                        p->d_node_off.p, p->d_line_tab.p, p->d_tabs.p, p->d_W.p, d_ts, ctx->d_tw[mix_lr(cl.log2p)],
                        nullptr, d_stress, d_mx, d_amx, d_flag, critical};
      const long long grid = (long long)cl.count * chunks;
      if (grid > 0x7fffffffLL) return fail(PROTON_EM_EINVAL, "too many work items");
      int rc = launch_eval_mix(ctx, cst, cl.log2p, ep, (int)grid);
      if (rc) return rc;
      continue;
    }
    const bool blue = cl.cls == CLS_BLUE;
    const int G = fft_group(blue, cl.log2p);
    const int npairs = (T + 1) / 2;
    const int max_chunks = (npairs + G - 1) / G;
    int chunks = (int)std::min<long long>(max_chunks, std::max<long long>(1, (target_items + cl.count - 1) / cl.count));
    int ppc = (npairs + chunks - 1) / chunks;
    ppc = ((ppc + G - 1) / G) * G;
    chunks = (npairs + ppc - 1) / ppc;
    em::EvalParams ep{p->d_class_lines.p + cl.offset, cl.count, chunks, ppc, T, p->d_num_nodes, p->d_a,
                      p->d_node_off.p, p->d_line_tab.p, p->d_tabs.p, p->d_W.p, d_ts, ctx->d_tw[cl.log2p],
                      ctx->d_pos[cl.log2p], d_stress, d_mx, d_amx, d_flag, critical};
    const long long grid = (long long)cl.count * chunks;
    if (grid > 0x7fffffffLL) return fail(PROTON_EM_EINVAL, "too many work items");
    int rc = launch_eval_fft(ctx, cst, blue, cl.log2p, ep, (int)grid);
    if (rc) return rc;
  }
  return eval_long_finish(p, long_list, n_long_tr, T, d_stress, d_mx, d_amx, d_flag, critical);
}

int proton_em_eval(proton_em_plan* p, const double* ts, int32_t T, const proton_em_opts* opts,
                   const proton_em_result* res, int location) try {
  if (!p || !res) return fail(PROTON_EM_EINVAL, "proton_em_eval: NULL argument");
  if (T < 0) return fail(PROTON_EM_EINVAL, "T < 0");
  if (T == 0 || p->L == 0) return PROTON_EM_OK;
  if (!ts || !res->max_stress || !res->argmax || !res->flag)
    return fail(PROTON_EM_EINVAL, "proton_em_eval: ts, max_stress, argmax and flag are required");
  proton_em_ctx* ctx = p->ctx;
  { int rc = set_device(ctx); if (rc) return rc; }
  if (location == PROTON_EM_DEVICE)
    return eval_device(p, ts, T, opts, res->stress, res->max_stress, res->argmax, res->flag);
  if (location != PROTON_EM_HOST) return fail(PROTON_EM_EINVAL, "bad location %d", location);
  const size_t LT = (size_t)p->L * T;
  DevBuf<double> d_ts, d_mx, d_stress;
  DevBuf<int> d_amx;
  DevBuf<unsigned char> d_flag;
  struct G2 { DevBuf<double>&a, &b, &c; DevBuf<int>& d; DevBuf<unsigned char>& e;
    ~G2() { a.free(); b.free(); c.free(); d.free(); e.free(); } } g2{d_ts, d_mx, d_stress, d_amx, d_flag};
  CU_TRY(d_ts.alloc(T)); CU_TRY(d_mx.alloc(LT)); CU_TRY(d_amx.alloc(LT)); CU_TRY(d_flag.alloc(LT));
  if (res->stress) CU_TRY(d_stress.alloc((size_t)p->total_nodes * T));
  CU_TRY(cudaMemcpyAsync(d_ts.p, ts, T * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  int rc = eval_device(p, d_ts.p, T, opts, res->stress ? d_stress.p : nullptr, d_mx.p, d_amx.p, d_flag.p);
  if (rc) { cudaStreamSynchronize(ctx->stream); return rc; }
  if (res->stress)
    CU_TRY(cudaMemcpyAsync(res->stress, d_stress.p, d_stress.bytes(), cudaMemcpyDeviceToHost, ctx->stream));
  CU_TRY(cudaMemcpyAsync(res->max_stress, d_mx.p, d_mx.bytes(), cudaMemcpyDeviceToHost, ctx->stream));
  CU_TRY(cudaMemcpyAsync(res->argmax, d_amx.p, d_amx.bytes(), cudaMemcpyDeviceToHost, ctx->stream));
  CU_TRY(cudaMemcpyAsync(res->flag, d_flag.p, d_flag.bytes(), cudaMemcpyDeviceToHost, ctx->stream));
  CU_TRY(cudaStreamSynchronize(ctx->stream));
  return PROTON_EM_OK;
} EM_ABI_CATCH("proton_em_eval")

int proton_em_analyze_host(proton_em_ctx* ctx, const proton_em_batch* b, const double* ts, int32_t T,
                           const proton_em_opts* opts, const proton_em_result* res, int64_t max_device_bytes) try {
  if (!ctx || !b || !res) return fail(PROTON_EM_EINVAL, "proton_em_analyze_host: NULL argument");
  if (T <= 0 || b->num_lines <= 0) return PROTON_EM_OK;
  if (!ts || !res->max_stress || !res->argmax || !res->flag)
    return fail(PROTON_EM_EINVAL, "proton_em_analyze_host: ts, max_stress, argmax and flag are required");
  { int rc = set_device(ctx); if (rc) return rc; }
  // Two device staging buffers for the stress field; a chunk is copied back while the next one is computed.
  static const int64_t chunk_mb = [] { const char* e = std::getenv("PROTON_EM_CHUNK_MB"); const long long v = e ? std::atoll(e) : 0; return (int64_t)(v > 0 ? v : 256); }();
  int64_t chunk_bytes = chunk_mb << 20;
  if (max_device_bytes > 0) chunk_bytes = std::min<int64_t>(chunk_bytes, std::max<int64_t>(max_device_bytes / 2, 1 << 20));
  const int64_t chunk_nodes = res->stress ? std::max<int64_t>(1, chunk_bytes / (8 * (int64_t)T)) : 0;
  Trace tr("analyze_host");
  proton_em_plan* plan = nullptr;
  { int rc = pack_impl(ctx, b, PROTON_EM_HOST, chunk_nodes, &plan); if (rc) return rc; }
  struct PG { proton_em_plan* p; ~PG() { Trace t2("analyze_host"); destroy_plan(p); t2.lap("destroy_plan"); } } pg{plan};
  tr.lap("pack");
  const int64_t L = plan->L;
  const size_t LT = (size_t)L * T;
  int64_t max_nodes = 0;
  for (const auto& ch : plan->chunks) max_nodes = std::max<int64_t>(max_nodes, plan->node_off[ch.l1] - plan->node_off[ch.l0]);
  const size_t need_field = res->stress ? (size_t)max_nodes * T : 0;     // reduce-only calls need no staging field
  if (ctx->io_stress[0].n < need_field || ctx->io_stress[1].n < need_field || ctx->io_mx.n < LT || ctx->io_ts.n < (size_t)T) {
    // something has to grow: earlier work of this context may still use the old buffers
    CU_TRY(cudaStreamSynchronize(ctx->stream));
    CU_TRY(cudaStreamSynchronize(ctx->copy_stream));
    CU_TRY(grow(ctx->io_ts, (size_t)T)); CU_TRY(grow(ctx->io_mx, LT)); CU_TRY(grow(ctx->io_amx, LT)); CU_TRY(grow(ctx->io_flag, LT));
    if (res->stress) { CU_TRY(grow(ctx->io_stress[0], (size_t)max_nodes * T)); CU_TRY(grow(ctx->io_stress[1], (size_t)max_nodes * T)); }
  }
  cudaStream_t st = ctx->stream, cs = ctx->copy_stream;
  CU_TRY(cudaMemcpyAsync(ctx->io_ts.p, ts, T * sizeof(double), cudaMemcpyHostToDevice, st));
  int rc = PROTON_EM_OK;
  for (int c = 0; c < (int)plan->chunks.size() && rc == PROTON_EM_OK; ++c) {
    const auto& ch = plan->chunks[c];
    const int bi = c & 1;
    const int64_t n0 = plan->node_off[ch.l0], n1 = plan->node_off[ch.l1];
    double* d_st = nullptr;
    if (res->stress) {
      if (c >= 2) CU_TRY(cudaStreamWaitEvent(st, ctx->ev_copied[bi], 0));   // the buffer's previous chunk has left
      d_st = ctx->io_stress[bi].p - n0 * (int64_t)T;                         // kernels index by global node offset
    }
    rc = eval_device(plan, ctx->io_ts.p, T, opts, d_st, ctx->io_mx.p, ctx->io_amx.p, ctx->io_flag.p, c, c + 1);
    if (rc) break;
    if (res->stress) {
      CU_TRY(cudaEventRecord(ctx->ev_done[bi], st));
      CU_TRY(cudaStreamWaitEvent(cs, ctx->ev_done[bi], 0));
      CU_TRY(cudaMemcpyAsync(res->stress + n0 * (int64_t)T, ctx->io_stress[bi].p, (size_t)(n1 - n0) * T * sizeof(double),
                             cudaMemcpyDeviceToHost, cs));
      CU_TRY(cudaEventRecord(ctx->ev_copied[bi], cs));
    }
  }
  if (rc) { cudaDeviceSynchronize(); return rc; }
  tr.lap("evals queued");
  CU_TRY(cudaMemcpyAsync(res->max_stress, ctx->io_mx.p, LT * sizeof(double), cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaMemcpyAsync(res->argmax, ctx->io_amx.p, LT * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaMemcpyAsync(res->flag, ctx->io_flag.p, LT * sizeof(unsigned char), cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  tr.lap("main stream done");
  CU_TRY(cudaStreamSynchronize(cs));
  tr.lap("copy stream done");
  return PROTON_EM_OK;
} EM_ABI_CATCH("proton_em_analyze_host")

int proton_em_analyze_history_host(proton_em_ctx* ctx, const proton_em_batch* b, const double* const* curden_seg,
                                   const double* seg_start, int32_t num_segments, const double* ts, int32_t T,
                                   const proton_em_opts* opts, const proton_em_result* res) try {
  if (!ctx || !b || !res || !curden_seg || !seg_start)
    return fail(PROTON_EM_EINVAL, "proton_em_analyze_history_host: NULL argument");
  if (num_segments <= 0) return fail(PROTON_EM_EINVAL, "proton_em_analyze_history_host: need at least one segment");
  if (T <= 0 || b->num_lines <= 0) return PROTON_EM_OK;
  if (!ts || !res->max_stress || !res->argmax || !res->flag)
    return fail(PROTON_EM_EINVAL, "proton_em_analyze_history_host: ts, max_stress, argmax and flag are required");
  if (seg_start[0] != 0.0) return fail(PROTON_EM_EINVAL, "the first segment must start at t = 0 (got %g)", seg_start[0]);
  for (int s = 1; s < num_segments; ++s)
    if (!(seg_start[s] > seg_start[s - 1]))
      return fail(PROTON_EM_EINVAL, "segment start times must increase (segment %d: %g after %g)", s, seg_start[s], seg_start[s - 1]);
  for (int s = 0; s < num_segments; ++s)
    if (!curden_seg[s]) return fail(PROTON_EM_EINVAL, "curden_seg[%d] is NULL", s);
  { int rc = set_device(ctx); if (rc) return rc; }
  const int64_t L = b->num_lines;
  if (!b->u_off) return fail(PROTON_EM_EINVAL, "batch.u_off is NULL");
  const int64_t nu = b->u_off[L];
  static const int64_t chunk_mb = [] { const char* e = std::getenv("PROTON_EM_CHUNK_MB"); const long long v = e ? std::atoll(e) : 0; return (int64_t)(v > 0 ? v : 256); }();
  const int64_t chunk_nodes = std::max<int64_t>(1, (chunk_mb << 20) / (8 * (int64_t)T));
  // one plan per segment, packed from the current INCREMENT of that segment
  std::vector<proton_em_plan*> plans;
  struct PG { std::vector<proton_em_plan*>& v; ~PG() { for (proton_em_plan* p : v) destroy_plan(p); } } pg{plans};
  std::vector<double> incr((size_t)nu);
  std::vector<int> live;                       // segments that start before the last timestamp
  double t_last = ts[0];
  for (int j = 1; j < T; ++j) t_last = std::max(t_last, ts[j]);
  for (int s = 0; s < num_segments; ++s) {
    if (s > 0 && !(seg_start[s] < t_last)) break;
    for (int64_t i = 0; i < nu; ++i) incr[(size_t)i] = curden_seg[s][i] - (s ? curden_seg[s - 1][i] : 0.0);
    proton_em_batch bs = *b;
    bs.curden = incr.data();
    proton_em_plan* plan = nullptr;
    int rc = pack_impl(ctx, &bs, PROTON_EM_HOST, chunk_nodes, &plan);     // copies the increment to the device
    if (rc) return rc;
    plans.push_back(plan);
    live.push_back(s);
  }
  const int S = (int)plans.size();
  proton_em_plan* p0 = plans[0];
  const size_t LT = (size_t)L * T;
  int64_t max_nodes = 0;
  for (const auto& ch : p0->chunks) max_nodes = std::max<int64_t>(max_nodes, p0->node_off[ch.l1] - p0->node_off[ch.l0]);
  const size_t field = (size_t)max_nodes * T;
  if (ctx->io_ts.n < (size_t)T * S || ctx->io_mx.n < LT || ctx->io_stress[0].n < field || ctx->io_stress[1].n < field ||
      ctx->io_tmp.n < field) {
    CU_TRY(cudaStreamSynchronize(ctx->stream));
    CU_TRY(cudaStreamSynchronize(ctx->copy_stream));
  }
  CU_TRY(grow(ctx->io_ts, (size_t)T * S)); CU_TRY(grow(ctx->io_mx, LT)); CU_TRY(grow(ctx->io_amx, LT)); CU_TRY(grow(ctx->io_flag, LT));
  CU_TRY(grow(ctx->io_stress[0], field)); CU_TRY(grow(ctx->io_stress[1], field)); CU_TRY(grow(ctx->io_tmp, field));
  cudaStream_t st = ctx->stream, cs = ctx->copy_stream;
  // timestamps of segment s: time since its start, clipped at 0 (the reference solution is 0 at t = 0)
  std::vector<double> tsh((size_t)T * S);
  for (int k = 0; k < S; ++k)
    for (int j = 0; j < T; ++j) tsh[(size_t)k * T + j] = std::max(0.0, ts[j] - seg_start[live[k]]);
  CU_TRY(cudaMemcpyAsync(ctx->io_ts.p, tsh.data(), tsh.size() * sizeof(double), cudaMemcpyHostToDevice, st));
  const double critical = opts ? opts->critical_stress : NAN;          // as in proton_em_eval: NaN never flags
  for (int c = 0; c < (int)p0->chunks.size(); ++c) {
    const auto& ch = p0->chunks[c];
    const int bi = c & 1;
    const int64_t n0 = p0->node_off[ch.l0], n1 = p0->node_off[ch.l1];
    const long long count = (long long)(n1 - n0) * T;
    if (c >= 2 && res->stress) CU_TRY(cudaStreamWaitEvent(st, ctx->ev_copied[bi], 0));
    double* d_total = ctx->io_stress[bi].p - n0 * (int64_t)T;             // kernels index by global node offset
    double* d_tmp = ctx->io_tmp.p - n0 * (int64_t)T;
    for (int k = 0; k < S; ++k) {
      int rc = eval_device(plans[k], ctx->io_ts.p + (size_t)k * T, T, opts, k == 0 ? d_total : d_tmp, ctx->io_mx.p,
                           ctx->io_amx.p, ctx->io_flag.p, c, c + 1);
      if (rc) { cudaDeviceSynchronize(); return rc; }
      if (k > 0) {
        const unsigned blocks = (unsigned)std::min<long long>((count + 255) / 256, (long long)ctx->sm_count * 16);
        em::em_accumulate_kernel<<<blocks, 256, 0, st>>>(ctx->io_stress[bi].p, ctx->io_tmp.p, count);
        ctx->launches++;
        CU_TRY(cudaGetLastError());
      }
    }
    if (S > 1) {   // the fused reductions of the last evaluation describe one increment only: redo them on the total
      const long long rows = (long long)(ch.l1 - ch.l0) * T;
      if (rows > 0x7fffffffLL) return fail(PROTON_EM_EINVAL, "chunk has too many (line, timestamp) rows");
      em::RowMaxParams rp{d_total, p0->d_node_off.p, p0->d_num_nodes, (long long)ch.l0, T, ctx->io_mx.p, ctx->io_amx.p,
                          ctx->io_flag.p, critical};
      em::em_rowmax_kernel<<<(unsigned)rows, 128, 0, st>>>(rp);
      ctx->launches++;
      CU_TRY(cudaGetLastError());
    }
    if (res->stress) {
      CU_TRY(cudaEventRecord(ctx->ev_done[bi], st));
      CU_TRY(cudaStreamWaitEvent(cs, ctx->ev_done[bi], 0));
      CU_TRY(cudaMemcpyAsync(res->stress + n0 * (int64_t)T, ctx->io_stress[bi].p, (size_t)count * sizeof(double),
                             cudaMemcpyDeviceToHost, cs));
      CU_TRY(cudaEventRecord(ctx->ev_copied[bi], cs));
    }
  }
  CU_TRY(cudaMemcpyAsync(res->max_stress, ctx->io_mx.p, LT * sizeof(double), cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaMemcpyAsync(res->argmax, ctx->io_amx.p, LT * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaMemcpyAsync(res->flag, ctx->io_flag.p, LT * sizeof(unsigned char), cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  CU_TRY(cudaStreamSynchronize(cs));
  return PROTON_EM_OK;
} EM_ABI_CATCH("proton_em_analyze_history_host")

static int debug_exp_impl(proton_em_ctx* ctx, const double* x, double* y, int64_t n, int which);

int proton_em_debug_line_table(proton_em_plan* p, int64_t line, int which, double* out, int64_t capacity, int64_t* count) try {
  if (!p || !out || !count) return fail(PROTON_EM_EINVAL, "proton_em_debug_line_table: NULL argument");
  if (line < 0 || line >= p->L) return fail(PROTON_EM_EINVAL, "line %lld out of range", (long long)line);
  proton_em_ctx* ctx = p->ctx;
  { int rc = set_device(ctx); if (rc) return rc; }
  const int N = p->num_nodes[line];
  const auto it = ctx->tabs.find(N);
  if (it == ctx->tabs.end()) return fail(PROTON_EM_EINVAL, "no tables for N = %d", N);
  const em::NTab& t = it->second;
  const void* src = nullptr;
  int64_t n = 0;
  switch (which) {
    case 0: src = t.lam; n = N; break;
    case 1: src = p->d_W.p + p->node_off[line]; n = N; break;
    case 2: src = t.pre; n = t.pre ? 2LL * N : 0; break;
    case 3: src = t.post; n = t.post ? 2LL * N : 0; break;
    case 4: n = 4LL * N; break;
    case 5: {                                     // transformed chirp H: P complex values of the line's class
      const ClassOf c = classify(N);
      const long long P = c.cls == CLS_BLUE ? (1LL << c.log2p) : c.cls == CLS_LBLUE ? (long long)c.log2p * em::kLongP2
                        : c.cls == CLS_MIX ? ((long long)mix_odd(c.log2p) << mix_lr(c.log2p)) : 0;
      src = t.hhat; n = t.hhat ? 2 * P : 0;
      break;
    }
    case 6: {                                     // column twiddles w_P^{+j}, j < row length (mixed and long classes)
      const ClassOf c = classify(N);
      if (c.cls == CLS_MIX) { src = ctx->d_twm[c.log2p]; n = src ? 2LL << mix_lr(c.log2p) : 0; }
      else if (c.cls == CLS_LBLUE || c.cls == CLS_LPOW2) { src = ctx->d_twl[c.log2p] ? ctx->d_twl[c.log2p] + em::kLongP2 : nullptr; n = src ? 2LL * em::kLongP2 : 0; }
      break;
    }
    default: return fail(PROTON_EM_EINVAL, "unknown table %d", which);
  }
  *count = n;
  if (n > capacity) return fail(PROTON_EM_EINVAL, "table has %lld doubles, the buffer %lld", (long long)n, (long long)capacity);
  if (n == 0) return PROTON_EM_OK;
  CU_TRY(cudaStreamSynchronize(ctx->stream));
  if (which == 4) {
    DevBuf<double> tmp;
    struct G { DevBuf<double>& b; ~G() { b.free(); } } g{tmp};
    CU_TRY(tmp.alloc((size_t)n));
    em::em_costab_probe_kernel<<<8, 256, 0, ctx->stream>>>(tmp.p, N);
    ctx->launches++;
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaMemcpyAsync(out, tmp.p, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CU_TRY(cudaStreamSynchronize(ctx->stream));
    return PROTON_EM_OK;
  }
  CU_TRY(cudaMemcpy(out, src, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost));
  if (which == 5) {
    // H is kept lane-major per row (em_fft.cuh: hh_pos); hand it out in its logical order
    const ClassOf c = classify(N);
    const int lr = c.cls == CLS_BLUE ? c.log2p : c.cls == CLS_MIX ? mix_lr(c.log2p) : em::kLongLog2P2;
    const long long R = 1LL << lr, rows = (n / 2) / R;
    const int radix = em::plan_radix(lr, em::plan_n(lr) - 1);
    std::vector<double> tmp((size_t)n);
    std::memcpy(tmp.data(), out, (size_t)n * sizeof(double));
    for (long long g = 0; g < rows; ++g)
      for (long long i = 0; i < R; ++i) {
        const long long at = g * R + em::hh_pos_of((int)R, radix, (int)i);
        out[2 * (g * R + i)] = tmp[(size_t)(2 * at)];
        out[2 * (g * R + i) + 1] = tmp[(size_t)(2 * at + 1)];
      }
  }
  return PROTON_EM_OK;
} EM_ABI_CATCH("proton_em_debug_line_table")

int proton_em_debug_class_of(int32_t num_nodes, int32_t* kind, int32_t* rows_log2, int32_t* rows) {
  if (num_nodes < 1 || !kind || !rows_log2 || !rows) return fail(PROTON_EM_EINVAL, "proton_em_debug_class_of: bad arguments");
  const ClassOf c = classify(num_nodes);
  *rows_log2 = 0; *rows = 0;
  switch (c.cls) {
    case CLS_SMALL: *kind = 0; break;
    case CLS_POW2: *kind = 1; *rows_log2 = c.log2p; *rows = 1; break;
    case CLS_BLUE: *kind = 2; *rows_log2 = c.log2p; *rows = 1; break;
    case CLS_MIX: *kind = 3; *rows_log2 = mix_lr(c.log2p); *rows = mix_odd(c.log2p); break;
    case CLS_LPOW2: *kind = 4; *rows_log2 = em::kLongLog2P2; *rows = c.log2p; break;
    case CLS_LBLUE: *kind = 5; *rows_log2 = em::kLongLog2P2; *rows = c.log2p; break;
    default: *kind = -1; break;
  }
  return PROTON_EM_OK;
}

int proton_em_debug_exp(proton_em_ctx* ctx, const double* x, double* y, int64_t n) { return debug_exp_impl(ctx, x, y, n, 0); }
int proton_em_debug_expm1(proton_em_ctx* ctx, const double* x, double* y, int64_t n) { return debug_exp_impl(ctx, x, y, n, 1); }

static int debug_exp_impl(proton_em_ctx* ctx, const double* x, double* y, int64_t n, int which) {
  if (!ctx || !x || !y || n < 0) return fail(PROTON_EM_EINVAL, "proton_em_debug_exp: bad argument");
  if (n == 0) return PROTON_EM_OK;
  { int rc = set_device(ctx); if (rc) return rc; }
  DevBuf<double> dx, dy;
  struct G3 { DevBuf<double>&a, &b; ~G3() { a.free(); b.free(); } } g3{dx, dy};
  CU_TRY(dx.alloc(n)); CU_TRY(dy.alloc(n));
  CU_TRY(cudaMemcpyAsync(dx.p, x, n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  if (which == 0) em::em_exp_probe_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(dx.p, dy.p, n, ctx->d_exp);
  else em::em_expm1_probe_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(dx.p, dy.p, n);
  ctx->launches++;
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaMemcpyAsync(y, dy.p, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CU_TRY(cudaStreamSynchronize(ctx->stream));
  return PROTON_EM_OK;
}

}  // extern "C"

// =================================================================================================
// proton_em_eval_line_noalloc — one line, no allocation of any kind.
//
// Who this is for: code that drives the GPU from a place where device memory must not be requested —
// a host application that handed its whole memory budget to its own pool at start-up, a plug-in
// running inside a framework's stream-ordered region where cudaMalloc would synchronise the device,
// or a latency-critical loop that cannot afford an allocator call failing half-way.  Such a caller
// cannot create a proton_em_ctx (that allocates tables) nor pack a plan (that allocates the modal
// weights), so this entry point keeps ALL of its working memory — the host-libm lambda table, the
// packed r = B*u, the modal weights and the bookkeeping words of the packing / forward / direct
// kernels — in one fixed array that is part of the library image.  Every GPU that loads the library
// carries that array (about 165 KB) until the process exits.
//
// Limits: num_nodes <= 4096, num_triplets <= 4096, num_curden <= 4096, T <= 65535.
// Concurrency: every call on a device uses the same array and leaves its contents behind; two calls
// whose GPU work overlaps (different streams, different host threads) race.  Serialise them.
// The first call on a device looks up the array's device address; make that call before starting any
// stream capture.
// =================================================================================================
namespace {

constexpr int kNoAllocMaxNodes = 4096;
constexpr int kNoAllocMaxTrip = 4096;
constexpr int kNoAllocMaxDevices = 64;

struct NoAllocScratch {
  double lam[kNoAllocMaxNodes];
  double W[kNoAllocMaxNodes];
  double prod[kNoAllocMaxTrip];
  double r_val[kNoAllocMaxTrip];
  int perm[kNoAllocMaxTrip];
  int r_idx[kNoAllocMaxTrip];
  long long node_off[2];
  long long blk_off[2];
  long long trip_off[2];
  long long u_off[2];
  double a_coeff[1];
  int num_nodes[1];
  int line_tab[1];
  int r_cnt[1];
  int error_flag[1];
  em::NTab tab[1];
  ulonglong2 exp_tab[em::kExpTableEntries];
};

static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];
static __device__ const unsigned long long g_noalloc_exp_table[2 * em::kExpTableEntries] = {
#include "exp_table.inc"
};

NoAllocScratch* g_noalloc_ptr[kNoAllocMaxDevices] = {};

// Fills the bookkeeping words from kernel arguments (no host-to-device copy) and zeroes what the
// following kernels expect to start at zero.
__global__ void em_noalloc_setup_kernel(NoAllocScratch* s, int N, double a, int ntrip, int ncur) {
  const int tid = threadIdx.x;
  for (int i = tid; i < em::kExpTableEntries; i += blockDim.x)
    s->exp_tab[i] = make_ulonglong2(g_noalloc_exp_table[2 * i], g_noalloc_exp_table[2 * i + 1]);
  if (tid == 0) {
    s->node_off[0] = 0; s->node_off[1] = N;
    s->blk_off[0] = 0; s->blk_off[1] = (N + em::kFwdBlock - 1) / em::kFwdBlock;
    s->trip_off[0] = 0; s->trip_off[1] = ntrip;
    s->u_off[0] = 0; s->u_off[1] = ncur;
    s->a_coeff[0] = a; s->num_nodes[0] = N; s->line_tab[0] = 0;
    s->r_cnt[0] = 0; s->error_flag[0] = 0;          // must start at zero on every call
    em::NTab t{};
    t.lam = s->lam; t.n = N;
    s->tab[0] = t;
  }
}

}  // namespace

extern "C" int proton_em_eval_line_noalloc(int device, void* stream_v, int32_t num_nodes, double a_coeff,
                                           int32_t num_triplets, const int32_t* d_trip_row, const int32_t* d_trip_col,
                                           const double* d_trip_val, int32_t num_curden, const double* d_curden,
                                           const double* d_ts, int32_t T, double critical_stress, double* d_stress,
                                           double* d_max_stress, int32_t* d_argmax, uint8_t* d_flag) {
  if (num_nodes <= 0 || num_nodes > kNoAllocMaxNodes)
    return fail(PROTON_EM_EINVAL, "proton_em_eval_line_noalloc: num_nodes = %d, the built-in scratch holds at most %d",
                num_nodes, kNoAllocMaxNodes);
  if (num_triplets < 0 || num_triplets > kNoAllocMaxTrip || num_curden < 0 || num_curden > kNoAllocMaxTrip)
    return fail(PROTON_EM_EINVAL, "proton_em_eval_line_noalloc: %d triplets / %d current densities, the built-in scratch "
                "holds at most %d of each", num_triplets, num_curden, kNoAllocMaxTrip);
  if (T < 0 || T > 65535) return fail(PROTON_EM_EINVAL, "proton_em_eval_line_noalloc: T = %d is outside [0, 65535]", T);
  if (T == 0) return PROTON_EM_OK;
  if (!d_ts || !d_max_stress || !d_argmax || !d_flag || (num_triplets > 0 && (!d_trip_row || !d_trip_col || !d_trip_val)) ||
      (num_curden > 0 && !d_curden))
    return fail(PROTON_EM_EINVAL, "proton_em_eval_line_noalloc: NULL pointer argument");
  if (device < 0 || device >= kNoAllocMaxDevices) return fail(PROTON_EM_EINVAL, "device %d out of range", device);
  // the caller's current device is put back on every way out
  struct DeviceGuard { int prev = -1; ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); } } dguard;
  if (cudaGetDevice(&dguard.prev) != cudaSuccess) dguard.prev = -1;
  CU_TRY(cudaSetDevice(device));
  if (!g_noalloc_ptr[device]) {
    // The array's name is only a host-side handle: ask the runtime where it lives on THIS device.
    void* ptr = nullptr;
    CU_TRY(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
    g_noalloc_ptr[device] = static_cast<NoAllocScratch*>(ptr);
  }
  NoAllocScratch* s = g_noalloc_ptr[device];
  cudaStream_t st = static_cast<cudaStream_t>(stream_v);
  // lambda_k from the host libm (same reason as in ensure_tables); a fixed per-thread host buffer, no heap, so host
  // threads that serve different devices do not share it (the copy below is staged before cudaMemcpyAsync returns).
  static thread_local double h_lam[kNoAllocMaxNodes];
  host_lambda(num_nodes, h_lam);
  CU_TRY(cudaMemcpyAsync(s->lam, h_lam, (size_t)num_nodes * sizeof(double), cudaMemcpyHostToDevice, st));
  em_noalloc_setup_kernel<<<1, 128, 0, st>>>(s, num_nodes, a_coeff, num_triplets, num_curden);
  CU_TRY(cudaGetLastError());
  em::PackParams pp{1, s->num_nodes, s->trip_off, d_trip_row, d_trip_col, d_trip_val, s->u_off, d_curden,
                    s->perm, s->prod, s->r_idx, s->r_val, s->r_cnt, s->error_flag};
  em::em_pack_kernel<<<1, 128, 0, st>>>(pp);
  CU_TRY(cudaGetLastError());
  em::ForwardParams fp{1, (num_nodes + em::kFwdBlock - 1) / em::kFwdBlock, s->num_nodes, s->a_coeff, s->node_off, s->blk_off,
                       s->trip_off, s->r_idx, s->r_val, s->r_cnt, s->line_tab, s->tab, s->W};
  em::em_forward_kernel<<<(fp.total_blocks + 127) / 128, 128, 0, st>>>(fp);
  CU_TRY(cudaGetLastError());
  em::DirectParams dp{1, 0, T, s->num_nodes, s->a_coeff, s->node_off, s->line_tab, s->tab, s->W, d_ts, s->exp_tab,
                      d_stress, d_max_stress, d_argmax, d_flag, critical_stress};
  em::em_eval_direct_kernel<<<(unsigned)T, 256, 0, st>>>(dp);
  CU_TRY(cudaGetLastError());
  int err = 0;
  CU_TRY(cudaMemcpyAsync(&err, s->error_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  if (err != 0)
    return fail(PROTON_EM_EINVAL, "proton_em_eval_line_noalloc: a B triplet has a row outside [0, num_nodes) or a column "
                "outside [0, num_curden)");
  return PROTON_EM_OK;
}
