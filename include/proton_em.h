/* proton_em.h — C ABI of libproton_em.so, the B200 (sm_100a) drop-in for the line-analysis hot path
 * of oaxelou/PROTON: what `bin/EMtool_analytical` computes per interconnect line.
 *
 * Plain C: pointers and sizes only, no torch / C++ types.  Every function returns 0 on success or
 * a negative PROTON_EM_E* code; the message is available from proton_em_last_error() (thread-local).
 * Nothing here throws or calls exit().
 *
 * Level 1 (process-compatible): proton_em_run_cfg() does exactly what the reference process does for
 * one analytical.cfg; bin/EMtool_analytical (built from proton_b200/csrc/emtool_main.cpp) wraps it.
 * Level 2 (batched, new — the reference has no batch API): pack a ragged batch of lines once, then
 * evaluate it at any number of timestamps.
 */
#ifndef PROTON_EM_H
#define PROTON_EM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PROTON_EM_OK 0
#define PROTON_EM_EINVAL (-1)   /* bad argument / malformed input (reference: message + exit(1)) */
#define PROTON_EM_ENOMEM (-2)
#define PROTON_EM_ECUDA (-3)    /* CUDA runtime error, or no usable device */
#define PROTON_EM_EIO (-4)      /* file missing / unreadable / unwritable */

#define PROTON_EM_HOST 0        /* pointer arguments are host memory */
#define PROTON_EM_DEVICE 1      /* pointer arguments are device memory on the ctx's device */

#define PROTON_EM_ALG_AUTO 0    /* short-line / FFT / long-line kernels chosen per line length */
#define PROTON_EM_ALG_DIRECT 1  /* O(N^2) cosine sums for every line (cross-check only, slow) */

typedef struct proton_em_ctx proton_em_ctx;
typedef struct proton_em_plan proton_em_plan;

/* A ragged batch of lines: for line l, the inputs the reference kernel reads from disk —
 *   num_nodes[l], a_coeff[l]      `num_nodes`, `A_coeff` of analytical.cfg
 *                                 (analytical_class_cli.py:53-57; read_file, bin/EMtool_analytical@0xaaf4)
 *   triplets [trip_off[l], trip_off[l+1])   rows of B.csv `row,col,value`
 *                                 (matrix_formulation_func_cli.py:240-252; tripletParser @0xb010)
 *   curden   [u_off[l], u_off[l+1])         lines of curden.csv (:268-273; analyical_wrapper @0xba39)
 * Triplets may be unsorted and may repeat a (row, col): repeats are summed, as Eigen's
 * setFromTriplets does in the reference (@0xbbfe). */
typedef struct proton_em_batch {
  int64_t num_lines;
  const int32_t* num_nodes;   /* [L] */
  const double* a_coeff;      /* [L] */
  const int64_t* trip_off;    /* [L+1] */
  const int32_t* trip_row;
  const int32_t* trip_col;
  const double* trip_val;
  const int64_t* u_off;       /* [L+1] */
  const double* curden;
} proton_em_batch;

typedef struct proton_em_opts {
  double critical_stress;     /* flag = max > critical (PROTON_cli.py:1063); NaN: all flags 0 */
  int32_t algorithm;          /* PROTON_EM_ALG_* */
  int32_t reserved;
} proton_em_opts;

/* Outputs for L lines at T timestamps.
 *   stress      sum_l N_l * T doubles; line l starts at node_off[l]*T and is [T][N_l] row-major, i.e.
 *               row j is the content of the reference's stress_<t_j>.txt (analyical_wrapper @0xbca3).
 *               May be NULL ("reduce-only": nothing but the reductions is written).
 *   max_stress  [L*T]  max over the nodes of a row (PROTON_cli.py:1058-1061)
 *   argmax      [L*T]  first node attaining it (PROTON_cli.py:861-867)
 *   flag        [L*T]  max_stress > critical_stress (PROTON_cli.py:1063) */
typedef struct proton_em_result {
  double* stress;
  double* max_stress;
  int32_t* argmax;
  uint8_t* flag;
} proton_em_result;

/* Context: one per (process, device); owns a stream, the constant tables and a cache of per-length
 * tables (lambda_k from the host libm, chirps, transformed chirps).  Not thread-safe; distinct
 * contexts are independent.  Plans point into their context: destroy every plan before its context. */
int proton_em_create(proton_em_ctx** out, int device);
int proton_em_destroy(proton_em_ctx* ctx);
const char* proton_em_last_error(void);
const char* proton_em_version(void);

/* Pack: validates the batch, computes r = B*u per line and the forward DCT, and leaves the
 * time-independent modal weights resident in device memory.  Replaces form_analytical_solution +
 * dct1d (bin/EMtool_analytical@0xc46c, @0xc156).  `location` says where the batch arrays live.
 * Synchronous (it builds host-side tables and reports input errors).
 * Line length limit of this build: num_nodes <= 65536 for any length, and the powers of two up to 131072; a longer
 * line is rejected with PROTON_EM_EINVAL and a message naming it (the reference has no such limit in principle, but
 * its dense N x N matrices stop it near N = 30 000 on a 16 GB host).  The same limit holds for proton_em_run_cfg and
 * proton_em_analyze_host. */
int proton_em_pack(proton_em_ctx* ctx, const proton_em_batch* batch, int location, proton_em_plan** out);
/* The same with an initial stress (restart from a stress field; no counterpart in the reference, whose
 * kernel starts from zero stress — bin/EMtool_analytical@0xc68c has no sigma_0 input).  `sigma0` holds one value per
 * node, line l at node_off[l] (the layout of one timestamp's rows of proton_em_result.stress, so the field evaluated at
 * one time can be handed back in as the state to continue from, possibly with other currents).  The plan keeps its own
 * copy.  The dense forward DCT-II of the field costs O(N^2) per line, once per pack.  Evaluation of such a plan returns
 *     sigma(t) = V exp(Lambda a t) V^T sigma0  +  the reference's closed form for the batch's currents. */
int proton_em_pack_sigma0(proton_em_ctx* ctx, const proton_em_batch* batch, const double* sigma0, int location,
                          proton_em_plan** out);
int proton_em_plan_destroy(proton_em_plan* plan);
int64_t proton_em_plan_total_nodes(const proton_em_plan* plan);
int64_t proton_em_plan_num_lines(const proton_em_plan* plan);
int32_t proton_em_plan_line_nodes(const proton_em_plan* plan, int64_t line);   /* N of one line; 0 if out of range */
/* Optional line names (the reference identifies a line by the stem of its csv, PROTON_cli.py:993): the plan copies
 * the n = num_lines strings; proton_em_plan_line_name returns NULL for a plan without names or a line out of range. */
int proton_em_plan_set_line_names(proton_em_plan* plan, const char* const* names, int64_t n);
const char* proton_em_plan_line_name(const proton_em_plan* plan, int64_t line);
/* Device scratch the plan holds (bytes) — modal weights, packed r, per-line tables. */
int64_t proton_em_plan_scratch_bytes(const proton_em_plan* plan);

/* Evaluate a packed batch at T timestamps (seconds).  Replaces calc_analytical_solution + idct1d
 * (@0xc68c, @0xc2a0) and the max / threshold consumers.  `ts` and the result arrays are device
 * pointers when location == PROTON_EM_DEVICE (work is queued on the ctx stream and the call
 * returns without synchronising), host pointers otherwise (the call copies and synchronises). */
int proton_em_eval(proton_em_plan* plan, const double* ts, int32_t T, const proton_em_opts* opts,
                   const proton_em_result* result, int location);
int proton_em_synchronize(proton_em_ctx* ctx);
/* CUDA stream (cudaStream_t) all of the ctx's work is queued on, for callers that time with events. */
void* proton_em_stream(proton_em_ctx* ctx);
/* Number of kernel launches issued by this ctx so far (bench.py reports it as gpu_launches). */
int64_t proton_em_launch_count(const proton_em_ctx* ctx);
/* Piecewise-constant current history (restart in time).  No counterpart in the reference, whose
 * kernel starts from zero stress and carries one constant current (bin/EMtool_analytical@0xc68c has no sigma_0 input).
 * Segment s applies the current densities curden_seg[s] (laid out like batch->curden; batch->curden itself is ignored)
 * from seg_start[s] on; seg_start[0] must be 0 and the times must increase.  The stress equation is linear, so
 *     sigma(t) = sum over s with seg_start[s] < t of  sigma_ref[curden_seg[s] - curden_seg[s-1]](t - seg_start[s]),
 * every term being the reference's closed form for a current increment switched on at seg_start[s] (the modal state
 * never leaves the DCT basis, so no forward transform of a stress field is needed).  Each increment is evaluated with
 * the kernels of proton_em_eval into a second field, added to the total, and max / argmax / flag are taken from the
 * total.  All pointers are host pointers; `result` as in proton_em_analyze_host (stress may be NULL). */
int proton_em_analyze_history_host(proton_em_ctx* ctx, const proton_em_batch* batch, const double* const* curden_seg,
                                   const double* seg_start, int32_t num_segments, const double* ts, int32_t T,
                                   const proton_em_opts* opts, const proton_em_result* result);

/* Per-kernel timing for bench.py's roofline (no counterpart in the reference).  While enabled, every launch of
 * the evaluation path is bracketed by a CUDA event pair on the stream it is launched on; the call synchronises
 * the device, folds the finished pairs into per-kind totals, and (enable != 0 after being off) clears them.
 * proton_em_profile_read returns name / total milliseconds / launch count of kernel kind 0, 1, ... and
 * PROTON_EM_EINVAL past the last kind. */
int proton_em_profile(proton_em_ctx* ctx, int enable);
int proton_em_profile_read(const proton_em_ctx* ctx, int kind, const char** name, double* total_ms, int64_t* launches);

/* One call for host callers: pack + eval + copy back, splitting the batch so that the device-side
 * stress buffer stays below max_device_bytes (0 = default).  This is the end-to-end path bench.py
 * times as `e2e`. */
int proton_em_analyze_host(proton_em_ctx* ctx, const proton_em_batch* batch, const double* ts, int32_t T,
                           const proton_em_opts* opts, const proton_em_result* result,
                           int64_t max_device_bytes);

/* Level 1: the reference process.  Reads `cfg_path` (keys input_files, num_nodes, A_coeff,
 * simulation_time; read_file @0xaaf4), <input_files>/curden.csv and <input_files>/B.csv, computes the
 * stress at simulation_time and writes <input_files with the first "input" replaced by "output">/
 * stress_<%f of t>.txt with one "%.18g" value per line (analyical_wrapper @0xb93c-0xbec0).  On failure
 * `msg` receives the text the reference prints to stdout before exit(1). */
int proton_em_run_cfg(proton_em_ctx* ctx, const char* cfg_path, char* msg, size_t msg_len);

/* Level 1, batched: n analytical.cfg files in one call — what PROTON_cli.analyze starts as one process per line
 * (PROTON_cli.py:1174-1181).  Every cfg is read like proton_em_run_cfg reads it; the lines that read cleanly form ONE
 * batch, evaluated once at the distinct simulation times they name; every line's stress_<t>.txt is written where the
 * reference process would write it.  status[i] (may be NULL) receives the code of cfg i.  Returns the number of cfgs
 * that failed (0 = all written), or a negative PROTON_EM_E* code if the batched evaluation itself failed; `msg` holds
 * the reference's stdout text of the first failure. */
int proton_em_run_cfgs(proton_em_ctx* ctx, const char* const* cfg_paths, int64_t n, int32_t* status, char* msg,
                       size_t msg_len);

/* Convenience: pack every discretised line of a PROTON project directory in one plan.  For every sub-directory
 * <project_dir>/input/<line>/<variant>/ — variant = "<technology>_<temperature>_<width>" as the reference composes it
 * (analytical_class_cli.py:30, e.g. "CuDD_378.0_0.1"); NULL or "" = the first variant, in byte order, that holds an
 * analytical.txt — reads what Matrix_Formulation_Class wrote there (matrix_formulation_func_cli.py:240-286):
 * analytical.txt (nx_total, Acoef; read like analytical_class_cli.py:31-41), B.csv and curden.csv (read like
 * proton_em_run_cfg reads them).  Lines are taken in byte order of their directory names, which become the plan's
 * line names.  A line whose files are missing or inconsistent is left out and counted in *skipped (may be NULL); `msg`
 * receives the text of the first such failure.  Returns PROTON_EM_EIO if no line could be read. */
int proton_em_pack_from_project(proton_em_ctx* ctx, const char* project_dir, const char* variant,
                                proton_em_plan** out, int64_t* skipped, char* msg, size_t msg_len);

/* The discretiser of a whole grid in one host call: what Matrix_Formulation_Class.matrix_formulation
 * does per line in step mode (matrix_formulation_func_cli.py:24-299) with the step policy of PROTON_cli.analyze_line_item
 * (PROTON_cli.py:1000-1014), over a flat table of segments.  Host arithmetic only — needs no context and no GPU — and
 * bit-identical to the Python formulation (proton_b200/formulation.py); its outputs are the arrays of proton_em_batch.
 *   seg_off [L+1]     segments of line l are [seg_off[l], seg_off[l+1]), in file order
 *   length_um, current [K]   columns 3 and 4 of the line's csv rows (um, A)
 *   lcoord, rcoord [K][2]    integer ids of left_node.split("_")[1], [2] and of the right node's (equal id <=> equal
 *                            string: the reference compares the strings, :148-157)
 *   named [K]         both node names non-empty (the reference's truthiness test, :148) */
typedef struct proton_em_segments {
  int64_t num_lines;
  const int64_t* seg_off;
  const double* length_um;
  const double* current;
  const int64_t* lcoord;
  const int64_t* rcoord;
  const uint8_t* named;
} proton_em_segments;
#define PROTON_EM_FORM_CORRUPT 1     /* "Corrupted CSV segments file." (a line without segments or without nodes) */
#define PROTON_EM_FORM_NOT_DENSE 2   /* "The given discretization is not dense enough for the geometry of the given line." */
#define PROTON_EM_FORM_INDEX 3       /* a via before the first node: Python's negative list index; use the Python twin */
/* step_um: one step per line in um, or NULL for the whole-grid policy 0.99 x the line's shortest segment.  kappa, b:
 * technology constants (:48-73); width2 = (width * 1e-6)^2 as the reference forms it.  Outputs (caller-allocated):
 * num_nodes, a_coeff [L]; trip_off [L+1]; trip_row, trip_col, trip_val with room for 2 K entries (a line has at most two
 * per segment); curden [K] (float32-rounded; u_off of the batch = seg_off).  Returns 0, a PROTON_EM_FORM_* code with
 * *bad_line (may be NULL) = the first offending line — the batch fails as a whole, nothing usable is written — or
 * PROTON_EM_EINVAL. */
int proton_em_formulate_segments(const proton_em_segments* table, const double* step_um, double kappa, double b,
                                 double width2, int32_t* num_nodes, double* a_coeff, int64_t* trip_off,
                                 int32_t* trip_row, int32_t* trip_col, double* trip_val, double* curden,
                                 int64_t* bad_line);

/* One line with no allocation anywhere (this entry point has no counterpart in the reference, which
 * allocates freely inside its own process).  It exists for callers that must not ask the CUDA runtime for
 * memory: hosts that gave all device memory to their own pool, plug-ins called where cudaMalloc would stall
 * the device, loops that cannot tolerate an allocator failure.  No context and no plan are needed; the
 * working memory of the packing, forward-DCT and direct evaluation kernels is a fixed ~165 KB array inside
 * the library image (present on every GPU that loads the library, for the life of the process).
 * All pointers are DEVICE pointers supplied by the caller; d_stress ([T][num_nodes]) may be NULL.
 * Limits: num_nodes, num_triplets, num_curden <= 4096 each and T <= 65535; larger values are rejected with
 * PROTON_EM_EINVAL.  The call queues its kernels on `stream` (a cudaStream_t, NULL = default stream),
 * then waits for them, so results are ready on return.  Every call shares the one array and leaves its
 * contents there: calls whose GPU work could overlap (other streams or threads, same device) race and must
 * be serialised by the caller.  Host side: the lambda table is staged through a per-thread host buffer, so threads
 * that serve different devices do not disturb each other; the caller's current device is restored on return.
 * The first call on a device looks up the array's address; make it before starting any stream capture.
 * Uses the O(N^2) direct evaluator, so it is meant for single lines, not bulk. */
int proton_em_eval_line_noalloc(int device, void* stream, int32_t num_nodes, double a_coeff,
                                int32_t num_triplets, const int32_t* d_trip_row, const int32_t* d_trip_col,
                                const double* d_trip_val, int32_t num_curden, const double* d_curden,
                                const double* d_ts, int32_t T, double critical_stress, double* d_stress,
                                double* d_max_stress, int32_t* d_argmax, uint8_t* d_flag);

/* Which kernel family evaluates a line of num_nodes nodes (host logic only, no GPU needed).  kind: 0 short-line DMMA
 * kernel; 1 in-CTA FFT, N = 2^rows_log2; 2 in-CTA Bluestein, P = 2^rows_log2; 3 in-CTA Bluestein with `rows` rows of
 * 2^rows_log2 points; 4 / 5 long path (power of two / Bluestein), `rows` rows of 4096 points; -1 longer than this build
 * supports. */
int proton_em_debug_class_of(int32_t num_nodes, int32_t* kind, int32_t* rows_log2, int32_t* rows);
/* Test hook: y[i] = the device exp() used by the modal evaluation (must equal the host libm's). */
int proton_em_debug_exp(proton_em_ctx* ctx, const double* x_host, double* y_host, int64_t n);
/* Test hook: y[i] = exp(x[i]) - 1.0 exactly as the fused kernels evaluate the modal factor: bit-identical to the
 * host libm's exp(x) - 1.0 where the reference cancels (|x| < ln2/256), within 2 ulp of exp elsewhere. */
int proton_em_debug_expm1(proton_em_ctx* ctx, const double* x_host, double* y_host, int64_t n);

/* Test hook for the CPU twin of the GPU arithmetic (oracle/em_twin.c, tests/test_twin.py): copies one of the tables
 * the kernels read for `line` of a packed plan to host memory.  which: 0 lambda_k [N]; 1 modal weights W_k [N];
 * 2 pre-twiddles [N complex = 2N doubles]; 3 Bluestein post-chirp [2N doubles]; 4 the cosine table of the short-line
 * kernel, cos(pi m / 2N), m < 4N; 5 the transformed chirp H of a Bluestein class [P complex]; 6 the column twiddles
 * w_P^{+j} of a mixed-radix or long class [row length, complex].  `count` receives the number of doubles. */
int proton_em_debug_line_table(proton_em_plan* plan, int64_t line, int which, double* out, int64_t capacity, int64_t* count);

#ifdef __cplusplus
}
#endif
#endif /* PROTON_EM_H */
