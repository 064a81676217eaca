/* baobzi.h -- C ABI of the B200-native libbaobzi.so (drop-in for the reference's include/baobzi.h).
 *
 * Every declaration in the "reference ABI" section is binary-compatible with the reference header
 * (citations are file:line in flatironinstitute/baobzi): same symbol names, same argument meaning,
 * same struct layouts, same error behaviour (NULL + a message on stderr; evaluation has no error
 * channel and yields NaN outside the semi-open domain [center - half_length, center + half_length)).
 * What differs is what happens behind the handle: on baobzi_init / baobzi_restore the fitted tree is
 * flattened into device arrays on a B200 and every evaluation entry point launches hand-written
 * sm_100a CUDA kernels.  There is no CPU evaluation path; without a usable CUDA device
 * baobzi_init / baobzi_restore fail (NULL).
 *
 * The "device extension" section is new: evaluation on device-resident points/results, node-index
 * queries (for parity testing) and multi-GPU point sharding.
 */
#ifndef BAOBZI_B200_BAOBZI_H
#define BAOBZI_B200_BAOBZI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ============================== reference ABI ============================================== */

/* .baobzi container version written/accepted (reference include/baobzi/header.h:4) */
#define BAOBZI_HEADER_VERSION 5

/* First msgpack object of a .baobzi file (reference include/baobzi/header.h:10-18). */
typedef struct {
    int dim;
    int order;
    int version;
} baobzi_header_t;

/* User callback y = f(x; data): x has `dim` entries, y has `output_dim` (reference baobzi.h:13). */
typedef void (*baobzi_input_func_t)(const double *x, double *y, const void *data);

/* How `tol` is interpreted by the fitter (reference baobzi.h:15-18). */
typedef enum {
    BAOBZI_TOL_RELATIVE = 0,
    BAOBZI_TOL_ABSOLUTE = 1,
} baobzi_tol_t;

/* Fit parameters (reference baobzi.h:21-38; field order and types are ABI).
 * Supported: dim in {1,2,3}; order in {6,8,10,12,14,16} (reference make_source.sh:11-13).
 * split_multi_eval is a CPU cache heuristic of the reference: accepted, persisted, ignored. */
struct baobzi_input_t {
    baobzi_input_func_t func;
    void *data;
    int dim;
    int output_dim;
    int order;
    double tol;
    double minimum_leaf_fraction;
    int split_multi_eval;
    int min_depth;
    int max_depth;
    baobzi_tol_t tol_type;
#ifdef __cplusplus
    /* same defaults as the reference's C++ constructor (baobzi.h:33-37) */
    baobzi_input_t()
        : func(nullptr), data(nullptr), dim(0), output_dim(1), order(8), tol(0.0), minimum_leaf_fraction(0.0),
          split_multi_eval(1), min_depth(0), max_depth(50), tol_type(BAOBZI_TOL_RELATIVE) {}
#endif
};
typedef struct baobzi_input_t baobzi_input_t;

/* The handle (reference baobzi.h:47-57).  Callers read DIM / OUTPUT_DIM / ORDER directly
 * (examples/C/baobzi_timing.c:44-45, src/python/__init__.py:95-97), so the layout is ABI.
 * `obj` is opaque: here it owns the host copy of the tree plus its device replicas. */
typedef struct {
    void *obj;
    int DIM;
    int OUTPUT_DIM;
    int ORDER;
    void (*eval)(const void *, const double *, double *);
    void (*eval_multi)(const void *, const double *, double *, int ntrg);
    void (*save)(const void *, const char *);
    void (*stats)(void *);
    void (*free)(void *);
} baobzi_struct;

typedef baobzi_struct *baobzi_t;

/* output_dim=1, order=8, split_multi_eval=1, max_depth=50, relative tol; everything else zero
 * (reference src/baobzi.cpp:16). */
extern const baobzi_input_t baobzi_input_default;

/* Fit `input->func` on the box center +- half_length, then flatten + upload (reference baobzi.cpp:147-184).
 * The callback runs on the host.  NULL + stderr message on: tol <= 0, NULL func, unsupported
 * (dim, order), max_depth exceeded, or no usable CUDA device. */
baobzi_t baobzi_init(const baobzi_input_t *input, const double *center, const double *half_length);

/* Read a version-5 .baobzi file, flatten + upload (reference baobzi.cpp:75-116).  NULL on failure. */
baobzi_t baobzi_restore(const char *filename);

/* Write a version-5 .baobzi file the reference can restore (reference baobzi_template.hpp:914-919). */
void baobzi_save(const baobzi_t func, const char *filename);

/* One point: x[DIM] -> y[OUTPUT_DIM] (reference baobzi.cpp:61).  Runs the same CUDA kernel with n = 1. */
void baobzi_eval(const baobzi_t func, const double *x, double *y);

/* Batched: x[ntrg][DIM] (point-major) -> res[ntrg], first output component (reference baobzi.cpp:63-65,
 * baobzi_template.hpp:866-889).  x / res are HOST pointers (pageable or pinned); the call pipelines
 * H2D copy, kernels and D2H copy in chunks and returns when res is complete.
 * Alignment: none beyond that of a double (8 bytes), as in the reference -- also for device pointers and for the
 * device entries below (a 2D point array that starts on an odd double is evaluated by kernels that load 8 bytes at
 * a time instead of the 16-byte point loads of the fast 2D paths). */
void baobzi_eval_multi(const baobzi_t func, const double *x, double *res, int ntrg);

/* Print a summary to stdout (reference baobzi.cpp:118-122); NULL-safe. */
void baobzi_stats(baobzi_t func);

/* Release host + device state; returns NULL; NULL-safe (reference baobzi.cpp:124-131). */
baobzi_t baobzi_free(baobzi_t func);

/* ============================== device extension =========================================== */

/* Batched evaluation with DEVICE pointers on the device that owns d_x (must be a device the handle
 * has a replica on; replicas are created lazily).  Asynchronous on `cuda_stream` (a cudaStream_t,
 * NULL = legacy default stream).  64-bit count.  Returns 0 or a cudaError_t value.
 * The first call on a device, and any call that needs a larger sort workspace than the calls before it, allocates
 * device memory (cudaMalloc / cudaFree: synchronous, not capturable in a CUDA graph); calls of the same or a smaller
 * size afterwards only enqueue kernels. */
int baobzi_eval_multi_device(const baobzi_t func, const double *d_x, double *d_res, int64_t ntrg, void *cuda_stream);

/* The same evaluation for callers that may not allocate: code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc.  This entry point has NO counterpart in the reference.
 * It groups the points by leaf like the binned path, but its scratch (keys, order, histogram) is one
 * fixed file-scope device array of 8.25 MB (baobzi_b200/csrc/noalloc.cu) that every GPU which loads the library holds
 * until the process exits, so the call itself makes no cudaMalloc / cudaFree or other allocating call: it enqueues
 * one memset and four kernels on `cuda_stream` and returns.  d_res[i] equals what baobzi_eval_multi_device writes, bit
 * for bit (NaN outside the domain).
 *   limits:  ntrg <= BAOBZI_NOALLOC_MAX_POINTS and a tree of at most BAOBZI_NOALLOC_MAX_LEAVES leaves; anything larger
 *            is rejected with cudaErrorInvalidValue and a line on stderr (split the batch, or use the entry above);
 *   set-up:  the handle must already have its replica on the device that owns d_x (baobzi_init / baobzi_restore
 *            create one on the then-current device; baobzi_device_view creates one anywhere) -- this call never does;
 *   sharing: ALL calls on one device, from any handle, use the same array, which keeps the last call's contents
 *            (the call clears what must start at zero).  Calls that overlap on the GPU -- different streams, or
 *            threads without ordering between them -- race on it: keep them on one stream or order them with events.
 * Returns 0 or a cudaError_t value. */
#define BAOBZI_NOALLOC_MAX_POINTS (1 << 20)
#define BAOBZI_NOALLOC_MAX_LEAVES (1 << 16)
int baobzi_eval_multi_device_noalloc(const baobzi_t func, const double *d_x, double *d_res, int64_t ntrg,
                                     void *cuda_stream);

/* All output components of a 1D multi-output tree, batched (extension): d_res[i * OUTPUT_DIM + k] is
 * what baobzi_eval (baobzi.cpp:61 -> template.hpp:851-860, 371-375) writes to y[k] for point i, bit for bit; NaN outside
 * the domain.  baobzi_eval_multi itself keeps the reference's behaviour (first component only).  DIM must be 1
 * (cudaErrorNotSupported otherwise).  Asynchronous on `cuda_stream`. */
int baobzi_eval_multi_device_all(const baobzi_t func, const double *d_x, double *d_res, int64_t ntrg, void *cuda_stream);

/* Evaluation fused with a consumer (no reference counterpart): *d_out = sum_i d_w[i] * f(d_x[i]) over
 * ntrg device-resident points, d_w == NULL for the plain sum, without an ntrg-sized result array (the values stay in
 * registers on the direct path, in an L2-sized scratch chunk on the binned / bucketed paths).  The f values are the
 * bit-exact ones of baobzi_eval_multi; a point outside the domain contributes NaN, as it would through res[].  The
 * summation order depends only on (ntrg, device, mode), so repeated calls give identical bits.  d_out is a device
 * pointer; asynchronous on `cuda_stream`. */
int baobzi_eval_multi_device_dot(const baobzi_t func, const double *d_x, const double *d_w, int64_t ntrg, double *d_out,
                                 void *cuda_stream);

/* Device-side hook: fills *view (a baobzi_device::View of include/baobzi_device.cuh, view_bytes = sizeof of it) with
 * the device layout of the replica on `device` (created on first use), for kernels of the caller that evaluate the
 * function in registers with baobzi_device::eval<DIM, ORDER>.  Valid until baobzi_free. */
int baobzi_device_view(const baobzi_t func, int device, void *view, size_t view_bytes);

/* Global node index per point, exactly Function::get_global_node_index of the reference
 * (baobzi_template.hpp:846-849; all nodes, subtree-major, breadth-first inside a subtree);
 * -1 outside the domain.  Device and host pointer variants. */
int baobzi_node_index_multi_device(const baobzi_t func, const double *d_x, int64_t *d_idx, int64_t ntrg,
                                   void *cuda_stream);
int baobzi_node_index_multi(const baobzi_t func, const double *x, int64_t *idx, int64_t ntrg);

/* Point-sharded evaluation across n_dev GPUs: shard k is d_x[k] / d_res[k] with ntrg[k] points, resident
 * on device devices[k].  The tree is replicated; there is no collective.  Launches are asynchronous on
 * per-device internal streams; baobzi_sync waits for all of them.  Returns 0 or a cudaError_t value. */
int baobzi_eval_multi_sharded(const baobzi_t func, int n_dev, const int *devices, const double *const *d_x,
                              double *const *d_res, const int64_t *ntrg);
int baobzi_sync(const baobzi_t func);

/* Tuning knobs (also readable from the environment at init/restore):
 *   mode: 0 = auto, 1 = direct (one fused kernel, thread per point), 2 = binned (points grouped by leaf in
 *   L2-resident chunks before a register-blocked evaluation), 3 = bucketed (points grouped by grid-cell bucket
 *   before a thread-per-point evaluation: trees larger than L2).  BAOBZI_B200_MODE=auto|direct|binned|bucketed.
 *   chunk: points per binned chunk (0 = default).  BAOBZI_B200_CHUNK. */
int baobzi_set_mode(baobzi_t func, int mode, int64_t chunk);

/* Size of the fitted tree behind the handle: info[0] = nodes, [1] = leaves, [2] = grid cells (subtrees),
 * [3] = doubles in the coefficient store, [4] = bytes of one device replica.  What baobzi_stats prints
 * (reference baobzi_template.hpp:582-603), machine readable.  Returns 0. */
int baobzi_tree_info(const baobzi_t func, int64_t info[5]);

/* How many GPUs a HOST-pointer baobzi_eval_multi call is spread over: 1 (default) = the current device only,
 * n = the current device and the next n - 1 visible ones, 0 = every visible device.  The call is cut into chunks
 * that worker threads on the chosen devices take from a shared counter (tree replicated lazily, no collective), so
 * an unmodified caller on an 8-GPU box uses eight PCIe links instead of one.  Also BAOBZI_B200_DEVICES=<n>|all at
 * init / restore.  Returns 0, or 1 for a negative count. */
int baobzi_set_host_devices(baobzi_t func, int n_devices);

/* Memory check for pools where compute-sanitizer is not available: with BAOBZI_B200_GUARD=1 in the environment when the
 * library is loaded, every device buffer of the evaluation path (sort workspaces, chunk buffers) is allocated between two
 * 4 KB bands of a known byte pattern.  This call synchronises, reads all bands back and returns in *bad_bytes how many
 * bytes changed (out-of-bounds writes next to a buffer) and in *buffers how many buffers are being watched.
 * Returns 0 or a cudaError_t value. */
int baobzi_b200_guard_check(int64_t *bad_bytes, int64_t *buffers);

/* Counters for tests / benchmarks: number of kernel launches issued through this handle so far. */
int64_t baobzi_launch_count(const baobzi_t func);

/* Per-kernel timing for benchmarks: between _begin and _end every kernel launched through this handle is
 * bracketed by CUDA events on its stream.  _end synchronises the current device and returns, per stage
 * (0 lookup, 1 scan, 2 scatter, 3 binned evaluation, 4 direct evaluation; arrays of BAOBZI_N_STAGES), the summed
 * elapsed milliseconds, the number of launches and the number of points processed. */
#define BAOBZI_N_STAGES 5
int baobzi_profile_begin(baobzi_t func);
int baobzi_profile_end(baobzi_t func, double *ms, int64_t *launches, int64_t *points);

/* Dependent-FMA microbenchmark on the current device: FP64 rate in TFLOP/s (roofline denominator of the
 * FP64-bound kernels).  Returns 0 or a cudaError_t value. */
int baobzi_b200_fp64_peak(double *tflops);
/* The same kernel run back to back for `seconds` (<= 30): the rate of the last windows, i.e. at the clocks the board
 * settles to under its power cap -- the denominator for a kernel timed inside a long FP64-bound step. */
int baobzi_b200_fp64_peak_sustained(double *tflops, double seconds);

/* Library identification string, e.g. "baobzi-b200 0.1 (sm_100a)". */
const char *baobzi_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif
