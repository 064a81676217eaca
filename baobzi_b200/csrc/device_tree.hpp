// device_tree.hpp -- the flattened, device-resident image of a fitted tree and the launch API of the
// sm_100a kernels that read it (kernels.cu).  Plain C++ header: no CUDA types beyond cudaStream_t.
//
// Device layout (one replica per GPU, all arrays in HBM, small enough to live in the 126 MB L2 for
// every benchmark tree except the deliberately oversized one):
//
//   bin_root [n_bins]  u32   global index of the root node of the subtree in grid cell `bin`
//                            (= the reference's subtree_node_offsets_, template.hpp:553, 764-767)
//   nodes    [n_nodes] 64 B  one NodeRec per node in the reference's global node order
//                            (template.hpp:763-778).  The first 32-byte sector holds everything a
//                            descent step needs (centre + global index of child 0), the second sector
//                            what only a leaf needs (inverse half lengths); a leaf visit is one 64-byte
//                            request, an internal-node visit one 32-byte sector.
//   coeffs   [n_coeffs] f64  the reference's coeffs_ (template.hpp:558), blocks 16-byte aligned, layout
//                            per leaf unchanged
//   bin_leaf [n_bins]  u32   dense leaf rank of an undivided grid cell (its root is a leaf), kNotLeaf otherwise:
//                            for such cells the binned path's lookup is the floor transform plus ONE 4-byte load
//   leaf_rank[n_nodes] u32   dense rank of a leaf among leaves in node order (kNotLeaf for internal
//                            nodes) -- bin key of the binned path
//   leaf_node[n_leaves] u32  inverse of leaf_rank
#pragma once

#include "baobzi/device_layout.hpp" // NodeRec, TreeView: the device ABI (include/baobzi/)

#include <cstdint>
#include <cuda_runtime_api.h>
#include <vector>

namespace bz {

struct HostTree;

// Optional per-launch timing (bench.py's roofline): when attached to a workspace / launch, every kernel is
// bracketed by CUDA events on the launching stream; elapsed times are summed per stage after a sync.
enum Stage { kStageLookup = 0, kStageScan = 1, kStageScatter = 2, kStageEval = 3, kStageDirect = 4, kNumStages = 5 };
struct StageProfile {
    struct Span {
        int stage;
        cudaEvent_t begin, end;
        int64_t points;
    };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> pool;
    cudaEvent_t get();
    void mark(int stage, int64_t points, cudaEvent_t b, cudaEvent_t e) { spans.push_back({stage, b, e, points}); }
    // sums (ms), launch counts and point counts per stage; recycles the events.  Caller must have synchronised.
    void collect(double ms[kNumStages], int64_t launches[kNumStages], int64_t points[kNumStages]);
    ~StageProfile();
};

// Workspace of the binned path for one chunk of points (device memory, reused chunk after chunk so it
// stays resident in L2).
struct BinWorkspace {
    int64_t chunk_capacity = 0; // points
    int64_t chunk_limit = 0;    // points per chunk wanted by this call (<= capacity; 0: capacity)
    uint32_t *key = nullptr;    // [chunk]            leaf rank per point (kNotLeaf = outside the domain)
    uint32_t *perm = nullptr;   // [chunk + padding]  point index (within the chunk) per sorted slot
    uint32_t *count = nullptr;  // [n_leaves + 1]     histogram (zero between chunks: the scan clears it)
    uint32_t *cursor = nullptr; // [n_leaves + 1]     per-leaf fill cursor of the scatter (cleared by the scan)
    uint32_t *cnt = nullptr;    // [n_leaves + 1]     dense copy of the histogram for the fill kernel
    int P = 2;                  // points per lane of the evaluation kernel (tile = 32 * P slots)
    int variant = -1;           // which instantiation of the evaluation kernel (bin_eval.cu: bin_eval_select)
    bool fused = false;         // lookup + scan + scatter as one cooperative kernel (binned.cu); count / cursor
                                // are then the two histogram buffers used by alternate chunks
    bool flip = false;          // which of the two the next chunk histograms into
    int sort_ctas = 0;          // grid of the cooperative kernel (all CTAs co-resident)
    uint32_t sort_sbin = 0;     // entries of the 16-bit bin_leaf table staged in the sort's shared memory (0: gathered via L1)
    uint32_t sort_coal = 0;     // per-CTA capacity of the shared-memory local sort (0: direct 4-byte scatter)
    uint32_t sort_skey = 0;     // per-CTA capacity of the shared-memory key array (0: keys go through key[])
    size_t sort_smem = 0;       // dynamic shared memory of the cooperative kernel
    uint2 *tile_info = nullptr; // [max_tiles] {node, coefficient offset / 16 B} of each tile's leaf
    int cshift = 0;             // count[] / cursor[] entries are spread 2^cshift words apart so that atomics of
                                // different leaves land in different L2 sectors / lines
    uint32_t *start = nullptr;  // [n_leaves + 1]     padded exclusive scan
    uint32_t *tile_leaf = nullptr; // [max_tiles]     leaf rank of each warp tile
    uint32_t *n_tiles = nullptr;   // [1]
    int64_t perm_capacity = 0, tile_capacity = 0;
    StageProfile *prof = nullptr; // not owned
    // Overlap of the (memory / latency bound) sort of chunk c+1 with the (FP64 bound) evaluation of chunk c: the
    // sort runs on its own stream into the second set of buffers below, sized to be co-resident with the
    // evaluation kernel's CTAs (binned.cu: launch_eval_binned).
    bool bucketed = false;      // workspace of the bucketed path (alloc_bucket_workspace)
    uint32_t n_keys = 0;        // buckets
    int bshift = 0;             // bucket = grid cell >> bshift
    // evaluation with the bucket's nodes + coefficient blocks staged in shared memory by TMA (bucket_stage.cu)
    bool stage_ok = false;
    int stage_sorted = 1; // CTA-local order of a bucket's points: 1 by leaf, 2 by grid cell
    int stage_pipe = 0;   // stages of the pipelined kernel (2 or 3), 0 = the queue kernel with one stage
    bool stage_minb2 = false;    // the pipelined kernel's 64-register instantiation (leaves room for a sort CTA)
    bool bucket_overlap = false; // both sort launches of chunk c+1 beside the evaluation of chunk c (binned.cu)
    uint32_t *alt_start = nullptr;
    uint4 *bucket_desc = nullptr; // [n_keys] {first node, nodes, first coefficient unit (16 B), units}
    int stage_grid = 0;
    size_t stage_smem = 0;
    uint32_t stage_tree_cap = 0, stage_hist_cap = 0, stage_meta_cap = 0;
    bool overlap = false, sort_overlap_ok = false;
    int sort_threads = 0;
    cudaStream_t sort_stream = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_sorted[2] = {nullptr, nullptr}, ev_evaluated[2] = {nullptr, nullptr};
    bool evaluated_valid[2] = {false, false};
    int half = 0;
    // Split sort (binned.cu: launch_eval_split): phase A of chunk c+1 on sort_stream beside phase B + scatter of chunk c.
    // key / alt_key and base_tbl[2] alternate between chunks; three histogram buffers rotate (A of chunk c fills
    // hist3[(rot+c)%3], B of chunk c reads it and clears hist3[(rot+c+2)%3], which chunk c+2 fills next).
    bool split = false;
    bool piggy = false;        // the evaluation launches carry the sort of the following chunks (launch_eval_piggy)
    int split_ctas = 0;
    uint32_t split_sbin = 0, piggy_sbin = 0;
    uint32_t split_coal = 0;   // scatter launch: capacity of the CTA-local order in shared memory (0: direct scatter)
    size_t split_smem_a = 0, split_smem_b = 0;
    uint32_t *count3 = nullptr, *base_tbl[2] = {nullptr, nullptr};
    int rot = 0;
    uint32_t *alt_key = nullptr, *alt_perm = nullptr, *alt_tile_leaf = nullptr, *alt_n_tiles = nullptr;
    uint2 *alt_tile_info = nullptr;
};

// Side jobs an evaluation launch of chunk c carries for the chunks behind it (bin_eval.cu: bin_eval3d_piggy_kernel).
struct PiggyArgs {
    // phase B of chunk c+1 (nb == 0: none)
    uint32_t nb;
    const uint32_t *key_b;    // packed keys (sort_common.cuh: leaf | rank << 15)
    const uint32_t *count_b;  // complete histogram of chunk c+1
    uint32_t *count_clear;    // the buffer chunk c+3 will histogram into
    const uint32_t *base_b;   // [CTA][leaf] bases reserved by phase A of chunk c+1
    uint32_t *perm_b, *tile_leaf_b, *n_tiles_b;
    uint2 *tile_info_b;
    // phase A of chunk c+2 (na == 0: none)
    uint32_t na;
    const double *xa;
    double *resa;
    uint32_t *key_a;
    uint32_t *count_a, *base_a;
    int cshift;
    uint32_t sbin_n; // entries of the 16-bit bin_leaf table staged in shared memory (0: gathered through L1)
};


struct DeviceTree {
    int device = -1;
    TreeView view{};
    void *blob = nullptr; // single allocation backing every array of `view`
    size_t blob_bytes = 0;
    int sm_count = 0;
    // 1D trees with a handful of leaves: the first-component coefficients [leaf rank][order] on the host, handed to
    // eval_sorted_1d_kernel as launch parameters (constant bank); empty otherwise
    std::vector<double> leaf_coef_1d;
};

// Device allocations of the evaluation path (sort workspaces, chunk buffers).  With BAOBZI_B200_GUARD=1 in the environment
// every buffer gets a 4 KB band of a known pattern in front of and behind it; guard_check() reads the bands back and
// counts the bytes that changed (out-of-bounds writes of any kernel into the neighbourhood of a buffer) -- the pool this
// was developed on has compute-sanitizer closed, so this is the memory check the GPU test-suite runs instead.
cudaError_t bz_malloc(void **p, size_t bytes);
cudaError_t bz_free(void *p);
int guard_check(int64_t *bad_bytes, int64_t *buffers); // synchronises the device; returns cudaError_t

// flatten + upload to the current device; returns cudaError_t (0 = ok)
int upload_tree(const HostTree &h, DeviceTree &out);
void free_tree(DeviceTree &dt);

int alloc_workspace(const DeviceTree &dt, int64_t chunk, BinWorkspace &ws);
void free_workspace(BinWorkspace &ws);

// ---- kernels (all asynchronous on `stream`; return cudaError_t of the launch; *launches += kernels launched)
// Fused lookup + evaluation, one thread per point.  node_idx may be nullptr; when given, receives the
// reference's global node index per point (-1 outside).  res may be nullptr (index query only).
int launch_eval_direct(const DeviceTree &dt, const double *d_x, double *d_res, int64_t *d_node_idx, int64_t n,
                       cudaStream_t stream, int64_t *launches, StageProfile *prof = nullptr);
// 1D trees that fit in shared memory (eval1d.cu): launch_eval_direct routes to this kernel from
// table_1d_min_points() points up (env BAOBZI_B200_TABLE1D_MIN; negative = never).
bool table_1d_fits(const DeviceTree &dt);
int64_t table_1d_min_points();
int launch_eval_table_1d(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream);
// ... and for trees of at most 16 leaves: CTA-local sort by leaf, coefficients as constant-bank operands
bool sorted_1d_fits(const DeviceTree &dt);
int launch_eval_sorted_1d(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream);
// Every output component of a 1D tree: d_res[i * out_dim + k] (what baobzi_eval writes to y[k], batched).
int launch_eval_all_1d(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream,
                       int64_t *launches);
// Dependent-DFMA microbenchmark on the current device: sustained FP64 FMA rate in TFLOP/s (2 flops per FMA).
int measure_fp64_peak(double *tflops, int repeats);
int measure_fp64_peak_sustained(double *tflops, double seconds);
// Scalar twin: x[dim] -> y[out_dim] (reference Function::eval(x, res), template.hpp:851-860).
int launch_eval_scalar(const DeviceTree &dt, const double *d_x, double *d_y, cudaStream_t stream, int64_t *launches,
                       unsigned long long *d_done = nullptr, unsigned long long seq = 0);
// Binned path over n points in chunks of ws.chunk_capacity: sort by leaf (lookup+histogram, padded scan, scatter), evaluation.
int launch_eval_binned(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, int64_t n,
                       cudaStream_t stream, int64_t *launches);
// Bucketed path (trees whose nodes + coefficients do not fit in L2, any dim): points of a chunk are grouped by
// grid-cell bucket with the fused sort kernel (key = cell >> shift, no tree access), then evaluated in that order.
int alloc_bucket_workspace(const DeviceTree &dt, int64_t chunk, BinWorkspace &ws);
int launch_eval_bucketed(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, int64_t n,
                         cudaStream_t stream, int64_t *launches);
int launch_eval_buckets(const DeviceTree &dt, const double *d_x, double *d_res, const uint32_t *d_perm,
                        const uint32_t *d_key_start, uint32_t *d_queue, uint32_t n_keys, int bshift, cudaStream_t stream);
// ... the same with each bucket's piece of the tree staged in shared memory (bucket_stage.cu); plan_bucket_stage decides
// per workspace whether it applies (ws.stage_ok)
int plan_bucket_stage(const DeviceTree &dt, BinWorkspace &ws);
int launch_eval_buckets_staged(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res,
                               cudaStream_t stream, int half = 0);
int launch_eval_perm(const DeviceTree &dt, const double *d_x, double *d_res, const uint32_t *d_perm,
                     const uint32_t *d_n_slots, int64_t n_max, cudaStream_t stream);
// Evaluation fused with a reduction (reduce.cu): per-CTA partial sums of w_i * f(x_i) (d_w may be null), dot_grid(dt)
// of them per launch; launch_dot consumes an array of values instead; launch_sum_partials folds m partials into *d_out.
int dot_grid(const DeviceTree &dt);
int launch_eval_dot(const DeviceTree &dt, const double *d_x, const double *d_w, int64_t n, double *d_partial,
                    cudaStream_t stream, int64_t *launches);
int launch_dot(const DeviceTree &dt, const double *d_y, const double *d_w, int64_t n, double *d_partial,
               cudaStream_t stream, int64_t *launches);
int launch_sum_partials(const double *d_partial, int64_t m, double *d_out, cudaStream_t stream, int64_t *launches);
// true when the binned path is implemented for this (dim, order)
bool binned_supported(int dim, int order);
bool binned_fits(const DeviceTree &dt);

} // namespace bz
