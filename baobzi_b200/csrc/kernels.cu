// kernels.cu -- sm_100a kernels of the batched evaluation path and their launchers.
//
// Reference path being replaced (include/baobzi_template.hpp): Function::eval(xp, res, n) 866-889 ->
// bounds test 872-873 -> get_linear_bin 799-820 -> get_node_index 482-493 -> Node::eval 366-369 ->
// cheb_eval 146-194.  See device_tree.hpp for the device layout and cheb.cuh for the arithmetic.
//
// Two execution strategies:
//   direct  one fused kernel, one thread per point: bounds -> grid cell -> descent -> affine map ->
//           Chebyshev sum with the leaf's coefficients fetched through L1/L2.  HBM-bound for 1D and
//           small 2D blocks, L2-gather-bound when every lane of a warp sits in a different leaf with a
//           multi-KB block (3D).
//   binned  for large blocks: per L2-resident chunk of points, (1) lookup + histogram by leaf,
//           (2) padded exclusive scan, (3) scatter of point indices into leaf-contiguous, warp-tile
//           aligned slots, (4) evaluation where every warp works on ONE leaf, so each 16-byte
//           coefficient load is a warp-uniform broadcast served from L1 and feeds 2*P FMAs
//           (P points per thread, register blocked).  FP64-pipe bound.
#include "device_tree.hpp"
#include "baobzi/eval_point.cuh"
#include "host_tree.hpp"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

namespace bz {

// ------------------------------------------------------------------------------------------ direct
constexpr int kDirectThreads = 128;

template <int DIM, int ORDER>
__global__ void __launch_bounds__(kDirectThreads)
eval_direct_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                   int64_t *__restrict__ node_idx, int64_t n) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        double p[DIM];
        load_point<DIM>(x, i, p);
        if (outside<DIM>(tv, p)) {
            if (res)
                __stcs(res + i, __longlong_as_double(0x7ff8000000000000ll)); // NAN, template.hpp:880-881
            if (node_idx)
                node_idx[i] = -1;
            continue;
        }
        NodeHead head;
        const uint32_t node = locate<DIM>(tv, p, head);
        if (node_idx)
            node_idx[i] = (int64_t)node;
        if (res)
            __stcs(res + i, eval_leaf<DIM, ORDER>(tv, node, head, p));
    }
}

// ------------------------------------------------------------------------------------------ bucketed
// Evaluation in the order of a permutation (binned.cu: points grouped by grid-cell bucket).  Consecutive
// threads then walk the same few subtrees and read neighbouring coefficient blocks, so the node and
// coefficient gathers of a tree that does not fit in L2 become L1 / L2 hits instead of DRAM sectors; the
// point itself is the only random access left (one 32-byte sector inside the L2-resident chunk).
constexpr int kPermThreads = 256;
template <int DIM, int ORDER>
__global__ void __launch_bounds__(kPermThreads)
eval_perm_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                 const uint32_t *__restrict__ perm, const uint32_t *__restrict__ n_slots_p) {
    // every CTA walks ONE contiguous range of slots, i.e. a few buckets one after the other: the bucket's nodes and
    // coefficient blocks are fetched from L2 once and then hit in this SM's L1
    const uint32_t n_slots = __ldg(n_slots_p);
    const uint32_t lo = (uint32_t)(((uint64_t)blockIdx.x * n_slots) / gridDim.x);
    const uint32_t hi = (uint32_t)(((uint64_t)(blockIdx.x + 1) * n_slots) / gridDim.x);
    for (uint32_t s = lo + threadIdx.x; s < hi; s += blockDim.x) {
        const uint32_t i = __ldcs(perm + s);
        double p[DIM];
        load_point_keep<DIM>(x, i, p);
        NodeHead head;
        const uint32_t node = locate<DIM>(tv, p, head);
        res[i] = eval_leaf<DIM, ORDER>(tv, node, head, p);
    }
}

// Bucket by bucket with a CTA-local counting sort by leaf: a bucket (a few grid cells) is taken from a work queue,
// its points are looked up, sorted by leaf in shared memory and evaluated in that order, so that the lanes of a
// warp sit in very few leaves and a coefficient load touches very few L1 lines.  (Thread per point in bucket order
// alone is bound by the L1 tag rate: 32 lanes x 16 bytes out of 32 different blocks per load, ncu in profiles/.)
constexpr int kBktThreads = 256, kBktBatch = 1024, kBktMaxLeaves = 2048;

template <int DIM, int ORDER>
__global__ void __launch_bounds__(kBktThreads)
eval_bucket_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                   const uint32_t *__restrict__ perm, const uint32_t *__restrict__ key_start, uint32_t *__restrict__ queue,
                   uint32_t n_keys, int bshift) {
    __shared__ double pts[DIM][kBktBatch];
    __shared__ uint32_t idx[kBktBatch], nodes[kBktBatch];
    __shared__ uint16_t order[kBktBatch];
    __shared__ uint32_t hist[kBktMaxLeaves];
    __shared__ uint32_t warp_sums[kBktThreads / 32];
    __shared__ uint32_t next_bucket;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int PER = kBktBatch / kBktThreads;       // points per thread and batch
    constexpr int HPER = kBktMaxLeaves / kBktThreads;  // histogram entries per thread in the scan
    for (;;) {
        __syncthreads();
        if (tid == 0)
            next_bucket = atomicAdd(queue, 1u);
        __syncthreads();
        const uint32_t b = next_bucket;
        if (b >= n_keys)
            break;
        const uint32_t s0 = __ldg(key_start + b), s1 = __ldg(key_start + b + 1);
        if (s0 == s1)
            continue;
        const uint32_t bin0 = b << bshift;
        const uint32_t bin1 = min(bin0 + (1u << bshift), tv.n_bins);
        const uint32_t leaf0 = __ldg(tv.bin_first_leaf + bin0);
        const uint32_t n_local = __ldg(tv.bin_first_leaf + bin1) - leaf0;
        const bool sorted = n_local <= (uint32_t)kBktMaxLeaves; // CTA uniform
        for (uint32_t bs = s0; bs < s1; bs += kBktBatch) {
            const uint32_t m = min((uint32_t)kBktBatch, s1 - bs);
            if (sorted)
                for (uint32_t l = tid; l < n_local; l += kBktThreads)
                    hist[l] = 0;
            __syncthreads();
            uint32_t lkey[PER], rank[PER];
#pragma unroll
            for (int u = 0; u < PER; ++u) {
                const uint32_t t = tid + u * kBktThreads;
                lkey[u] = 0xFFFFFFFFu;
                if (t < m) {
                    const uint32_t i = __ldcs(perm + bs + t);
                    double p[DIM];
                    load_point_keep<DIM>(x, i, p);
#pragma unroll
                    for (int j = 0; j < DIM; ++j)
                        pts[j][t] = p[j];
                    idx[t] = i;
                    NodeHead head;
                    const uint32_t node = locate<DIM>(tv, p, head);
                    nodes[t] = node;
                    if (sorted) {
                        lkey[u] = load_tail(tv.nodes, node).leaf_rank - leaf0;
                        rank[u] = atomicAdd(hist + lkey[u], 1u);
                    } else {
                        order[t] = (uint16_t)t;
                    }
                }
            }
            __syncthreads();
            if (sorted) { // exclusive scan of hist[0, n_local)
                uint32_t v[HPER], sum = 0;
#pragma unroll
                for (int e = 0; e < HPER; ++e) {
                    const uint32_t l = tid * HPER + e;
                    v[e] = l < n_local ? hist[l] : 0u;
                    sum += v[e];
                }
                uint32_t incl = sum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o)
                        incl += y;
                }
                if (lane == 31)
                    warp_sums[warp] = incl;
                __syncthreads();
                uint32_t base = incl - sum;
                for (uint32_t w = 0; w < warp; ++w)
                    base += warp_sums[w];
#pragma unroll
                for (int e = 0; e < HPER; ++e) {
                    const uint32_t l = tid * HPER + e;
                    if (l < n_local)
                        hist[l] = base;
                    base += v[e];
                }
                __syncthreads();
#pragma unroll
                for (int u = 0; u < PER; ++u)
                    if (lkey[u] != 0xFFFFFFFFu)
                        order[hist[lkey[u]] + rank[u]] = (uint16_t)(tid + u * kBktThreads);
                __syncthreads();
            }
#pragma unroll 1
            for (uint32_t t = tid; t < m; t += kBktThreads) {
                const uint32_t j = order[t];
                double p[DIM];
#pragma unroll
                for (int d = 0; d < DIM; ++d)
                    p[d] = pts[d][j];
                const uint32_t node = nodes[j];
                const NodeHead head = load_head(tv.nodes, node);
                res[idx[j]] = eval_leaf<DIM, ORDER>(tv, node, head, p);
            }
            __syncthreads();
        }
    }
}

// Scalar twin (template.hpp:851-860 -> Node::eval(x, res, coeffs) 371-375): out_dim values, block k of
// ORDER... the reference indexes `coeffs + coeff_offset + k*ORDER` for every DIM; only DIM == 1 is
// meaningful and for DIM > 1 we evaluate component 0 only.
template <int DIM, int ORDER>
__device__ __forceinline__ void eval_scalar_body(const TreeView &tv, const double *__restrict__ x, double *__restrict__ y) {
    const uint32_t k = threadIdx.x;
    if (k >= tv.out_dim)
        return;
    double p[DIM];
#pragma unroll
    for (int j = 0; j < DIM; ++j)
        p[j] = x[j];
    if (outside<DIM>(tv, p)) {
        y[k] = __longlong_as_double(0x7ff8000000000000ll);
        return;
    }
    NodeHead head;
    const uint32_t node = locate<DIM>(tv, p, head);
    if constexpr (DIM == 1) {
        const NodeTail tail = load_tail(tv.nodes, node);
        if (k >= tail.out_dim)
            return;
        const double u = __dmul_rn(__dsub_rn(p[0], head.c0), tail.ih0);
        y[k] = cheb_eval_1d<ORDER>(u, tv.coeffs + 2ull * head.coeff16 + (size_t)k * ORDER);
    } else {
        if (k == 0)
            y[0] = eval_leaf<DIM, ORDER>(tv, node, head, p);
    }
}

// `done` (optional, mapped pinned memory like x and y): set to `seq` once every component is written and visible to
// the host, which polls it instead of paying a stream synchronisation per point.
template <int DIM, int ORDER>
__global__ void eval_scalar_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ y,
                                   volatile unsigned long long *done, unsigned long long seq) {
    eval_scalar_body<DIM, ORDER>(tv, x, y);
    if (done) {
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            *done = seq;
            __threadfence_system();
        }
    }
}

// Batched twin of the scalar multi-output entry: every output component of a 1D tree for every
// point, res[i * out_dim + k] = the value baobzi_eval writes to y[k] (block k at +k*ORDER, template.hpp:371-375).
// One thread per (point, component): the lookup is redone per component (cache hits), the stores are contiguous.
template <int ORDER>
__global__ void __launch_bounds__(kDirectThreads)
eval_all_1d_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res, int64_t n) {
    const int64_t total = n * (int64_t)tv.out_dim;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int64_t i = e / tv.out_dim;
        const uint32_t k = (uint32_t)(e - i * tv.out_dim);
        double p[1] = {__ldg(x + i)};
        double y = __longlong_as_double(0x7ff8000000000000ll);
        if (!outside<1>(tv, p)) {
            NodeHead head;
            const uint32_t node = locate<1>(tv, p, head);
            const NodeTail tail = load_tail(tv.nodes, node);
            if (k < tail.out_dim) {
                const double u = __dmul_rn(__dsub_rn(p[0], head.c0), tail.ih0);
                y = cheb_eval_1d<ORDER>(u, tv.coeffs + 2ull * head.coeff16 + (size_t)k * ORDER);
            }
        }
        __stcs(res + e, y);
    }
}

#define BZ_DISPATCH_ORDER(DIM, order, CALL)                                                                            \
    switch (order) {                                                                                                   \
    case 6: { constexpr int O = 6; constexpr int D = DIM; CALL; } break;                                               \
    case 8: { constexpr int O = 8; constexpr int D = DIM; CALL; } break;                                               \
    case 10: { constexpr int O = 10; constexpr int D = DIM; CALL; } break;                                             \
    case 12: { constexpr int O = 12; constexpr int D = DIM; CALL; } break;                                             \
    case 14: { constexpr int O = 14; constexpr int D = DIM; CALL; } break;                                             \
    case 16: { constexpr int O = 16; constexpr int D = DIM; CALL; } break;                                             \
    default: return (int)cudaErrorInvalidValue;                                                                        \
    }

#define BZ_DISPATCH(dim, order, CALL)                                                                                  \
    switch (dim) {                                                                                                     \
    case 1: BZ_DISPATCH_ORDER(1, order, CALL) break;                                                                   \
    case 2: BZ_DISPATCH_ORDER(2, order, CALL) break;                                                                   \
    case 3: BZ_DISPATCH_ORDER(3, order, CALL) break;                                                                   \
    default: return (int)cudaErrorInvalidValue;                                                                        \
    }

int launch_eval_direct(const DeviceTree &dt, const double *d_x, double *d_res, int64_t *d_node_idx, int64_t n,
                       cudaStream_t stream, int64_t *launches, StageProfile *prof) {
    if (n <= 0)
        return 0;
    cudaEvent_t eb = nullptr, ee = nullptr;
    if (prof) {
        eb = prof->get();
        ee = prof->get();
        cudaEventRecord(eb, stream);
    }
    const int64_t table_min = table_1d_min_points();
    if (dt.view.dim == 1 && d_res && !d_node_idx && table_min >= 0 && n >= table_min && table_1d_fits(dt)) {
        // 1D: the tree lives in shared memory (eval1d.cu)
        const int rc = launch_eval_table_1d(dt, d_x, d_res, n, stream);
        if (rc != 0)
            return rc;
    } else {
        // enough CTAs to fill every SM several times over, capped so the grid-stride loop amortises launch cost
        const int64_t want = (n + kDirectThreads - 1) / kDirectThreads;
        const int64_t cap = (int64_t)dt.sm_count * 64;
        const unsigned grid = (unsigned)(want < cap ? want : cap);
        BZ_DISPATCH(dt.view.dim, dt.view.order,
                    (eval_direct_kernel<D, O><<<grid, kDirectThreads, 0, stream>>>(dt.view, d_x, d_res, d_node_idx, n)));
    }
    if (prof) {
        cudaEventRecord(ee, stream);
        prof->mark(kStageDirect, n, eb, ee);
    }
    if (launches)
        *launches += 1;
    return (int)cudaGetLastError();
}

int launch_eval_perm(const DeviceTree &dt, const double *d_x, double *d_res, const uint32_t *d_perm,
                     const uint32_t *d_n_slots, int64_t n_max, cudaStream_t stream) {
    const int64_t want = (n_max + kPermThreads - 1) / kPermThreads;
    const int64_t cap = (int64_t)dt.sm_count * 4;
    const unsigned grid = (unsigned)(want < cap ? (want > 0 ? want : 1) : cap);
    BZ_DISPATCH(dt.view.dim, dt.view.order,
                (eval_perm_kernel<D, O><<<grid, kPermThreads, 0, stream>>>(dt.view, d_x, d_res, d_perm, d_n_slots)));
    return (int)cudaGetLastError();
}

int launch_eval_buckets(const DeviceTree &dt, const double *d_x, double *d_res, const uint32_t *d_perm,
                        const uint32_t *d_key_start, uint32_t *d_queue, uint32_t n_keys, int bshift, cudaStream_t stream) {
    const unsigned grid = (unsigned)dt.sm_count * 6;
    BZ_DISPATCH(dt.view.dim, dt.view.order,
                (eval_bucket_kernel<D, O><<<grid, kBktThreads, 0, stream>>>(dt.view, d_x, d_res, d_perm, d_key_start, d_queue,
                                                                          n_keys, bshift)));
    return (int)cudaGetLastError();
}

int launch_eval_all_1d(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream,
                       int64_t *launches) {
    if (dt.view.dim != 1)
        return (int)cudaErrorInvalidValue;
    if (n <= 0)
        return 0;
    const int64_t want = (n * dt.view.out_dim + kDirectThreads - 1) / kDirectThreads;
    const int64_t cap = (int64_t)dt.sm_count * 64;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    BZ_DISPATCH_ORDER(1, dt.view.order,
                      ((void)D, eval_all_1d_kernel<O><<<grid, kDirectThreads, 0, stream>>>(dt.view, d_x, d_res, n)));
    if (launches)
        *launches += 1;
    return (int)cudaGetLastError();
}

int launch_eval_scalar(const DeviceTree &dt, const double *d_x, double *d_y, cudaStream_t stream, int64_t *launches,
                       unsigned long long *d_done, unsigned long long seq) {
    const unsigned threads = dt.view.out_dim < 1 ? 1 : (dt.view.out_dim > 1024 ? 1024 : dt.view.out_dim);
    BZ_DISPATCH(dt.view.dim, dt.view.order,
                (eval_scalar_kernel<D, O><<<1, threads, 0, stream>>>(dt.view, d_x, d_y, d_done, seq)));
    if (launches)
        *launches += 1;
    return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------ profiling
cudaEvent_t StageProfile::get() {
    if (!pool.empty()) {
        cudaEvent_t e = pool.back();
        pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

void StageProfile::collect(double ms[kNumStages], int64_t launches[kNumStages], int64_t points[kNumStages]) {
    for (int s = 0; s < kNumStages; ++s) {
        ms[s] = 0.0;
        launches[s] = 0;
        points[s] = 0;
    }
    for (const Span &sp : spans) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, sp.begin, sp.end) == cudaSuccess) {
            ms[sp.stage] += t;
            launches[sp.stage] += 1;
            points[sp.stage] += sp.points;
        }
    }
    // adjacent spans may share an event: recycle each one once
    std::vector<cudaEvent_t> evs;
    for (const Span &sp : spans) {
        evs.push_back(sp.begin);
        evs.push_back(sp.end);
    }
    std::sort(evs.begin(), evs.end());
    evs.erase(std::unique(evs.begin(), evs.end()), evs.end());
    pool.insert(pool.end(), evs.begin(), evs.end());
    spans.clear();
}

StageProfile::~StageProfile() {
    double ms[kNumStages];
    int64_t a[kNumStages], b[kNumStages];
    collect(ms, a, b);
    for (cudaEvent_t e : pool)
        cudaEventDestroy(e);
}

// ------------------------------------------------------------------------------------------ FP64 peak
// 16 independent dependent-FMA chains per thread: the roofline denominator for the FP64-bound kernels
// (MEASURED_PEAKS.json carries only HBM and bf16 tensor numbers).
__global__ void __launch_bounds__(256) dfma_peak_kernel(double *out, double a, double b, int iters) {
    double v[16];
#pragma unroll
    for (int k = 0; k < 16; ++k)
        v[k] = (double)(threadIdx.x + k) * 1e-3;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k)
            v[k] = __fma_rn(v[k], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 16; ++k)
        s += v[k];
    if (s == 12345.678)
        out[0] = s;
}

int measure_fp64_peak(double *tflops, int repeats) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double *d = nullptr;
    cudaError_t e = cudaMalloc((void **)&d, 8);
    if (e != cudaSuccess)
        return (int)e;
    cudaEvent_t b, f;
    cudaEventCreate(&b);
    cudaEventCreate(&f);
    const int iters = 8192, grid = sms * 8, threads = 256;
    dfma_peak_kernel<<<grid, threads>>>(d, 0.999999, 1e-9, 64); // warm-up
    double best = 0.0;
    for (int r = 0; r < repeats; ++r) {
        cudaEventRecord(b);
        dfma_peak_kernel<<<grid, threads>>>(d, 0.999999, 1e-9, iters);
        cudaEventRecord(f);
        e = cudaEventSynchronize(f);
        if (e != cudaSuccess)
            break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, b, f);
        const double flops = 2.0 * 16.0 * iters * (double)grid * threads;
        const double t = flops / (ms * 1e-3) / 1e12;
        if (t > best)
            best = t;
    }
    cudaEventDestroy(b);
    cudaEventDestroy(f);
    cudaFree(d);
    *tflops = best;
    return (int)e;
}

// The same kernel back to back for `seconds`: the rate of the last windows, i.e. at the clocks the board settles to
// under its power cap.  A kernel timed inside a long FP64-bound step runs at those clocks, not at the burst clock.
int measure_fp64_peak_sustained(double *tflops, double seconds) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double *d = nullptr;
    cudaError_t e = cudaMalloc((void **)&d, 8);
    if (e != cudaSuccess)
        return (int)e;
    cudaEvent_t b, f;
    cudaEventCreate(&b);
    cudaEventCreate(&f);
    const int iters = 8192, grid = sms * 8, threads = 256, per_window = 40;
    const double flops = 2.0 * 16.0 * iters * (double)grid * threads * per_window;
    double elapsed = 0.0, last[3] = {0.0, 0.0, 0.0};
    int windows = 0;
    if (seconds > 30.0)
        seconds = 30.0;
    while (elapsed < seconds || windows < 3) {
        cudaEventRecord(b);
        for (int r = 0; r < per_window; ++r)
            dfma_peak_kernel<<<grid, threads>>>(d, 0.999999, 1e-9, iters);
        cudaEventRecord(f);
        e = cudaEventSynchronize(f);
        if (e != cudaSuccess)
            break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, b, f);
        elapsed += ms * 1e-3;
        last[windows % 3] = flops / (ms * 1e-3) / 1e12;
        ++windows;
    }
    cudaEventDestroy(b);
    cudaEventDestroy(f);
    cudaFree(d);
    *tflops = windows >= 3 ? (last[0] + last[1] + last[2]) / 3.0 : 0.0;
    return (int)e;
}

} // namespace bz
