// noalloc.cu -- baobzi_eval_multi_device_noalloc: the grouped (sort-by-leaf, then evaluate) path with its scratch in
// a fixed file-scope device array, for callers that may not allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
//
// The ordinary grouped path (binned.cu) keeps its keys / order / histogram in a BinWorkspace that grows with
// cudaMalloc.  Here the same three arrays live in g_scratch below: kMaxPoints keys, kMaxPoints slots and
// kMaxLeaves + 1 counters, 8.25 MB that every GPU which loads the library holds until the process exits.  The entry
// point enqueues one memset and four kernels on the caller's stream and calls nothing that allocates.
//
//   na_lookup_kernel   point -> leaf node (include/baobzi/lookup.cuh, reference template.hpp:799-820, 482-493);
//                      node[i] to scratch, histogram by dense leaf rank; NaN to res[i] for points outside
//   na_scan_kernel     exclusive scan of the histogram, in place (one CTA)
//   na_scatter_kernel  order[cursor[rank]++] = i
//   na_eval_kernel     slot t: i = order[t], evaluates point i in leaf node[i] with eval_leaf (eval_point.cuh), the
//                      arithmetic of the direct kernel, so results equal baobzi_eval_multi_device's bit for bit
//
// ONE array per device, shared by every call and every handle: it keeps the last call's contents (the counters are
// cleared at the start of each call, everything else is overwritten before it is read), and two calls that overlap
// on the GPU -- different streams, or different host threads without ordering between them -- race on it.
#include "../../include/baobzi.h"
#include "../../include/baobzi/eval_point.cuh"

#include <cstdint>
#include <cstdio>
#include <mutex>

namespace bz {

// abi.cu: the tree descriptor of the replica of `obj` on `dev` if one exists; never creates (= never allocates)
bool find_replica_view(const void *obj, int dev, TreeView *out);

namespace {

constexpr int64_t kMaxPoints = BAOBZI_NOALLOC_MAX_POINTS;
constexpr uint32_t kMaxLeaves = BAOBZI_NOALLOC_MAX_LEAVES;
constexpr size_t kNodeOff = 0;                                       // uint32 node[kMaxPoints]
constexpr size_t kOrderOff = kNodeOff + sizeof(uint32_t) * kMaxPoints; // uint32 order[kMaxPoints]
constexpr size_t kCountOff = kOrderOff + sizeof(uint32_t) * kMaxPoints; // uint32 count[kMaxLeaves + 1] (+ padding)
constexpr size_t kScratchBytes = kCountOff + sizeof(uint32_t) * (kMaxLeaves + 64);

// In host code this name is only a handle for cudaGetSymbolAddress (scratch_of below); kernels get the address.
static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char *g_scratch_addr[kMaxDevices] = {};

// device address of g_scratch on the CURRENT device, resolved once per device
cudaError_t scratch_of(int dev, unsigned char **out) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[dev]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
        if (e != cudaSuccess)
            return e;
        g_scratch_addr[dev] = static_cast<unsigned char *>(p);
    }
    *out = g_scratch_addr[dev];
    return cudaSuccess;
}

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

template <int DIM>
__global__ void __launch_bounds__(256) na_lookup_kernel(const TreeView tv, const double *__restrict__ x,
                                                        double *__restrict__ res, uint32_t n,
                                                        uint32_t *__restrict__ node, uint32_t *__restrict__ count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t rank = kNotLeaf;
    if (i < n) {
        double p[DIM];
        load_point_keep<DIM>(x, i, p);
        if (outside<DIM>(tv, p)) {
            res[i] = __longlong_as_double(0x7ff8000000000000ll);
            node[i] = kNotLeaf;
        } else {
            NodeHead head;
            const uint32_t nd = locate<DIM>(tv, p, head);
            node[i] = nd;
            rank = load_tail(tv.nodes, nd).leaf_rank;
        }
    }
    // one atomic per distinct leaf in the warp
    const uint32_t peers = __match_any_sync(0xffffffffu, rank);
    if (rank != kNotLeaf && (peers & lanemask_lt()) == 0)
        atomicAdd(count + rank, (uint32_t)__popc(peers));
}

// count[0 .. nl) -> exclusive prefix sums in place; count[nl] = points inside the domain.  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024) na_scan_kernel(uint32_t *__restrict__ count, uint32_t nl) {
    __shared__ uint32_t warp_sum[32];
    const uint32_t per = (nl + 1023u) / 1024u, lo = threadIdx.x * per, hi = min(lo + per, nl);
    uint32_t s = 0;
    for (uint32_t l = lo; l < hi; ++l)
        s += count[l];
    uint32_t inc = s;
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, inc, d);
        if ((threadIdx.x & 31) >= d)
            inc += v;
    }
    if ((threadIdx.x & 31) == 31)
        warp_sum[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t w = warp_sum[threadIdx.x];
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, w, d);
            if (threadIdx.x >= d)
                w += v;
        }
        warp_sum[threadIdx.x] = w;
    }
    __syncthreads();
    uint32_t base = inc - s + ((threadIdx.x >> 5) ? warp_sum[(threadIdx.x >> 5) - 1] : 0u);
    for (uint32_t l = lo; l < hi; ++l) {
        const uint32_t c = count[l];
        count[l] = base;
        base += c;
    }
    if (threadIdx.x == 1023)
        count[nl] = warp_sum[31];
}

__global__ void __launch_bounds__(256) na_scatter_kernel(const TreeView tv, uint32_t n, const uint32_t *__restrict__ node,
                                                         uint32_t *__restrict__ cursor, uint32_t *__restrict__ order) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t rank = kNotLeaf;
    if (i < n) {
        const uint32_t nd = node[i];
        if (nd != kNotLeaf)
            rank = load_tail(tv.nodes, nd).leaf_rank;
    }
    const uint32_t peers = __match_any_sync(0xffffffffu, rank);
    const int leader = __ffs(peers) - 1;
    uint32_t base = 0;
    if (rank != kNotLeaf && (peers & lanemask_lt()) == 0)
        base = atomicAdd(cursor + rank, (uint32_t)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (rank != kNotLeaf)
        order[base + __popc(peers & lanemask_lt())] = i;
}

// `total` = count[n_leaves] of na_scan_kernel, which the scatter's cursors never touch
template <int DIM, int ORDER>
__global__ void __launch_bounds__(128) na_eval_kernel(const TreeView tv, const double *__restrict__ x,
                                                      double *__restrict__ res, const uint32_t *__restrict__ node,
                                                      const uint32_t *__restrict__ order,
                                                      const uint32_t *__restrict__ total) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= *total)
        return;
    const uint32_t i = order[t], nd = node[i];
    double p[DIM];
    load_point_keep<DIM>(x, i, p);
    const NodeHead head = load_head(tv.nodes, nd);
    res[i] = eval_leaf<DIM, ORDER>(tv, nd, head, p);
}

template <int DIM, int ORDER>
cudaError_t enqueue(const TreeView &tv, const double *d_x, double *d_res, uint32_t n, unsigned char *scratch,
                    cudaStream_t st) {
    uint32_t *node = reinterpret_cast<uint32_t *>(scratch + kNodeOff);
    uint32_t *order = reinterpret_cast<uint32_t *>(scratch + kOrderOff);
    uint32_t *count = reinterpret_cast<uint32_t *>(scratch + kCountOff);
    cudaError_t e = cudaMemsetAsync(count, 0, sizeof(uint32_t) * (tv.n_leaves + 1), st); // must start at zero
    if (e != cudaSuccess)
        return e;
    const uint32_t g256 = (n + 255u) / 256u, g128 = (n + 127u) / 128u;
    na_lookup_kernel<DIM><<<g256, 256, 0, st>>>(tv, d_x, d_res, n, node, count);
    na_scan_kernel<<<1, 1024, 0, st>>>(count, tv.n_leaves);
    na_scatter_kernel<<<g256, 256, 0, st>>>(tv, n, node, count, order);
    na_eval_kernel<DIM, ORDER><<<g128, 128, 0, st>>>(tv, d_x, d_res, node, order, count + tv.n_leaves);
    return cudaGetLastError();
}

template <int DIM>
cudaError_t enqueue_order(const TreeView &tv, const double *d_x, double *d_res, uint32_t n, unsigned char *scratch,
                          cudaStream_t st) {
    switch (tv.order) {
    case 6: return enqueue<DIM, 6>(tv, d_x, d_res, n, scratch, st);
    case 8: return enqueue<DIM, 8>(tv, d_x, d_res, n, scratch, st);
    case 10: return enqueue<DIM, 10>(tv, d_x, d_res, n, scratch, st);
    case 12: return enqueue<DIM, 12>(tv, d_x, d_res, n, scratch, st);
    case 14: return enqueue<DIM, 14>(tv, d_x, d_res, n, scratch, st);
    case 16: return enqueue<DIM, 16>(tv, d_x, d_res, n, scratch, st);
    default: return cudaErrorNotSupported;
    }
}

int reject(const char *why, long long got, long long limit) {
    std::fprintf(stderr, "Baobzi (b200) baobzi_eval_multi_device_noalloc: %s (%lld, limit %lld)\n", why, got, limit);
    return (int)cudaErrorInvalidValue;
}

} // namespace
} // namespace bz

extern "C" int baobzi_eval_multi_device_noalloc(const baobzi_t func, const double *d_x, double *d_res, int64_t ntrg,
                                                void *cuda_stream) {
    using namespace bz;
    if (!func || !func->obj)
        return reject("NULL handle", 0, 0);
    if (ntrg <= 0)
        return 0;
    if (ntrg > kMaxPoints)
        return reject("ntrg exceeds BAOBZI_NOALLOC_MAX_POINTS: split the batch, or use baobzi_eval_multi_device",
                      (long long)ntrg, (long long)kMaxPoints);
    cudaPointerAttributes xa, ra;
    if (cudaPointerGetAttributes(&xa, d_x) != cudaSuccess || cudaPointerGetAttributes(&ra, d_res) != cudaSuccess ||
        (xa.type != cudaMemoryTypeDevice && xa.type != cudaMemoryTypeManaged) ||
        (ra.type != cudaMemoryTypeDevice && ra.type != cudaMemoryTypeManaged) || xa.device != ra.device) {
        cudaGetLastError();
        std::fprintf(stderr, "Baobzi (b200) baobzi_eval_multi_device_noalloc: d_x and d_res must be device pointers "
                             "on one device\n");
        return (int)cudaErrorInvalidDevicePointer;
    }
    const int dev = xa.device;
    if (dev < 0 || dev >= kMaxDevices)
        return reject("device ordinal out of range", dev, kMaxDevices - 1);
    TreeView tv;
    if (!find_replica_view(func->obj, dev, &tv)) {
        std::fprintf(stderr, "Baobzi (b200) baobzi_eval_multi_device_noalloc: the handle has no replica on device %d yet "
                             "and creating one allocates; call baobzi_device_view(func, %d, ...) once at set-up\n", dev, dev);
        return (int)cudaErrorInvalidValue;
    }
    if (tv.n_leaves > kMaxLeaves)
        return reject("the tree has more leaves than BAOBZI_NOALLOC_MAX_LEAVES: use baobzi_eval_multi_device",
                      (long long)tv.n_leaves, (long long)kMaxLeaves);
    int prev = 0;
    cudaGetDevice(&prev);
    if (prev != dev)
        cudaSetDevice(dev);
    unsigned char *scratch = nullptr;
    cudaError_t e = scratch_of(dev, &scratch);
    if (e != cudaSuccess) {
        std::fprintf(stderr, "Baobzi (b200) baobzi_eval_multi_device_noalloc: cudaGetSymbolAddress: %s\n",
                     cudaGetErrorString(e));
    } else {
        cudaStream_t st = (cudaStream_t)cuda_stream;
        switch (tv.dim) {
        case 1: e = enqueue_order<1>(tv, d_x, d_res, (uint32_t)ntrg, scratch, st); break;
        case 2: e = enqueue_order<2>(tv, d_x, d_res, (uint32_t)ntrg, scratch, st); break;
        case 3: e = enqueue_order<3>(tv, d_x, d_res, (uint32_t)ntrg, scratch, st); break;
        default: e = cudaErrorNotSupported;
        }
    }
    if (prev != dev)
        cudaSetDevice(prev);
    return (int)e;
}
