// bucket_stage.cu -- bucketed path, evaluation with the bucket's piece of the tree staged in shared memory by TMA.
//
// A bucket (2^bshift consecutive grid cells) owns ONE contiguous run of node records and ONE contiguous run of
// coefficient blocks (device_tree.cu stores both cell by cell).  eval_bucket_kernel (kernels.cu) reaches them through
// per-lane gathers: on a tree larger than L2 with chunks small enough for the points to stay in L2 (4 Mi points) it
// reads the tree at ~2 TB/s and that binds.  Here a CTA has one thread issue cp.async.bulk copies of
// the bucket's two runs into shared memory (completion on an mbarrier) and does everything else -- subtree descent,
// CTA-local counting sort, Chebyshev sums -- out of shared memory.  The tree is read from DRAM exactly once per chunk,
// in large sequential pieces; the only scattered global accesses left are the 16-byte point gather and the 8-byte
// result store, both inside the L2-resident chunk.  Arithmetic is eval_leaf's (eval_point.cuh): identical results.
//
// Two kernels share the per-batch routine:
//   eval_bucket_staged_kernel  one stage per CTA, buckets from a work queue; bucket -> perm -> points -> tree is a
//                              chain of dependent global latencies that only other CTAs of the SM can hide
//   eval_bucket_pipe_kernel    a ring of NS stages per CTA: while bucket k is evaluated, the tree pieces and the points
//                              of buckets k+1 .. k+NS-1 are in flight (cp.async.bulk for the tree, one cp.async per point,
//                              both completing on the stage's mbarrier) and the point indices of bucket k+NS are being
//                              loaded into registers.  Buckets are dealt round robin (ids known in advance).
// Usable when every bucket's runs fit the stage (plan_bucket_stage decides per workspace); otherwise the workspace
// keeps eval_bucket_kernel.  BAOBZI_B200_BUCKET_STAGE=0 switches both off, =2 makes "does not fit" an error (tests);
// BAOBZI_B200_BUCKET_PIPE=0 keeps the queue kernel, =2 / =3 asks for that many stages.
#include "device_tree.hpp"

#include "baobzi/cheb.cuh"
#include "baobzi/lookup.cuh"

#include <cstdio>
#include <cstdlib>

namespace bz {
namespace {

constexpr int kStgThreads = 256, kStgBatch = 512;   // queue kernel
constexpr int kPipeThreads = 512, kPipeBatch = 1024; // pipelined kernel
constexpr int kStgMaxLeaves = 2048, kStgHeader = 128;
constexpr uint32_t kStgPiece = 16384; // bytes per bulk copy (several in flight instead of one long one)
constexpr size_t kStgMaxSmem = 227 * 1024;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// The tree piece is read once per chunk and then dead, and one chunk's pieces add up to three times the L2: marked
// evict-first so that they do not push the chunk's points, indices and results (read / written at random by the same
// kernel) out of L2.  Without the hint the kernel reads 765 MB from DRAM per 4 Mi-point chunk of the 383 MB benchmark tree.
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void tma_bulk_g2s_hint(void *dst, const void *src, uint32_t bytes, uint64_t *bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
}
__device__ __forceinline__ void cp_async_16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_4(void *dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_8(void *dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
// the mbarrier gets one arrival from this thread once all its cp.async issued so far have landed (the arrival is
// part of the count the barrier was initialised with)
__device__ __forceinline__ void cp_async_arrive(uint64_t *bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

struct SmemLoad {
    __device__ __forceinline__ double2 operator()(const double2 *p) const { return *p; }
};

// load_head / load_tail of lookup.cuh for a record in shared memory
__device__ __forceinline__ NodeHead head_s(const NodeRec *nodes, uint32_t local) {
    const double2 *p = reinterpret_cast<const double2 *>(nodes + local);
    const double2 a = p[0], b = p[1];
    NodeHead h;
    h.c0 = a.x;
    h.c1 = a.y;
    h.c2 = b.x;
    const unsigned long long w = (unsigned long long)__double_as_longlong(b.y);
    h.child0 = (uint32_t)w;
    h.coeff16 = (uint32_t)(w >> 32);
    return h;
}
__device__ __forceinline__ NodeTail tail_s(const NodeRec *nodes, uint32_t local) {
    const double2 *p = reinterpret_cast<const double2 *>(nodes + local) + 2;
    const double2 a = p[0], b = p[1];
    NodeTail t;
    t.ih0 = a.x;
    t.ih1 = a.y;
    t.ih2 = b.x;
    const unsigned long long w = (unsigned long long)__double_as_longlong(b.y);
    t.out_dim = (uint32_t)w;
    t.leaf_rank = (uint32_t)(w >> 32);
    return t;
}

// {first node, nodes, first coefficient unit (16 B), coefficient units} of every bucket; maxima[0] = largest stage in
// bytes, maxima[1] = most leaves in one bucket
__global__ void bucket_describe_kernel(const TreeView tv, uint32_t n_keys, int bshift, uint32_t blk16, uint4 *desc,
                                       uint32_t *maxima) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_keys)
        return;
    const uint32_t bin0 = b << bshift, bin1 = min(bin0 + (1u << bshift), tv.n_bins);
    const uint32_t node0 = tv.bin_root[bin0], node1 = bin1 < tv.n_bins ? tv.bin_root[bin1] : tv.n_nodes;
    const uint32_t leaf0 = tv.bin_first_leaf[bin0], leaf1 = tv.bin_first_leaf[bin1];
    uint32_t c0 = 0, cn = 0;
    if (leaf1 > leaf0) { // blocks ascend with the node order (device_tree.cu: upload_tree)
        c0 = tv.nodes[tv.leaf_node[leaf0]].coeff16;
        cn = tv.nodes[tv.leaf_node[leaf1 - 1]].coeff16 + blk16 - c0;
    }
    desc[b] = make_uint4(node0, node1 - node0, c0, cn);
    atomicMax(maxima, (node1 - node0) * 64u + cn * 16u);
    atomicMax(maxima + 1, leaf1 - leaf0);
}

// Thread 0: the bucket's node run and coefficient run into `tree_s`, completion (transaction bytes) on `bar`.
__device__ __forceinline__ void issue_tree(const TreeView &tv, const uint4 d, unsigned char *tree_s, uint64_t *bar) {
    const uint32_t nbytes = d.y * 64u, cbytes = d.w * 16u;
    mbar_expect_tx(bar, nbytes + cbytes);
    const uint64_t pol = l2_evict_first_policy();
    const unsigned char *gn = reinterpret_cast<const unsigned char *>(tv.nodes + d.x);
    for (uint32_t o = 0; o < nbytes; o += kStgPiece)
        tma_bulk_g2s_hint(tree_s + o, gn + o, min(kStgPiece, nbytes - o), bar, pol);
    const unsigned char *gc = reinterpret_cast<const unsigned char *>(tv.coeffs + 2ull * d.z);
    for (uint32_t o = 0; o < cbytes; o += kStgPiece)
        tma_bulk_g2s_hint(tree_s + nbytes + o, gc + o, min(kStgPiece, cbytes - o), bar, pol);
}

// Work arrays of one CTA (behind the stages in shared memory).
struct BatchWork {
    uint32_t *nloc;      // [B] BYBIN: grid cell of the point relative to the bucket's first; else its leaf node, relative
    uint32_t *hist;      // [hist_cap]
    uint16_t *order;     // [B]
    uint32_t *warp_sums; // [T / 32]
};

// One batch of m <= B points of a bucket whose tree piece sits in shared memory: pts[t][DIM] / idx[t] (point and its
// index within the chunk, visible to every thread on entry) -> res[idx[t]].  CTA-local counting sort, then evaluation in
// sorted order.  BYBIN: the key is the grid cell (computed from the point alone) and the descent happens AFTER the
// sort: the lanes of a warp then walk the same one or two subtrees, so their reads of node records (64-byte stride:
// 16-way bank conflicts when 32 lanes read 32 different records) and of coefficient blocks are broadcasts.  Good for
// trees with a handful of leaves per cell (the > L2 benchmark tree: 4); trees with deep subtrees sort by leaf (descent
// before the sort, in arrival order).  No barrier at the end: the caller synchronises before the arrays are reused.
template <int DIM, int ORDER, bool BYBIN, int T, int B>
__device__ __forceinline__ void process_batch(const TreeView &tv, double *__restrict__ res, uint32_t m, const double *pts,
                                              const uint32_t *idx, const BatchWork &w, const NodeRec *nodes_s,
                                              const double *coef_s, const uint4 d, uint32_t bin0, uint32_t leaf0,
                                              uint32_t n_local, const uint32_t *roots) {
    // `roots`: the bucket's slice of bin_root in shared memory, or null (then read through L1).  hist[0, n_local) is
    // zero on entry and again on exit.
    constexpr int PER = B / T, HPER = kStgMaxLeaves / T;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t lkey[PER], rank[PER];
#pragma unroll
    for (int u = 0; u < PER; ++u) {
        const uint32_t t = tid + u * T;
        lkey[u] = 0xFFFFFFFFu;
        if (t < m) {
            double p[DIM];
#pragma unroll
            for (int j = 0; j < DIM; ++j)
                p[j] = pts[t * DIM + j];
            const uint32_t lin = linear_bin<DIM>(tv, p);
            if (BYBIN) {
                lkey[u] = lin - bin0;
                w.nloc[t] = lkey[u];
            } else { // locate() of lookup.cuh with the records in shared memory
                uint32_t node = (roots ? roots[lin - bin0] : __ldg(tv.bin_root + lin)) - d.x;
                NodeHead head = head_s(nodes_s, node);
                while (head.child0 != kLeafMark) {
                    uint32_t bits = (uint32_t)(p[0] > head.c0);
                    if (DIM > 1)
                        bits |= (uint32_t)(p[1] > head.c1) << 1;
                    if (DIM > 2)
                        bits |= (uint32_t)(p[2] > head.c2) << 2;
                    node = head.child0 - d.x + bits;
                    head = head_s(nodes_s, node);
                }
                w.nloc[t] = node;
                lkey[u] = tail_s(nodes_s, node).leaf_rank - leaf0;
            }
            rank[u] = atomicAdd(w.hist + lkey[u], 1u);
        }
    }
    __syncthreads();
    if (n_local <= 32u) { // every warp scans the few counters itself: no second pass through shared memory, no barrier
        const uint32_t v = lane < n_local ? w.hist[lane] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o)
                incl += y;
        }
        const uint32_t excl = incl - v;
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const uint32_t base = __shfl_sync(0xffffffffu, excl, lkey[u] & 31u);
            if (lkey[u] != 0xFFFFFFFFu)
                w.order[base + rank[u]] = (uint16_t)(tid + u * T);
        }
        __syncthreads();
    } else { // exclusive scan of hist[0, n_local)
        uint32_t v[HPER], sum = 0;
#pragma unroll
        for (int e = 0; e < HPER; ++e) {
            const uint32_t l = tid * HPER + e;
            v[e] = l < n_local ? w.hist[l] : 0u;
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
            w.warp_sums[warp] = incl;
        __syncthreads();
        uint32_t base = incl - sum;
        for (uint32_t k = 0; k < warp; ++k)
            base += w.warp_sums[k];
#pragma unroll
        for (int e = 0; e < HPER; ++e) {
            const uint32_t l = tid * HPER + e;
            if (l < n_local)
                w.hist[l] = base;
            base += v[e];
        }
        __syncthreads();
#pragma unroll
        for (int u = 0; u < PER; ++u)
            if (lkey[u] != 0xFFFFFFFFu)
                w.order[w.hist[lkey[u]] + rank[u]] = (uint16_t)(tid + u * T);
        __syncthreads();
    }
    for (uint32_t l = tid; l < n_local; l += T) // for the next batch (last read before the barrier above)
        w.hist[l] = 0;
#pragma unroll 1
    for (uint32_t t = tid; t < m; t += T) { // eval_leaf of eval_point.cuh out of the stage
        const uint32_t j = w.order[t];
        double p[DIM];
#pragma unroll
        for (int k = 0; k < DIM; ++k)
            p[k] = pts[j * DIM + k];
        uint32_t node = w.nloc[j];
        NodeHead head;
        if (BYBIN) { // nloc holds the grid cell: descend now, next to lanes that walk the same subtree
            node = (roots ? roots[node] : __ldg(tv.bin_root + bin0 + node)) - d.x;
            head = head_s(nodes_s, node);
            while (head.child0 != kLeafMark) {
                uint32_t bits = (uint32_t)(p[0] > head.c0);
                if (DIM > 1)
                    bits |= (uint32_t)(p[1] > head.c1) << 1;
                if (DIM > 2)
                    bits |= (uint32_t)(p[2] > head.c2) << 2;
                node = head.child0 - d.x + bits;
                head = head_s(nodes_s, node);
            }
        } else {
            head = head_s(nodes_s, node);
        }
        const NodeTail tail = tail_s(nodes_s, node);
        const double *C = coef_s + 2ull * (head.coeff16 - d.z);
        const double u0 = __dmul_rn(__dsub_rn(p[0], head.c0), tail.ih0);
        const double u1 = __dmul_rn(__dsub_rn(p[1], head.c1), tail.ih1);
        double y;
        if constexpr (DIM == 2) {
            y = cheb_eval_2d<ORDER>(u0, u1, C, SmemLoad());
        } else {
            const double u2 = __dmul_rn(__dsub_rn(p[2], head.c2), tail.ih2);
            y = cheb_eval_3d<ORDER>(u0, u1, u2, C, SmemLoad());
        }
        res[idx[j]] = y;
    }
}

// Plain (synchronous) load of one batch of a bucket's points into pts / idx.
template <int DIM, int T, int B>
__device__ __forceinline__ void load_batch(const double *__restrict__ x, const uint32_t *__restrict__ perm, uint32_t bs,
                                           uint32_t m, double *pts, uint32_t *idx) {
    constexpr int PER = B / T;
    const uint32_t tid = threadIdx.x;
    uint32_t ii[PER];
    double p[PER][DIM];
#pragma unroll
    for (int u = 0; u < PER; ++u) {
        const uint32_t t = tid + u * T;
        ii[u] = t < m ? __ldcs(perm + bs + t) : 0u;
    }
#pragma unroll
    for (int u = 0; u < PER; ++u) // every load of a thread is issued before the first is used
        if (tid + u * T < m)
            load_point_keep<DIM>(x, ii[u], p[u]);
#pragma unroll
    for (int u = 0; u < PER; ++u) {
        const uint32_t t = tid + u * T;
        if (t < m) {
#pragma unroll
            for (int j = 0; j < DIM; ++j)
                pts[t * DIM + j] = p[u][j];
            idx[t] = ii[u];
        }
    }
}

// ------------------------------------------------------------------------------------------ queue kernel (one stage)
template <int DIM, int ORDER, bool BYBIN>
__global__ void __launch_bounds__(kStgThreads)
eval_bucket_staged_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                          const uint32_t *__restrict__ perm, const uint32_t *__restrict__ key_start,
                          uint32_t *__restrict__ queue, const uint4 *__restrict__ desc, uint32_t n_keys, int bshift,
                          uint32_t tree_cap, uint32_t hist_cap) {
    constexpr int B = kStgBatch, T = kStgThreads;
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem);
    uint32_t *next_bucket = reinterpret_cast<uint32_t *>(smem + 8);
    unsigned char *tree_s = smem + kStgHeader;
    double *pts = reinterpret_cast<double *>(tree_s + tree_cap); // [B][DIM]
    uint32_t *idx = reinterpret_cast<uint32_t *>(pts + DIM * B);  // [B] point index within the chunk
    BatchWork w;
    w.warp_sums = reinterpret_cast<uint32_t *>(smem + 16);
    w.nloc = idx + B;
    w.hist = w.nloc + B;
    w.order = reinterpret_cast<uint16_t *>(w.hist + hist_cap);
    const uint32_t tid = threadIdx.x;
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init_fence();
    }
    for (uint32_t l = tid; l < hist_cap; l += T)
        w.hist[l] = 0;
    uint32_t parity = 0;
    for (;;) {
        __syncthreads(); // the previous bucket's stage and batch arrays are no longer read
        if (tid == 0)
            *next_bucket = atomicAdd(queue, 1u);
        __syncthreads();
        const uint32_t b = *next_bucket;
        if (b >= n_keys)
            break;
        const uint32_t s0 = __ldg(key_start + b), s1 = __ldg(key_start + b + 1);
        if (s0 == s1)
            continue;
        const uint4 d = __ldg(desc + b);
        const NodeRec *nodes_s = reinterpret_cast<const NodeRec *>(tree_s);
        const double *coef_s = reinterpret_cast<const double *>(tree_s + d.y * 64u);
        if (tid == 0)
            issue_tree(tv, d, tree_s, bar);
        const uint32_t bin0 = b << bshift, bin1 = min(bin0 + (1u << bshift), tv.n_bins);
        const uint32_t leaf0 = __ldg(tv.bin_first_leaf + bin0);
        const uint32_t n_local = BYBIN ? bin1 - bin0 : __ldg(tv.bin_first_leaf + bin1) - leaf0;
        bool staged = false;
        for (uint32_t bs = s0; bs < s1; bs += B) {
            const uint32_t m = min((uint32_t)B, s1 - bs);
            load_batch<DIM, T, B>(x, perm, bs, m, pts, idx);
            if (!staged) { // every thread observes the completion itself: the copied bytes are then visible to it
                mbar_wait(bar, parity);
                parity ^= 1u;
                staged = true;
            }
            __syncthreads();
            process_batch<DIM, ORDER, BYBIN, T, B>(tv, res, m, pts, idx, w, nodes_s, coef_s, d, bin0, leaf0, n_local, nullptr);
            if (bs + B < s1)
                __syncthreads(); // the next batch overwrites pts / idx and the work arrays
        }
    }
}

// ------------------------------------------------------------------------------------------ pipelined kernel (NS stages)
// CTA c evaluates buckets c, c + grid, c + 2 grid, ...  Stage s = k mod NS holds bucket k of the CTA: its tree piece,
// the first B of its points and their indices.  Iteration k: start the copies of bucket k + NS - 1 (its point indices
// were loaded into registers an iteration ago), start loading the indices of bucket k + NS, wait for stage s, evaluate.
// A bucket with more than B points takes its further batches with plain loads (tree piece already staged).
// MINB = 2 caps the kernel at 64 registers per thread: a 512-thread sort CTA then fits the SM beside it (overlap mode).
template <int DIM, int ORDER, bool BYBIN, int NS, int MINB>
__global__ void __launch_bounds__(kPipeThreads, MINB)
eval_bucket_pipe_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                        const uint32_t *__restrict__ perm, const uint32_t *__restrict__ key_start,
                        const uint4 *__restrict__ desc, uint32_t n_keys, int bshift, uint32_t tree_cap, uint32_t hist_cap,
                        uint32_t meta_cap) {
    constexpr int B = kPipeBatch, T = kPipeThreads, PER = B / T;
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem); // [NS]
    const uint32_t cells_cap = 1u << bshift;
    // stage: tree piece | points [B][DIM] | point indices [B] | the bucket's slice of bin_root [cells_cap]
    const uint32_t stage_bytes = tree_cap + B * DIM * 8u + B * 4u + ((cells_cap * 4u + 127u) & ~127u);
    unsigned char *stage0 = smem + kStgHeader;
    BatchWork w;
    w.warp_sums = reinterpret_cast<uint32_t *>(smem + 32);
    w.nloc = reinterpret_cast<uint32_t *>(stage0 + (size_t)NS * stage_bytes);
    w.hist = w.nloc + B;
    w.order = reinterpret_cast<uint16_t *>(w.hist + hist_cap);
    // every bucket of this CTA, described once: {first slot, end slot, desc.x .. desc.w, first leaf, leaves}
    uint32_t *meta = reinterpret_cast<uint32_t *>(reinterpret_cast<unsigned char *>(w.order) + ((B * 2u + 15u) & ~15u));
    const uint32_t tid = threadIdx.x, G = gridDim.x;
    const uint32_t n_mine = blockIdx.x < n_keys ? min((n_keys - blockIdx.x + G - 1) / G, meta_cap) : 0u;
    if (tid == 0) {
        for (int s = 0; s < NS; ++s)
            mbar_init(bars + s, T + 1); // thread 0's expect_tx arrival + one cp.async arrival per thread
        mbar_init_fence();
    }
    for (uint32_t l = tid; l < hist_cap; l += T)
        w.hist[l] = 0;
    for (uint32_t k = tid; k < n_mine; k += T) {
        const uint32_t b = blockIdx.x + k * G;
        const uint4 d = __ldg(desc + b);
        const uint32_t bin0 = b << bshift, bin1 = min(bin0 + cells_cap, tv.n_bins);
        const uint32_t leaf0 = __ldg(tv.bin_first_leaf + bin0);
        uint32_t *mk = meta + 8 * k;
        mk[0] = __ldg(key_start + b);
        mk[1] = __ldg(key_start + b + 1);
        mk[2] = d.x;
        mk[3] = d.y;
        mk[4] = d.z;
        mk[5] = d.w;
        mk[6] = leaf0;
        mk[7] = __ldg(tv.bin_first_leaf + bin1) - leaf0;
    }
    __syncthreads();
    auto load_perm = [&](uint32_t k, uint32_t (&pi)[PER]) {
        const uint32_t s0 = meta[8 * k], m = min((uint32_t)B, meta[8 * k + 1] - s0);
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const uint32_t t = tid + u * T;
            pi[u] = t < m ? __ldcs(perm + s0 + t) : 0u;
        }
    };
    auto issue = [&](uint32_t k, const uint32_t (&pi)[PER]) {
        const uint32_t b = blockIdx.x + k * G, s = k % NS;
        const uint32_t *mk = meta + 8 * k;
        const uint32_t m = min((uint32_t)B, mk[1] - mk[0]);
        unsigned char *st = stage0 + (size_t)s * stage_bytes;
        double *pts = reinterpret_cast<double *>(st + tree_cap);
        uint32_t *idx = reinterpret_cast<uint32_t *>(pts + B * DIM);
        uint32_t *roots = idx + B;
        if (tid == 0) {
            if (m)
                issue_tree(tv, make_uint4(mk[2], mk[3], mk[4], mk[5]), st, bars + s);
            else
                mbar_expect_tx(bars + s, 0u);
        }
        if (m) {
            const uint32_t bin0 = b << bshift, cells = min(cells_cap, tv.n_bins - bin0);
            for (uint32_t c = tid; c < cells; c += T)
                cp_async_4(roots + c, tv.bin_root + bin0 + c);
        }
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const uint32_t t = tid + u * T;
            if (t < m) {
                idx[t] = pi[u];
                if constexpr (DIM == 2) { // the bucketed path never sees a 2D array on an odd double (abi.cu: misaligned_2d)
                    cp_async_16(pts + 2 * t, x + 2ull * pi[u]);
                } else {
#pragma unroll
                    for (int j = 0; j < DIM; ++j)
                        cp_async_8(pts + DIM * t + j, x + (size_t)DIM * pi[u] + j);
                }
            }
        }
        cp_async_arrive(bars + s);
    };
    uint32_t pa[PER];
    for (uint32_t k = 0; k + 1 < (uint32_t)NS && k < n_mine; ++k) {
        load_perm(k, pa);
        issue(k, pa);
    }
    if ((uint32_t)NS - 1 < n_mine)
        load_perm(NS - 1, pa);
    __syncthreads();
    for (uint32_t k = 0; k < n_mine; ++k) {
        const uint32_t s = k % NS;
        if (k + NS - 1 < n_mine)
            issue(k + NS - 1, pa); // into the stage bucket k - 1 left at the barrier that ended the last iteration
        if (k + NS < n_mine)
            load_perm(k + NS, pa);
        const uint32_t b = blockIdx.x + k * G;
        const uint32_t *mk = meta + 8 * k;
        const uint32_t s0 = mk[0], s1 = mk[1];
        mbar_wait(bars + s, (k / NS) & 1u);
        if (s0 != s1) {
            const uint4 d = make_uint4(mk[2], mk[3], mk[4], mk[5]);
            unsigned char *st = stage0 + (size_t)s * stage_bytes;
            const NodeRec *nodes_s = reinterpret_cast<const NodeRec *>(st);
            const double *coef_s = reinterpret_cast<const double *>(st + d.y * 64u);
            double *pts = reinterpret_cast<double *>(st + tree_cap);
            uint32_t *idx = reinterpret_cast<uint32_t *>(pts + B * DIM);
            const uint32_t *roots = idx + B;
            const uint32_t bin0 = b << bshift, bin1 = min(bin0 + cells_cap, tv.n_bins);
            const uint32_t n_local = BYBIN ? bin1 - bin0 : mk[7];
            for (uint32_t bs = s0; bs < s1; bs += B) {
                const uint32_t m = min((uint32_t)B, s1 - bs);
                if (bs != s0) {
                    __syncthreads();
                    load_batch<DIM, T, B>(x, perm, bs, m, pts, idx);
                    __syncthreads();
                }
                process_batch<DIM, ORDER, BYBIN, T, B>(tv, res, m, pts, idx, w, nodes_s, coef_s, d, bin0, mk[6], n_local, roots);
            }
        }
        __syncthreads();
    }
    // buckets beyond the description table (a grid much smaller than planned): never with the planned launch
}

// ------------------------------------------------------------------------------------------ host
size_t queue_smem(int dim, uint32_t tree_cap, uint32_t hist_cap) {
    return (size_t)kStgHeader + tree_cap + (size_t)kStgBatch * (dim * 8 + 4 + 4) + (size_t)hist_cap * 4 +
           (size_t)kStgBatch * 2;
}
size_t pipe_smem(int dim, int ns, uint32_t tree_cap, uint32_t hist_cap, int bshift, uint32_t meta_cap) {
    const size_t roots = (((size_t)4 << bshift) + 127) & ~(size_t)127;
    return (size_t)kStgHeader + (size_t)ns * (tree_cap + (size_t)kPipeBatch * (dim * 8 + 4) + roots) + (size_t)kPipeBatch * 4 +
           (size_t)hist_cap * 4 + (((size_t)kPipeBatch * 2 + 15) & ~(size_t)15) + (size_t)meta_cap * 32;
}

template <int DIM, int ORDER>
const void *kernel_of(int sorted, int pipe, bool overlap) {
    const bool bybin = sorted == 2;
    if (pipe == 3)
        return bybin ? (const void *)eval_bucket_pipe_kernel<DIM, ORDER, true, 3, 1> : (const void *)eval_bucket_pipe_kernel<DIM, ORDER, false, 3, 1>;
    if (pipe == 2 && overlap)
        return bybin ? (const void *)eval_bucket_pipe_kernel<DIM, ORDER, true, 2, 2> : (const void *)eval_bucket_pipe_kernel<DIM, ORDER, false, 2, 2>;
    if (pipe == 2)
        return bybin ? (const void *)eval_bucket_pipe_kernel<DIM, ORDER, true, 2, 1> : (const void *)eval_bucket_pipe_kernel<DIM, ORDER, false, 2, 1>;
    return bybin ? (const void *)eval_bucket_staged_kernel<DIM, ORDER, true> : (const void *)eval_bucket_staged_kernel<DIM, ORDER, false>;
}

template <int DIM, int ORDER>
int configure(const DeviceTree &dt, BinWorkspace &ws) {
    const void *kern = kernel_of<DIM, ORDER>(ws.stage_sorted, ws.stage_pipe, ws.stage_minb2);
    const int threads = ws.stage_pipe ? kPipeThreads : kStgThreads;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws.stage_smem);
    if (e != cudaSuccess)
        return (int)e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, ws.stage_smem);
    if (e != cudaSuccess)
        return (int)e;
    ws.stage_grid = ws.stage_pipe ? (per_sm > 0 ? dt.sm_count : 0) : dt.sm_count * per_sm;
    return 0;
}

template <int DIM, int ORDER>
int launch(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st, int half) {
    uint32_t grid = (uint32_t)ws.stage_grid;
    grid = grid < ws.n_keys ? grid : ws.n_keys;
    const TreeView &tv = dt.view;
    const bool bybin = ws.stage_sorted == 2;
    const uint32_t *perm = half ? ws.alt_perm : ws.perm, *start = half ? ws.alt_start : ws.start;
    uint32_t *queue = (half ? ws.alt_n_tiles : ws.n_tiles) + 1;
#define BZ_PIPE(BB, NS, MB)                                                                                            \
    eval_bucket_pipe_kernel<DIM, ORDER, BB, NS, MB><<<grid, kPipeThreads, ws.stage_smem, st>>>(                        \
        tv, d_x, d_res, perm, start, ws.bucket_desc, ws.n_keys, ws.bshift, ws.stage_tree_cap, ws.stage_hist_cap,       \
        ws.stage_meta_cap)
#define BZ_QUEUE(BB)                                                                                                   \
    eval_bucket_staged_kernel<DIM, ORDER, BB><<<grid, kStgThreads, ws.stage_smem, st>>>(                               \
        tv, d_x, d_res, perm, start, queue, ws.bucket_desc, ws.n_keys, ws.bshift, ws.stage_tree_cap, ws.stage_hist_cap)
    if (ws.stage_pipe == 3) {
        if (bybin)
            BZ_PIPE(true, 3, 1);
        else
            BZ_PIPE(false, 3, 1);
    } else if (ws.stage_pipe == 2 && ws.stage_minb2) {
        if (bybin)
            BZ_PIPE(true, 2, 2);
        else
            BZ_PIPE(false, 2, 2);
    } else if (ws.stage_pipe == 2) {
        if (bybin)
            BZ_PIPE(true, 2, 1);
        else
            BZ_PIPE(false, 2, 1);
    } else {
        if (bybin)
            BZ_QUEUE(true);
        else
            BZ_QUEUE(false);
    }
#undef BZ_PIPE
#undef BZ_QUEUE
    return (int)cudaGetLastError();
}

#define BZ_STAGE_ORDERS(D, order, CALL)                                                                                \
    switch (order) {                                                                                                   \
    case 6: { constexpr int O = 6; constexpr int DD = D; CALL; } break;                                                \
    case 8: { constexpr int O = 8; constexpr int DD = D; CALL; } break;                                                \
    case 10: { constexpr int O = 10; constexpr int DD = D; CALL; } break;                                              \
    case 12: { constexpr int O = 12; constexpr int DD = D; CALL; } break;                                              \
    case 14: { constexpr int O = 14; constexpr int DD = D; CALL; } break;                                              \
    case 16: { constexpr int O = 16; constexpr int DD = D; CALL; } break;                                              \
    default: rc = (int)cudaErrorNotSupported;                                                                          \
    }

} // namespace

// Called by alloc_bucket_workspace once n_keys / bshift are fixed: describes the buckets on the device and decides
// whether a staged kernel serves this workspace (ws.stage_ok) and which.  Returns cudaError_t; "does not fit" is not
// an error unless BAOBZI_B200_BUCKET_STAGE=2.
int plan_bucket_stage(const DeviceTree &dt, BinWorkspace &ws) {
    ws.stage_ok = false;
    const int dim = dt.view.dim, order = dt.view.order;
    if (dim < 2 || dim > 3)
        return 0;
    bool required = false;
    if (const char *e = std::getenv("BAOBZI_B200_BUCKET_STAGE")) {
        if (std::atoi(e) == 0)
            return 0;
        required = std::atoi(e) == 2;
    }
    const int no_fit = required ? (int)cudaErrorInvalidConfiguration : 0;
    uint32_t blk = 1;
    for (int j = 0; j < dim; ++j)
        blk *= (uint32_t)order;
    uint32_t *maxima = nullptr;
    cudaError_t e = bz_malloc((void **)&ws.bucket_desc, sizeof(uint4) * ws.n_keys);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&maxima, 2 * sizeof(uint32_t));
    uint32_t h[2] = {0, 0};
    if (e == cudaSuccess)
        e = cudaMemcpy(maxima, h, sizeof h, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        bucket_describe_kernel<<<(ws.n_keys + 127) / 128, 128>>>(dt.view, ws.n_keys, ws.bshift, blk / 2, ws.bucket_desc, maxima);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess)
        e = cudaMemcpy(h, maxima, sizeof h, cudaMemcpyDeviceToHost); // also orders the kernel before any other stream's use
    bz_free(maxima);
    if (e != cudaSuccess)
        return (int)e;
    if (h[0] >= (1u << 20)) // mbarrier transaction counts are below 2^20 bytes
        return no_fit;
    ws.stage_tree_cap = (h[0] + 127u) / 128u * 128u;
    // CTA-local order: 2 = by grid cell (few leaves per cell), 1 = by leaf; BAOBZI_B200_BUCKET_SORT overrides
    const uint32_t cells = 1u << ws.bshift;
    ws.stage_sorted = (uint64_t)dt.view.n_leaves <= 8ull * dt.view.n_bins ? 2 : 1;
    if (const char *s = std::getenv("BAOBZI_B200_BUCKET_SORT"))
        ws.stage_sorted = std::atoi(s) == 2 ? 2 : 1;
    if (ws.stage_sorted == 2 && cells > (uint32_t)kStgMaxLeaves)
        ws.stage_sorted = 1;
    if (ws.stage_sorted == 1 && h[1] > (uint32_t)kStgMaxLeaves)
        ws.stage_sorted = cells <= (uint32_t)kStgMaxLeaves ? 2 : 0;
    if (ws.stage_sorted == 0)
        return no_fit;
    ws.stage_hist_cap = ((ws.stage_sorted == 2 ? cells : h[1]) + 7u) / 8u * 8u;
    // stages: two when they fit, else the queue kernel; BAOBZI_B200_BUCKET_PIPE = 0 | 2 | 3 overrides
    int want = 2; // two stages measured 1 % ahead of three on the benchmark tree at every chunk size (1.935e10 / 1.917e10)
    if (const char *s = std::getenv("BAOBZI_B200_BUCKET_PIPE"))
        want = std::atoi(s);
    // sort launches beside the evaluation (binned.cu: launch_eval_bucket_overlap): two stages, 64 registers per thread
    ws.bucket_overlap = false;
    if (const char *s = std::getenv("BAOBZI_B200_BUCKET_OVERLAP"))
        ws.bucket_overlap = std::atoi(s) != 0;
    if (ws.bucket_overlap && dim == 2 && want >= 2)
        want = 2;
    else
        ws.bucket_overlap = false;
    ws.stage_minb2 = ws.bucket_overlap; // which instantiation is configured and launched (the schedule may still fall back)
    ws.stage_pipe = 0;
    // one CTA per SM (512 threads, most of the shared memory): each describes its ceil(n_keys / grid) buckets up front
    const uint32_t pipe_grid = (uint32_t)dt.sm_count < ws.n_keys ? (uint32_t)dt.sm_count : ws.n_keys;
    ws.stage_meta_cap = (ws.n_keys + pipe_grid - 1) / pipe_grid;
    for (int ns = want >= 3 ? 3 : want; ns >= 2 && ws.bshift <= 11; --ns)
        if (pipe_smem(dim, ns, ws.stage_tree_cap, ws.stage_hist_cap, ws.bshift, ws.stage_meta_cap) <= kStgMaxSmem) {
            ws.stage_pipe = ns;
            break;
        }
    ws.stage_smem = ws.stage_pipe ? pipe_smem(dim, ws.stage_pipe, ws.stage_tree_cap, ws.stage_hist_cap, ws.bshift, ws.stage_meta_cap)
                                  : queue_smem(dim, ws.stage_tree_cap, ws.stage_hist_cap);
    if (ws.stage_smem > kStgMaxSmem)
        return no_fit;
    int rc = 0;
    if (dim == 2) {
        BZ_STAGE_ORDERS(2, order, rc = (configure<DD, O>(dt, ws)))
    } else {
        BZ_STAGE_ORDERS(3, order, rc = (configure<DD, O>(dt, ws)))
    }
    if (rc == (int)cudaErrorNotSupported)
        return no_fit;
    if (rc)
        return rc;
    ws.stage_ok = ws.stage_grid > 0;
    // The one-stage kernel lives off the other CTAs of its SM: with fewer than three of them it is behind the gathering
    // kernel (2D 1 M-leaf tree, 95 KB stages, two CTAs per SM: 1.11e10 - 1.18e10 against 1.32e10; 3D 249 K-leaf tree, 105 KB
    // stages, one CTA: 4.8e9 against 6.2e9), with three it is ahead (1.47e10 - 1.53e10).
    if (ws.stage_ok && !ws.stage_pipe && !required && ws.stage_grid < 3 * dt.sm_count)
        ws.stage_ok = false;
    if (!ws.stage_ok || ws.stage_pipe != 2)
        ws.bucket_overlap = false;
    return 0;
}

int launch_eval_buckets_staged(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res,
                               cudaStream_t stream, int half) {
    int rc = 0;
    if (dt.view.dim == 2) {
        BZ_STAGE_ORDERS(2, dt.view.order, rc = (launch<DD, O>(dt, ws, d_x, d_res, stream, half)))
    } else {
        BZ_STAGE_ORDERS(3, dt.view.order, rc = (launch<DD, O>(dt, ws, d_x, d_res, stream, half)))
    }
    return rc;
}

} // namespace bz
