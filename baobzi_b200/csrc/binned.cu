// binned.cu -- the "group points by leaf, then evaluate" path for trees whose per-leaf coefficient block
// is large (3D: 1.7 KB at order 6 ... 32 KB at order 16).
//
// Why: in the direct kernel every lane of a warp sits in a different leaf, so a warp pulls 32 different
// multi-KB coefficient blocks through L2 for 32 results -- 4 KB of L2 traffic per evaluation at 3D
// order 8, two orders of magnitude above the 32 algorithmic bytes, and far beyond what L2 can deliver
// at the FP64 roofline.  Grouping makes every warp work inside ONE leaf: a coefficient is then a
// warp-uniform 16-byte load (one L1 wavefront, broadcast to 32 lanes) feeding 2*P FMAs.
//
// Per chunk of points (sized so that the chunk's points, keys, slots and results stay L2 resident and
// the grouping never touches HBM; DESIGN.md has the arithmetic):
//   K1 bin_lookup   thread per point: bounds -> grid cell -> descent -> leaf rank; key[i] = rank,
//                   count[rank] += 1 (fire-and-forget RED); out-of-domain points get NaN right here
//   K2 bin_scan     one CTA: start[] = exclusive scan of counts rounded up to the warp-tile size,
//                   tile_leaf[] = leaf of every tile, padding slots of perm[] = invalid, cursors reset
//   K3 bin_scatter  thread per point: perm[start[key] + cursor[key]++] = i
//   K4 bin_eval     warp per tile (32*P consecutive slots of one leaf): gather the P points per lane,
//                   evaluate with register-blocked, reference-ordered arithmetic, scatter the results
#include "device_tree.hpp"

#include <cstdio>
#include <cstdlib>

#include "lookup.cuh"

namespace bz {

constexpr uint32_t kInvalidSlot = 0xFFFFFFFFu;
constexpr int kBinThreads = 256;

int bin_eval_default_P(int order);
bool bin_eval_default_smem(int order);
bool bin_eval_variant_exists(int order, int P, bool smem);
bool bin_eval_pipe_exists(int order, int P);
int launch_bin_eval3d(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st);

// ------------------------------------------------------------------------------------------ K1
template <int DIM>
__global__ void __launch_bounds__(kBinThreads)
bin_lookup_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                  uint32_t *__restrict__ key, uint32_t *__restrict__ count, int cshift, uint32_t n) {
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        double p[DIM];
        load_point_keep<DIM>(x, i, p); // keep in L2: K4 gathers these again
        if (outside<DIM>(tv, p)) {
            res[i] = __longlong_as_double(0x7ff8000000000000ll);
            key[i] = kNotLeaf;
            continue;
        }
        NodeHead head;
        const uint32_t node = locate<DIM>(tv, p, head);
        const uint32_t rank = load_tail(tv.nodes, node).leaf_rank;
        key[i] = rank;
        atomicAdd(count + ((size_t)rank << cshift), 1u);
    }
}

// ------------------------------------------------------------------------------------------ K1/K3, privatised
// Same job as K1/K3 with one shared-memory atomic per point instead of two global ones: every CTA owns a
// contiguous range of the chunk, histograms it in shared memory (the atomic's return value is the point's rank
// inside (CTA, leaf)), then reserves its share of every leaf with ONE global atomic per non-empty (CTA, leaf).
// K3 becomes atomic-free: slot = start[leaf] + base[CTA][leaf] + rank.  Used when the histogram fits in smem.
template <int DIM>
__global__ void __launch_bounds__(kBinThreads)
bin_lookup2_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                   uint32_t *__restrict__ key, uint32_t *__restrict__ rank, uint32_t *__restrict__ count, int cshift,
                   uint32_t *__restrict__ cta_base, uint32_t n) {
    extern __shared__ uint32_t hist[];
    const uint32_t nl = tv.n_leaves;
    for (uint32_t l = threadIdx.x; l < nl; l += blockDim.x)
        hist[l] = 0;
    __syncthreads();
    const uint32_t lo = (uint32_t)(((uint64_t)blockIdx.x * n) / gridDim.x);
    const uint32_t hi = (uint32_t)(((uint64_t)(blockIdx.x + 1) * n) / gridDim.x);
    // two points per thread per trip: the lookup is a chain of dependent L2 loads, so independent chains are
    // what keeps the memory system busy
    for (uint32_t i0 = lo + threadIdx.x; i0 < hi; i0 += 2 * blockDim.x) {
        const uint32_t i1 = i0 + blockDim.x;
        const bool has1 = i1 < hi;
        double p0[DIM], p1[DIM];
        load_point_keep<DIM>(x, i0, p0);
        load_point_keep<DIM>(x, has1 ? i1 : i0, p1);
        const bool out0 = outside<DIM>(tv, p0), out1 = outside<DIM>(tv, p1) || !has1;
        uint32_t leaf0 = kNotLeaf, leaf1 = kNotLeaf;
        if (!out0)
            leaf0 = locate_leaf<DIM>(tv, p0);
        if (!out1)
            leaf1 = locate_leaf<DIM>(tv, p1);
        if (out0)
            res[i0] = __longlong_as_double(0x7ff8000000000000ll);
        else
            rank[i0] = atomicAdd(hist + leaf0, 1u);
        key[i0] = leaf0;
        if (has1) {
            if (outside<DIM>(tv, p1))
                res[i1] = __longlong_as_double(0x7ff8000000000000ll);
            else
                rank[i1] = atomicAdd(hist + leaf1, 1u);
            key[i1] = leaf1;
        }
    }
    __syncthreads();
    for (uint32_t l = threadIdx.x; l < nl; l += blockDim.x) {
        const uint32_t h = hist[l];
        if (h)
            cta_base[(size_t)blockIdx.x * nl + l] = atomicAdd(count + ((size_t)l << cshift), h);
    }
}

__global__ void __launch_bounds__(kBinThreads)
bin_scatter2_kernel(const uint32_t *__restrict__ key, const uint32_t *__restrict__ rank,
                    const uint32_t *__restrict__ start, const uint32_t *__restrict__ cta_base, uint32_t n_leaves,
                    uint32_t *__restrict__ perm, uint32_t n) {
    const uint32_t lo = (uint32_t)(((uint64_t)blockIdx.x * n) / gridDim.x);
    const uint32_t hi = (uint32_t)(((uint64_t)(blockIdx.x + 1) * n) / gridDim.x);
    const uint32_t *base = cta_base + (size_t)blockIdx.x * n_leaves;
    for (uint32_t i0 = lo + threadIdx.x; i0 < hi; i0 += 4 * blockDim.x) {
        uint32_t k[4], r[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t i = i0 + u * blockDim.x;
            k[u] = i < hi ? __ldcs(key + i) : kNotLeaf;
            r[u] = i < hi ? __ldcs(rank + i) : 0u;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (k[u] != kNotLeaf)
                perm[__ldg(start + k[u]) + __ldg(base + k[u]) + r[u]] = i0 + u * blockDim.x;
    }
}

// ------------------------------------------------------------------------------------------ K2
// Single CTA: exclusive scan of the per-leaf counts rounded up to the warp-tile size.  count[] in: histogram;
// out: zero (ready for the next chunk's K1).  cursor[] out: zero (for K3).  cnt[] out: dense copy for K2b.
__global__ void __launch_bounds__(1024)
bin_scan_kernel(uint32_t TILE, uint32_t n_leaves, uint32_t *__restrict__ count, uint32_t *__restrict__ cursor, int cshift,
                uint32_t *__restrict__ cnt_out, uint32_t *__restrict__ start, uint32_t *__restrict__ n_tiles) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry_s;
    if (threadIdx.x == 0)
        carry_s = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t base = 0; base < n_leaves; base += blockDim.x) {
        const uint32_t l = base + threadIdx.x;
        const uint32_t cnt = l < n_leaves ? count[(size_t)l << cshift] : 0u;
        const uint32_t padded = (cnt + TILE - 1) / TILE * TILE;
        uint32_t v = padded; // inclusive scan of `padded` across the CTA
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o)
                v += t;
        }
        if (lane == 31)
            warp_sums[warp] = v;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o)
                    w += t;
            }
            warp_sums[lane] = w;
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        if (l < n_leaves) {
            start[l] = carry + (warp ? warp_sums[warp - 1] : 0u) + v - padded;
            cnt_out[l] = cnt;
            count[(size_t)l << cshift] = 0;
            cursor[(size_t)l << cshift] = 0;
        }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1)
            carry_s = carry + warp_sums[31];
        __syncthreads();
    }
    if (threadIdx.x == 0)
        *n_tiles = carry_s / TILE;
}

// K2b: one warp per leaf marks the padding slots of perm[] invalid and records the leaf of each of its tiles.
__global__ void __launch_bounds__(kBinThreads)
bin_fill_kernel(const TreeView tv, uint32_t TILE, uint32_t n_leaves, const uint32_t *__restrict__ cnt,
                const uint32_t *__restrict__ start, uint32_t *__restrict__ tile_leaf, uint2 *__restrict__ tile_info,
                uint32_t *__restrict__ perm) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t l = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (l >= n_leaves)
        return;
    const uint32_t c = cnt[l], s0 = start[l];
    const uint32_t padded = (c + TILE - 1) / TILE * TILE;
    for (uint32_t s = s0 + c + lane; s < s0 + padded; s += 32)
        perm[s] = kInvalidSlot;
    const uint32_t node = __ldg(tv.leaf_node + l);
    const uint2 info = make_uint2(node, load_head(tv.nodes, node).coeff16);
    for (uint32_t t = s0 / TILE + lane; t < (s0 + padded) / TILE; t += 32) {
        tile_leaf[t] = l;
        tile_info[t] = info;
    }
}

// ------------------------------------------------------------------------------------------ K3
__global__ void __launch_bounds__(kBinThreads)
bin_scatter_kernel(const uint32_t *__restrict__ key, const uint32_t *__restrict__ start, uint32_t *__restrict__ cursor,
                   int cshift, uint32_t *__restrict__ perm, uint32_t n) {
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint32_t k = key[i];
        if (k == kNotLeaf)
            continue;
        const uint32_t pos = __ldg(start + k) + atomicAdd(cursor + ((size_t)k << cshift), 1u);
        perm[pos] = i;
    }
}

// ------------------------------------------------------------------------------------------ host side
bool binned_supported(int dim, int order) { return dim == 3 && order >= 6 && order <= 16 && order % 2 == 0; }

int alloc_workspace(const DeviceTree &dt, int64_t chunk, BinWorkspace &ws) {
    if (chunk < 1)
        chunk = 1;
    if (chunk > 0x7FFFFFFFll)
        chunk = 0x7FFFFFFFll;
    int P = bin_eval_default_P(dt.view.order);
    bool smem = bin_eval_default_smem(dt.view.order);
    if (const char *e = std::getenv("BAOBZI_B200_P")) // tuning overrides (benchmarks)
        if (bin_eval_variant_exists(dt.view.order, std::atoi(e), smem))
            P = std::atoi(e);
    if (const char *e = std::getenv("BAOBZI_B200_SMEM"))
        if (bin_eval_variant_exists(dt.view.order, P, std::atoi(e) != 0))
            smem = std::atoi(e) != 0;
    ws.P = P;
    ws.smem_coeffs = smem;
    ws.pipelined = bin_eval_pipe_exists(dt.view.order, P);
    if (const char *e = std::getenv("BAOBZI_B200_PIPE"))
        ws.pipelined = ws.pipelined && std::atoi(e) != 0;
    const int64_t tile = 32 * P;
    const int64_t nl = dt.view.n_leaves;
    ws.perm_capacity = chunk + nl * tile;
    ws.tile_capacity = ws.perm_capacity / tile + 1;
    // spread the per-leaf counters: up to 32 words (one 128-byte line) apart while the arrays stay <= 4 MB
    ws.cshift = 0;
    while (ws.cshift < 5 && ((nl + 1) << (ws.cshift + 1)) <= (1 << 20))
        ws.cshift++;
    const size_t cwords = (size_t)(nl + 1) << ws.cshift;
    ws.privatized = (size_t)nl * sizeof(uint32_t) <= 160 * 1024;
    if (const char *e = std::getenv("BAOBZI_B200_PRIVATE"))
        ws.privatized = ws.privatized && std::atoi(e) != 0;
    ws.sort_ctas = dt.sm_count * 4;
    if (const char *e = std::getenv("BAOBZI_B200_SORT_CTAS"))
        if (std::atoi(e) > 0)
            ws.sort_ctas = dt.sm_count * std::atoi(e);
    cudaError_t e = cudaMalloc((void **)&ws.key, sizeof(uint32_t) * chunk);
    if (e == cudaSuccess && ws.privatized)
        e = cudaMalloc((void **)&ws.rank, sizeof(uint32_t) * chunk);
    if (e == cudaSuccess && ws.privatized)
        e = cudaMalloc((void **)&ws.cta_base, sizeof(uint32_t) * (size_t)ws.sort_ctas * (nl + 1));
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&ws.perm, sizeof(uint32_t) * ws.perm_capacity);
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&ws.count, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&ws.cursor, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&ws.cnt, sizeof(uint32_t) * (nl + 1));
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&ws.start, sizeof(uint32_t) * (nl + 1));
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&ws.tile_leaf, sizeof(uint32_t) * ws.tile_capacity);
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&ws.n_tiles, sizeof(uint32_t));
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&ws.tile_info, sizeof(uint2) * ws.tile_capacity);
    if (e == cudaSuccess)
        e = cudaMemset(ws.count, 0, sizeof(uint32_t) * cwords);
    if (e != cudaSuccess) {
        free_workspace(ws);
        return (int)e;
    }
    ws.chunk_capacity = chunk;
    return 0;
}

void free_workspace(BinWorkspace &ws) {
    cudaFree(ws.key);
    cudaFree(ws.perm);
    cudaFree(ws.count);
    cudaFree(ws.cursor);
    cudaFree(ws.cnt);
    cudaFree(ws.start);
    cudaFree(ws.tile_leaf);
    cudaFree(ws.n_tiles);
    cudaFree(ws.rank);
    cudaFree(ws.cta_base);
    cudaFree(ws.tile_info);
    ws = BinWorkspace();
}

static int launch_chunk_3d(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, uint32_t n,
                           cudaStream_t st) {
    const uint32_t TILE = 32u * (uint32_t)ws.P;
    const TreeView &tv = dt.view;
    const unsigned cap = (unsigned)dt.sm_count * 16;
    unsigned g1 = (n + kBinThreads - 1) / kBinThreads;
    g1 = g1 < cap ? g1 : cap;
    StageProfile *pf = ws.prof;
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    if (pf) {
        for (auto &e : ev)
            e = pf->get();
        cudaEventRecord(ev[0], st);
    }
    const size_t hist_bytes = (size_t)tv.n_leaves * sizeof(uint32_t);
    if (ws.privatized) {
        static bool configured = false;
        if (!configured && hist_bytes > 48 * 1024) {
            cudaFuncSetAttribute(bin_lookup2_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
            configured = true;
        }
        bin_lookup2_kernel<3><<<ws.sort_ctas, kBinThreads, hist_bytes, st>>>(tv, d_x, d_res, ws.key, ws.rank, ws.count,
                                                                           ws.cshift, ws.cta_base, n);
    } else {
        bin_lookup_kernel<3><<<g1, kBinThreads, 0, st>>>(tv, d_x, d_res, ws.key, ws.count, ws.cshift, n);
    }
    if (pf)
        cudaEventRecord(ev[1], st);
    bin_scan_kernel<<<1, 1024, 0, st>>>(TILE, tv.n_leaves, ws.count, ws.cursor, ws.cshift, ws.cnt, ws.start, ws.n_tiles);
    bin_fill_kernel<<<(tv.n_leaves * 32 + kBinThreads - 1) / kBinThreads, kBinThreads, 0, st>>>(
        tv, TILE, tv.n_leaves, ws.cnt, ws.start, ws.tile_leaf, ws.tile_info, ws.perm);
    if (pf)
        cudaEventRecord(ev[2], st);
    if (ws.privatized)
        bin_scatter2_kernel<<<ws.sort_ctas, kBinThreads, 0, st>>>(ws.key, ws.rank, ws.start, ws.cta_base, tv.n_leaves, ws.perm, n);
    else
        bin_scatter_kernel<<<g1, kBinThreads, 0, st>>>(ws.key, ws.start, ws.cursor, ws.cshift, ws.perm, n);
    if (pf)
        cudaEventRecord(ev[3], st);
    int rc = launch_bin_eval3d(dt, ws, d_x, d_res, st);
    if (pf) {
        cudaEventRecord(ev[4], st);
        pf->spans.push_back({kStageLookup, ev[0], ev[1], (int64_t)n});
        pf->spans.push_back({kStageScan, ev[1], ev[2], (int64_t)n});
        pf->spans.push_back({kStageScatter, ev[2], ev[3], (int64_t)n});
        pf->spans.push_back({kStageEval, ev[3], ev[4], (int64_t)n});
    }
    return rc ? rc : (int)cudaGetLastError();
}

int launch_eval_binned(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, int64_t n,
                       cudaStream_t stream, int64_t *launches) {
    if (n <= 0)
        return 0;
    if (dt.view.dim != 3 || n > ws.chunk_capacity)
        return (int)cudaErrorInvalidValue;
    int rc = launch_chunk_3d(dt, ws, d_x, d_res, (uint32_t)n, stream);
    if (launches)
        *launches += 5;
    return rc;
}

} // namespace bz
