// sort_common.cuh -- pieces of the leaf sort shared by the sort kernels (binned.cu) and by the evaluation kernel that
// carries the sort of the following chunks as a side job (bin_eval.cu: bin_eval3d_piggy_kernel).
#pragma once

#include "baobzi/lookup.cuh"
#include "device_tree.hpp"

namespace bz {

constexpr uint32_t kInvalidSlot = 0xFFFFFFFFu;
// Key format between the two launches of the split sort (and the side jobs of the piggyback schedule): leaf rank in
// the low 15 bits (kKeyNoLeaf = outside the domain), the point's rank inside (CTA, leaf) above them.  Limits, checked
// where the workspace is built: fewer than 32 767 leaves, at most 2^17 points per CTA and chunk.
constexpr int kKeyLeafBits = 15;
constexpr uint32_t kKeyNoLeaf = (1u << kKeyLeafBits) - 1u;
constexpr int kSortBatch = 6; // histogram counters (loads / atomics) in flight per thread in the scan phases

// BYBIN (the bucketed path of large 1D / 2D trees): the key is the grid cell number >> bshift instead of the leaf,
// so the sort touches no tree data; TILE is 1 and no tile descriptors are written (tile_info == nullptr).
constexpr uint32_t kBigLeafCap = 128;

// Shared-memory counter increment for a warp whose lanes may hold the same key (k == kNotLeaf: no increment).
// Returns the value the lane's own atomicAdd(tbl + k, 1) would have returned, up to the order inside the warp.
// `heavy` (warp uniform) says most lanes share keys -- clustered or ordered input, where 32 same-address ATOMS
// per warp and 16 warps on one bank made the sort 5x slower than on uniform points: the lanes of each key then
// elect a leader that adds the group's size once.  All 32 lanes must call this together.
__device__ __forceinline__ uint32_t grouped_add(uint32_t *tbl, uint32_t k, bool heavy, uint32_t lane) {
    if (!heavy)
        return k != kNotLeaf ? atomicAdd(tbl + k, 1u) : 0u;
    const uint32_t grp = __match_any_sync(0xffffffffu, k);
    const int leader = __ffs(grp) - 1;
    uint32_t base = 0;
    if ((int)lane == leader && k != kNotLeaf)
        base = atomicAdd(tbl + k, (uint32_t)__popc(grp));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + (uint32_t)__popc(grp & ((1u << lane) - 1u));
}

// warp uniform: do at least half of the lanes hold the key of their right neighbour?
__device__ __forceinline__ bool keys_repeat(uint32_t k, uint32_t lane) {
    const uint32_t nb = __shfl_sync(0xffffffffu, k, (lane + 1) & 31);
    return __popc(__ballot_sync(0xffffffffu, nb == k && k != kNotLeaf)) >= 16;
}


// CTA-wide, after the histogram count[] of the chunk is complete and tbl[l] holds this CTA's reserved base inside
// leaf l: padded exclusive scan of the histogram (thread t owns leaves [t*per, (t+1)*per)), tbl[l] += start[l];
// the CTA's share of the leaves (l % gridDim == blockIdx) gets its padding slots marked, its tile descriptors
// written and its counter in count_next cleared (the buffer a later chunk histograms into); leaves holding a large
// share of the chunk get their descriptors from the whole grid.  Ends with a __syncthreads.
// warp_sums[32], big[kBigLeafCap], n_big: shared scratch; *n_big must be 0 on entry (and visible).
template <bool BYBIN>
__device__ __forceinline__ void sort_scan_describe(const TreeView &tv, uint32_t *tbl, const uint32_t *__restrict__ count,
                                                   uint32_t *__restrict__ count_next, int cshift, uint32_t TILE,
                                                   uint32_t *__restrict__ perm, uint32_t *__restrict__ tile_leaf,
                                                   uint2 *__restrict__ tile_info, uint32_t *__restrict__ n_tiles, uint32_t nl,
                                                   uint32_t *__restrict__ key_start, uint32_t *warp_sums, uint4 *big,
                                                   uint32_t *n_big_p) {
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    uint32_t &n_big = *n_big_p;
    // ---- phase B: padded exclusive scan of the global histogram, thread t owns leaves [t*per, (t+1)*per)
    const uint32_t per = (nl + nthr - 1) / nthr;
    const uint32_t l0 = min(tid * per, nl), l1 = min(l0 + per, nl);
    uint32_t mine = 0;
    for (uint32_t lb = l0; lb < l1; lb += kSortBatch) { // kSortBatch counters requested together, not one by one
        uint32_t c[kSortBatch];
#pragma unroll
        for (int j = 0; j < kSortBatch; ++j)
            c[j] = lb + j < l1 ? __ldcg(count + ((size_t)(lb + j) << cshift)) : 0u;
#pragma unroll
        for (int j = 0; j < kSortBatch; ++j)
            mine += (c[j] + TILE - 1) / TILE * TILE;
    }
    const uint32_t lane = tid & 31, warp = tid >> 5;
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o)
            incl += t;
    }
    if (lane == 31)
        warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < nthr / 32 ? warp_sums[lane] : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o)
                w += t;
        }
        if (lane < nthr / 32)
            warp_sums[lane] = w;
    }
    __syncthreads();
    uint32_t start = (warp ? warp_sums[warp - 1] : 0u) + incl - mine;
    // A leaf holding a large share of the chunk (clustered input: up to every tile of it) gets its tile descriptors
    // from the whole grid, not from one thread: 65 K tiles of one leaf took 0.6 ms per chunk that way.  "Large" is
    // more than 1/kBigLeafCap of all tiles, so the list below cannot overflow and is the same in every CTA.
    const uint32_t big_tiles = max(256u, warp_sums[nthr / 32 - 1] / TILE / kBigLeafCap);
    const uint32_t start0 = start;
    bool any_big = false;
    for (uint32_t lb = l0; lb < l1; lb += kSortBatch) {
      uint32_t cb[kSortBatch];
#pragma unroll
      for (int j = 0; j < kSortBatch; ++j)
          cb[j] = lb + j < l1 ? __ldcg(count + ((size_t)(lb + j) << cshift)) : 0u;
#pragma unroll
      for (int j = 0; j < kSortBatch; ++j) {
        const uint32_t l = lb + j;
        if (l >= l1)
            break;
        const uint32_t c = cb[j];
        const uint32_t padded = (c + TILE - 1) / TILE * TILE;
        const bool is_big = !BYBIN && padded / TILE > big_tiles;
        any_big = any_big || is_big;
        if (l % gridDim.x == blockIdx.x) { // this CTA's leaves: padding slots, tile descriptors, counter reset
            if (!BYBIN) {
                for (uint32_t s = start + c; s < start + padded; ++s)
                    perm[s] = kInvalidSlot;
                if (!is_big) {
                    const uint32_t node = __ldg(tv.leaf_node + l);
                    const uint2 info = make_uint2(node, load_head(tv.nodes, node).coeff16);
                    for (uint32_t t = start / TILE; t < (start + padded) / TILE; ++t) {
                        tile_leaf[t] = l;
                        tile_info[t] = info;
                    }
                }
            }
            count_next[(size_t)l << cshift] = 0; // the buffer the NEXT chunk histograms into
        }
        if (BYBIN && blockIdx.x == 0)
            key_start[l] = start; // first slot of every bucket, for the evaluation kernel
        tbl[l] += start;
        start += padded;
      }
    }
    if (any_big) { // rare: list this thread's big leaves (kept out of the loop above: an atomic there serialises its loads)
        uint32_t s2 = start0;
        for (uint32_t l = l0; l < l1; ++l) {
            const uint32_t c = __ldcg(count + ((size_t)l << cshift));
            const uint32_t padded = (c + TILE - 1) / TILE * TILE;
            if (padded / TILE > big_tiles) // every CTA lists the same (fewer than kBigLeafCap) leaves
                big[atomicAdd(&n_big, 1u)] = make_uint4(l, s2 / TILE, padded / TILE, __ldg(tv.leaf_node + l));
            s2 += padded;
        }
    }
    if (blockIdx.x == 0 && tid == nthr - 1) {
        *n_tiles = (warp_sums[nthr / 32 - 1]) / TILE;
        if (BYBIN) {
            key_start[nl] = warp_sums[nthr / 32 - 1];
            n_tiles[1] = 0; // work queue of the evaluation kernel
        }
    }
    __syncthreads();
    if (!BYBIN) {
        for (uint32_t e = 0; e < n_big; ++e) {
            const uint4 b = big[e]; // {leaf, first tile, tiles, node}
            const uint2 info = make_uint2(b.w, load_head(tv.nodes, b.w).coeff16);
            for (uint32_t t = blockIdx.x * nthr + tid; t < b.z; t += gridDim.x * nthr) {
                tile_leaf[b.y + t] = b.x;
                tile_info[b.y + t] = info;
            }
        }
    }

}

} // namespace bz
