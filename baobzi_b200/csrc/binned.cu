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
// K1-K3 as written above are the fallback for trees whose histogram exceeds shared memory.  Otherwise they are ONE
// kernel, bin_sort_fused_kernel, launched either once per chunk (cooperative, a grid barrier between lookup and
// scan + scatter) or, for trees from kSortSplitMinLeaves(2D) leaves whenever a call spans several chunks, as two ordinary
// launches per chunk on two streams: the lookup of chunk c+1 runs beside the scan + scatter of chunk c
// (launch_eval_split; DESIGN.md 4.2 has the measurements behind every threshold).
#include "device_tree.hpp"

#include <cstdio>
#include <cstdlib>
#include <utility>

#include "baobzi/lookup.cuh"
#include "sort_common.cuh"

#include <cooperative_groups.h>

namespace bz {

constexpr int kBinThreads = 256;

int bin_eval_select(int order, int *P);
int bin_eval2d_P(int order);
int launch_bin_eval2d(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st);
bool bin_eval_footprint(int variant, int *cta_regs, size_t *cta_smem);
int launch_bin_eval3d(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st);
bool bin_eval_is_default(int variant);
int bin_eval_piggy_P(int order);
size_t bin_eval_piggy_smem(int order, uint32_t n_leaves, uint32_t sbin_n);
int launch_bin_eval3d_piggy(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st,
                            const PiggyArgs &pg);

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

// ------------------------------------------------------------------------------------------ fused sort
// K1 + K2 + K2b + K3 as ONE cooperative kernel (used whenever the per-leaf histogram fits in shared memory).
// Every CTA owns a contiguous range of the chunk.
//   phase A  lookup: leaf rank per point, histogram of the range in shared memory, key[i] = leaf; then one global
//            atomic per non-empty (CTA, leaf) reserves the CTA's share of the leaf: base = atomicAdd(count[leaf], h)
//   -- grid barrier --
//   phase B  every CTA scans the (now complete) histogram itself -- a few thousand counters out of L2, cheaper than
//            a single-CTA scan kernel plus a launch; the leaves are dealt round-robin to the CTAs for the padding
//            slots and the tile descriptors.  tbl[leaf] = start[leaf] + base[leaf], and the scatter is
//            perm[tbl[leaf]++] = i with a shared-memory atomic: no rank array, no global atomics, no table gathers.
// HBM traffic: the points once (8*DIM B) + perm (4 B); key[] (4 B written, 4 B read) lives in L2.
constexpr int kSortThreads = 512;
constexpr int kSortThreadsA = 768; // most threads the lookup launch of the split sort may be given
constexpr size_t kFusedSortMaxSmem = 160 * 1024;
constexpr int kSortSplitDefault = 1;
constexpr int64_t kSortSplitMinLeaves = 2048;
constexpr int64_t kSortSplitMinLeaves2D = 128; // split sort (phase A of the next chunk beside the scatter of this one): CTAs per SM and kernel, 0 = off
constexpr int kSortPts = 4; // independent lookups in flight per thread (the lookup is latency bound, not issue bound)

template <int DIM, bool BYBIN, int PHASE = 0, int APTS = kSortPts>
__global__ void __launch_bounds__(PHASE == 1 ? kSortThreadsA : kSortThreads)
bin_sort_fused_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                      uint32_t *__restrict__ key, uint32_t *__restrict__ count, uint32_t *__restrict__ count_next,
                      int cshift, uint32_t TILE, uint32_t *__restrict__ perm, uint32_t *__restrict__ tile_leaf,
                      uint2 *__restrict__ tile_info, uint32_t *__restrict__ n_tiles, uint32_t n, uint32_t skey_cap,
                      uint32_t n_keys, int bshift, uint32_t *__restrict__ key_start, uint32_t coal_cap,
                      uint32_t sbin_n, uint32_t /*phase: the template argument*/, uint32_t *__restrict__ base_tbl) {
    constexpr uint32_t phase = PHASE;
    // phase 0: everything in one cooperative launch (grid barrier between A and B).  Split mode: phase 1 = A alone
    // (keys through key[], the CTA's reserved bases saved in base_tbl[CTA][leaf]), phase 2 = B + scatter alone (ordinary
    // launches of the same grid).  A streams points from HBM, the scatter is bound by the rate at which the chip takes
    // scattered 4-byte stores: launched on two streams, A of chunk c+1 runs beside the scatter of chunk c.
    constexpr bool do_a = phase != 2;
    extern __shared__ uint32_t tbl[]; // [n_keys] histogram -> base -> next free slot
    __shared__ uint32_t warp_sums[32];
    __shared__ uint4 big[kBigLeafCap];
    __shared__ uint32_t n_big;
    if (threadIdx.x == 0)
        n_big = 0;
    const uint32_t nl = n_keys;
    // keys of the CTA's own range as 16-bit leaf ranks in shared memory when they fit (skey_cap entries follow
    // tbl[]): the scatter then never goes back to global memory for them
    uint16_t *skey = reinterpret_cast<uint16_t *>(tbl + nl);
    // sbin_n > 0: the bin_leaf table as 16-bit ranks in shared memory behind the keys.  The gather of phase A then costs
    // a few bank-conflict wavefronts per warp instead of one L1 wavefront per distinct 128-byte line (~28 of them for 32
    // random grid cells), and the LSU wavefront rate is what bounds that phase.
    uint16_t *sbin = skey + skey_cap;
    // coal_cap > 0: the scatter goes through a CTA-local counting sort in shared memory (cnt[], loc[], sorted[] follow
    // tbl[]), and the CTA's run of every leaf leaves as ONE merged store per 8 slots instead of one 4-byte store per
    // point (the scattered 4-byte stores were 40 % of this kernel: L2 takes them one partial sector at a time)
    uint32_t *cnt = tbl + nl, *loc = cnt + nl;
    uint16_t *sorted = reinterpret_cast<uint16_t *>(loc + nl + 1);
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    for (uint32_t l = tid; l < nl; l += nthr)
        tbl[l] = do_a ? 0u : base_tbl[(size_t)blockIdx.x * nl + l];
    for (uint32_t l = tid; l < sbin_n; l += nthr)
        sbin[l] = (uint16_t)__ldg(tv.bin_leaf + l); // kNotLeaf -> 0xFFFF (n_leaves < 0xFFFF when sbin_n > 0)
    __syncthreads();
    // split mode: keys travel through key[] as 16-bit ranks and come back as 128-bit loads, so ranges start at multiples of 8
    const uint32_t lo_raw = (uint32_t)(((uint64_t)blockIdx.x * n) / gridDim.x);
    const uint32_t hi_raw = (uint32_t)(((uint64_t)(blockIdx.x + 1) * n) / gridDim.x);
    const uint32_t lo = PHASE ? lo_raw & ~7u : lo_raw;
    const uint32_t hi = PHASE && blockIdx.x + 1 != gridDim.x ? hi_raw & ~7u : hi_raw;
    const bool use_skey = hi - lo <= skey_cap; // CTA uniform
    const bool coalesce = PHASE == 0 && coal_cap > 0 && hi - lo <= coal_cap; // (PHASE 2 has its own use of coal_cap below)

    // ---- phase A
    const uint32_t lane_id = tid & 31;
    constexpr int PA = PHASE == 1 ? APTS : kSortPts; // points in flight per thread in the lookup loop
    for (uint32_t b0 = lo; do_a && b0 < hi; b0 += PA * nthr) { // CTA-uniform trip count: the warps stay converged
        const uint32_t i0 = b0 + tid;
        double p[PA][DIM];
        bool have[PA];
#pragma unroll
        for (int u = 0; u < PA; ++u) {
            const uint32_t i = i0 + u * nthr;
            have[u] = i < hi;
            load_point_keep<DIM>(x, have[u] ? i : lo, p[u]); // default cache policy: K4 gathers these again
        }
        uint32_t leaf[PA], lin[PA];
#pragma unroll
        for (int u = 0; u < PA; ++u) { // all table lookups in flight before any (rare) descent
            lin[u] = linear_bin<DIM>(tv, p[u]);
            if (BYBIN) {
                leaf[u] = lin[u] >> bshift;
            } else if (sbin_n) {
                const uint32_t v = sbin[lin[u]];
                leaf[u] = v == 0xFFFFu ? kNotLeaf : v;
            } else {
                leaf[u] = __ldg(tv.bin_leaf + lin[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < PA; ++u) {
            if (!have[u] || outside<DIM>(tv, p[u]))
                leaf[u] = kNotLeaf;
            else if (!BYBIN && leaf[u] == kNotLeaf)
                leaf[u] = walk_to_leaf<DIM>(tv, lin[u], p[u]);
        }
        const bool heavy = keys_repeat(leaf[0], lane_id);
#pragma unroll
        for (int u = 0; u < PA; ++u) {
            const uint32_t i = i0 + u * nthr;
            // rank of the point inside (this CTA, its leaf): what the histogram update returns (lanes past the end hold kNotLeaf)
            const uint32_t rank = grouped_add(tbl, leaf[u], heavy, lane_id);
            if (have[u]) {
                if (PHASE == 1) // split mode: leaf and rank travel together, the scatter launch needs no atomics
                    key[i] = leaf[u] == kNotLeaf ? kKeyNoLeaf : (leaf[u] | (rank << kKeyLeafBits));
                else if (use_skey && !coalesce)
                    skey[i - lo] = (uint16_t)leaf[u]; // kNotLeaf -> 0xFFFF (n_leaves < 0xFFFF when skey_cap > 0)
                else
                    key[i] = leaf[u];
                if (leaf[u] == kNotLeaf)
                    res[i] = __longlong_as_double(0x7ff8000000000000ll);
            }
        }
    }
    __syncthreads();
    // kSortBatch atomics in flight per thread: one at a time (a round trip to L2 each, nl / nthr of them per thread)
    // was 18 % of the kernel on a launch-sized 2D chunk
    for (uint32_t lb = tid; do_a && lb < nl; lb += kSortBatch * nthr) {
        uint32_t h[kSortBatch], r[kSortBatch];
#pragma unroll
        for (int j = 0; j < kSortBatch; ++j) {
            const uint32_t l = lb + j * nthr;
            h[j] = l < nl ? tbl[l] : 0u;
        }
#pragma unroll
        for (int j = 0; j < kSortBatch; ++j)
            r[j] = h[j] ? atomicAdd(count + ((size_t)(lb + j * nthr) << cshift), h[j]) : 0u;
#pragma unroll
        for (int j = 0; j < kSortBatch; ++j) {
            const uint32_t l = lb + j * nthr;
            if (l < nl) {
                if (coalesce)
                    cnt[l] = h[j];
                tbl[l] = r[j];
                if (phase == 1) {
                    base_tbl[(size_t)blockIdx.x * nl + l] = r[j];
                    base_tbl[(size_t)(gridDim.x + blockIdx.x) * nl + l] = h[j]; // this CTA's own count: the scatter launch sorts locally with it
                }
            }
        }
    }
    if (phase == 1)
        return;
    if (phase == 0) {
        __threadfence();
        cooperative_groups::this_grid().sync();
    }

    // ---- phase B: padded exclusive scan of the global histogram, padding slots, tile descriptors (sort_common.cuh)
    sort_scan_describe<BYBIN>(tv, tbl, count, count_next, cshift, TILE, perm, tile_leaf, tile_info, n_tiles, nl, key_start,
                              warp_sums, big, &n_big);
    const uint32_t lane = tid & 31, warp = tid >> 5;
    const uint32_t per = (nl + nthr - 1) / nthr;
    const uint32_t l0 = min(tid * per, nl), l1 = min(l0 + per, nl);
    // ---- scatter
    if (coalesce) {
        uint32_t mine2 = 0;
        for (uint32_t l = l0; l < l1; ++l)
            mine2 += cnt[l];
        uint32_t incl2 = mine2;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl2, o);
            if (lane >= o)
                incl2 += t;
        }
        if (lane == 31)
            warp_sums[warp] = incl2;
        __syncthreads();
        uint32_t off = incl2 - mine2;
        for (uint32_t w = 0; w < warp; ++w)
            off += warp_sums[w];
        for (uint32_t l = l0; l < l1; ++l) {
            loc[l] = off;
            off += cnt[l];
        }
        if (l1 == nl)
            loc[nl] = off; // every thread at or past the end holds the total
        __syncthreads();
        for (uint32_t i0 = lo + tid; i0 < hi; i0 += kSortPts * nthr) {
            uint32_t k[kSortPts];
#pragma unroll
            for (int u = 0; u < kSortPts; ++u) {
                const uint32_t i = i0 + u * nthr;
                k[u] = i < hi ? __ldcs(key + i) : kNotLeaf;
            }
#pragma unroll
            for (int u = 0; u < kSortPts; ++u)
                if (k[u] != kNotLeaf)
                    sorted[loc[k[u]] + atomicSub(cnt + k[u], 1u) - 1u] = (uint16_t)(i0 + u * nthr - lo);
        }
        __syncthreads();
        const uint32_t sub = tid & 7;
        for (uint32_t l = tid >> 3; l < nl; l += nthr >> 3) { // 8 lanes per run: their stores merge into one sector
            const uint32_t b = loc[l], len = loc[l + 1] - b, g = tbl[l];
            for (uint32_t e = sub; e < len; e += 8)
                perm[g + e] = lo + sorted[b + e];
        }
        return;
    }
    if constexpr (PHASE == 2) {
        // Keys come from global memory (written by phase A's launch, mostly evicted from L2 by the points it streamed
        // afterwards), eight per thread and iteration as two 128-bit loads requested two iterations ahead.  A key carries
        // the leaf AND the point's rank inside (CTA, leaf) -- the value phase A's histogram atomic returned -- so the slot is
        // tbl[leaf] + rank: a plain shared-memory load.  (Round 1 took a second shared atomic per point here; at 2 cycles
        // per lane the two atomics per point WERE the sort's duration: 16.8 M of them per 8.4 M-point chunk = 115 us.)
        constexpr uint32_t KPT = 8;
        const uint32_t per_it = KPT * nthr;
        // Coalescing scatter (coal_cap > 0: the CTA's range fits the shared-memory staging).  What bounds the direct
        // scatter below is not arithmetic or atomics but the rate at which the chip takes scattered 4-byte stores
        // (ncu source view: every iteration waits for the 4096 stores of the previous one to drain; 8.4 M of them =
        // 46 us, scripts/micro/scatter_rate.cu).  With the rank in the key a CTA-local counting sort costs no atomic at
        // all: local position = (exclusive scan of the CTA's own counts)[leaf] + rank.  The CTA's points then leave as
        // one run per leaf -- consecutive slots, written by neighbouring lanes: one or two sector writes per run
        // instead of one per point.
        if (coal_cap > 0 && hi - lo <= coal_cap) {
            uint32_t *loc = tbl + nl; // [nl + 1] first local slot of every leaf
            uint16_t *sorted = reinterpret_cast<uint16_t *>(loc + nl + 1 + ((nl + 1) & 1));
            const uint32_t *cntg = base_tbl + (size_t)(gridDim.x + blockIdx.x) * nl;
            uint32_t mine2 = 0;
            for (uint32_t l = l0; l < l1; ++l)
                mine2 += cntg[l];
            uint32_t incl2 = mine2;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl2, o);
                if (lane >= o)
                    incl2 += t;
            }
            if (lane == 31)
                warp_sums[warp] = incl2;
            __syncthreads();
            uint32_t off = incl2 - mine2;
            for (uint32_t w = 0; w < warp; ++w)
                off += warp_sums[w];
            for (uint32_t l = l0; l < l1; ++l) {
                loc[l] = off;
                off += cntg[l];
            }
            if (l1 == nl)
                loc[nl] = off; // every thread at or past the end holds the total
            __syncthreads();
            for (uint32_t b0 = lo; b0 < hi; b0 += per_it) { // keys -> local order (16-bit offsets inside the CTA's range)
                const uint32_t i0 = b0 + KPT * tid;
                uint4 qa = make_uint4(kKeyNoLeaf, kKeyNoLeaf, kKeyNoLeaf, kKeyNoLeaf), qb = qa;
                if (i0 < hi)
                    qa = __ldcs(reinterpret_cast<const uint4 *>(key + i0));
                if (i0 + 4 < hi)
                    qb = __ldcs(reinterpret_cast<const uint4 *>(key + i0 + 4));
                const uint32_t w[KPT] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
                for (int j = 0; j < (int)KPT; ++j) {
                    const uint32_t leaf = w[j] & kKeyNoLeaf;
                    if (leaf != kKeyNoLeaf && i0 + j < hi)
                        sorted[loc[leaf] + (w[j] >> kKeyLeafBits)] = (uint16_t)(i0 + j - lo);
                }
            }
            __syncthreads();
            // copy-out: G lanes per run (8 when the runs are short, a whole warp when they are long)
            const uint32_t G = loc[nl] >= 24u * nl ? 32u : 8u, sub = tid & (G - 1);
            for (uint32_t l = tid / G; l < nl; l += nthr / G) {
                const uint32_t b = loc[l], len = loc[l + 1] - b, g = tbl[l];
                for (uint32_t e = sub; e < len; e += G)
                    perm[g + e] = lo + sorted[b + e];
            }
            return;
        }
        auto ld = [&](uint32_t b, int half) -> uint4 {
            const uint32_t i = b + KPT * tid + 4 * half;
            return b < hi && i < hi ? __ldcs(reinterpret_cast<const uint4 *>(key + i)) : make_uint4(kKeyNoLeaf, kKeyNoLeaf, kKeyNoLeaf, kKeyNoLeaf);
        };
        uint4 q0a = ld(lo, 0), q0b = ld(lo, 1), q1a = ld(lo + per_it, 0), q1b = ld(lo + per_it, 1);
        for (uint32_t b0 = lo; b0 < hi; b0 += per_it) {
            const uint32_t w[KPT] = {q0a.x, q0a.y, q0a.z, q0a.w, q0b.x, q0b.y, q0b.z, q0b.w};
            q0a = q1a;
            q0b = q1b;
            q1a = ld(b0 + 2 * per_it, 0);
            q1b = ld(b0 + 2 * per_it, 1);
            const uint32_t i0 = b0 + KPT * tid;
            uint32_t slot[KPT];
            bool ok[KPT];
#pragma unroll
            for (int j = 0; j < (int)KPT; ++j) {
                const uint32_t leaf = w[j] & kKeyNoLeaf;
                ok[j] = leaf != kKeyNoLeaf && i0 + j < hi;
                slot[j] = ok[j] ? tbl[leaf] + (w[j] >> kKeyLeafBits) : 0u;
            }
#pragma unroll
            for (int j = 0; j < (int)KPT; ++j)
                if (ok[j])
                    perm[slot[j]] = i0 + j;
        }
        return;
    }
    uint32_t kn[kSortPts]; // keys that come from global memory are requested one iteration ahead
    if (!use_skey) {
#pragma unroll
        for (int u = 0; u < kSortPts; ++u) {
            const uint32_t i = lo + tid + u * nthr;
            kn[u] = i < hi ? __ldcs(key + i) : kNotLeaf;
        }
    }
    for (uint32_t b0 = lo; b0 < hi; b0 += kSortPts * nthr) {
        const uint32_t i0 = b0 + tid;
        uint32_t k[kSortPts];
#pragma unroll
        for (int u = 0; u < kSortPts; ++u) {
            const uint32_t i = i0 + u * nthr;
            if (use_skey) {
                const uint32_t v = i < hi ? skey[i - lo] : 0xFFFFu;
                k[u] = v == 0xFFFFu ? kNotLeaf : v;
            } else {
                k[u] = kn[u];
            }
        }
        if (!use_skey) {
#pragma unroll
            for (int u = 0; u < kSortPts; ++u) {
                const uint32_t i = i0 + kSortPts * nthr + u * nthr;
                kn[u] = i < hi ? __ldcs(key + i) : kNotLeaf;
            }
        }
        if (!keys_repeat(k[0], lane)) { // the usual case, kept free of branches: all atomics in flight, then the stores
            uint32_t slot[kSortPts];
#pragma unroll
            for (int u = 0; u < kSortPts; ++u)
                slot[u] = k[u] != kNotLeaf ? atomicAdd(tbl + k[u], 1u) : 0u;
#pragma unroll
            for (int u = 0; u < kSortPts; ++u)
                if (k[u] != kNotLeaf)
                    perm[slot[u]] = i0 + u * nthr;
        } else {
#pragma unroll
            for (int u = 0; u < kSortPts; ++u) {
                const uint32_t slot = grouped_add(tbl, k[u], true, lane);
                if (k[u] != kNotLeaf)
                    perm[slot] = i0 + u * nthr;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------ K2
// Single CTA: exclusive scan of the per-leaf counts rounded up to the warp-tile size.  count[] in: histogram;
// out: zero (ready for the next chunk's K1).  cursor[] out: zero (for K3).  cnt[] out: dense copy for K2b.
constexpr int kScanItems = 8; // sub-blocks of 1024 leaves per pass: their counters are all requested before the first scan
__global__ void __launch_bounds__(1024)
bin_scan_kernel(uint32_t TILE, uint32_t n_leaves, uint32_t *__restrict__ count, uint32_t *__restrict__ cursor, int cshift,
                uint32_t *__restrict__ cnt_out, uint32_t *__restrict__ start, uint32_t *__restrict__ n_tiles) {
    __shared__ uint32_t warp_sums[2][32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t carry = 0; // every thread tracks it: no shared carry, two barriers per 1024 leaves
    int buf = 0;
    for (uint32_t base = 0; base < n_leaves; base += blockDim.x * kScanItems) {
        uint32_t cnt[kScanItems];
#pragma unroll
        for (int e = 0; e < kScanItems; ++e) { // coalesced, kScanItems loads in flight
            const uint32_t l = base + e * blockDim.x + threadIdx.x;
            cnt[e] = l < n_leaves ? count[(size_t)l << cshift] : 0u;
        }
#pragma unroll
        for (int e = 0; e < kScanItems; ++e) {
            const uint32_t l = base + e * blockDim.x + threadIdx.x;
            if (base + e * blockDim.x >= n_leaves) // CTA uniform
                break;
            const uint32_t padded = (cnt[e] + TILE - 1) / TILE * TILE;
            uint32_t v = padded;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o)
                    v += t;
            }
            if (lane == 31)
                warp_sums[buf][warp] = v;
            __syncthreads();
            uint32_t w = warp_sums[buf][lane]; // every warp scans the 32 warp totals itself
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o)
                    w += t;
            }
            const uint32_t before = __shfl_sync(0xffffffffu, w, warp ? warp - 1 : 0);
            const uint32_t total = __shfl_sync(0xffffffffu, w, 31);
            if (l < n_leaves) {
                start[l] = carry + (warp ? before : 0u) + v - padded;
                cnt_out[l] = cnt[e];
                count[(size_t)l << cshift] = 0;
                cursor[(size_t)l << cshift] = 0;
            }
            carry += total;
            buf ^= 1; // the other buffer: no second barrier needed before the next sub-block writes
        }
    }
    if (threadIdx.x == 0)
        *n_tiles = carry / TILE;
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
// 2D needs the fused sort (per-leaf histogram in shared memory); 3D falls back to the four-kernel sort
bool binned_fits(const DeviceTree &dt) {
    return dt.view.dim == 3 || (size_t)dt.view.n_leaves * sizeof(uint32_t) <= kFusedSortMaxSmem;
}

bool binned_supported(int dim, int order) { return (dim == 2 || dim == 3) && order >= 6 && order <= 16 && order % 2 == 0; }

// third histogram buffer, second key buffer, the per-CTA base tables, side stream and events of the split sort
static cudaError_t alloc_split_buffers(BinWorkspace &ws, size_t cwords, int64_t chunk, size_t n_keys) {
    cudaError_t e = bz_malloc((void **)&ws.count3, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = cudaMemset(ws.count3, 0, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.alt_key, sizeof(uint32_t) * chunk);
    for (int h = 0; h < 2 && e == cudaSuccess; ++h)
        e = bz_malloc((void **)&ws.base_tbl[h], sizeof(uint32_t) * 2 * (size_t)ws.split_ctas * n_keys); // bases | counts
    if (e == cudaSuccess)
        e = cudaStreamCreateWithFlags(&ws.sort_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess)
        e = cudaEventCreateWithFlags(&ws.ev_begin, cudaEventDisableTiming);
    for (int h = 0; h < 2 && e == cudaSuccess; ++h) {
        e = cudaEventCreateWithFlags(&ws.ev_sorted[h], cudaEventDisableTiming);
        if (e == cudaSuccess)
            e = cudaEventCreateWithFlags(&ws.ev_evaluated[h], cudaEventDisableTiming);
    }
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(nullptr);
    return e;
}

int alloc_workspace(const DeviceTree &dt, int64_t chunk, BinWorkspace &ws) {
    if (chunk < 1)
        chunk = 1;
    if (chunk > 0x7FFFFFFFll)
        chunk = 0x7FFFFFFFll;
    int P = 0;
    const bool is2d = dt.view.dim == 2;
    if (is2d) {
        P = bin_eval2d_P(dt.view.order);
        ws.variant = -1;
    } else {
        ws.variant = bin_eval_select(dt.view.order, &P);
        if (ws.variant < 0)
            return (int)cudaErrorInvalidValue;
    }
    ws.P = P;
    const void *sort_fn = is2d ? (const void *)bin_sort_fused_kernel<2, false> : (const void *)bin_sort_fused_kernel<3, false>;
    const int64_t tile = 32 * P;
    const int64_t nl = dt.view.n_leaves;
    ws.perm_capacity = chunk + nl * tile;
    ws.tile_capacity = ws.perm_capacity / tile + 1;
    // spread the per-leaf counters: up to 32 words (one 128-byte line) apart while the arrays stay <= 4 MB
    ws.cshift = 0;
    while (ws.cshift < 5 && ((nl + 1) << (ws.cshift + 1)) <= (1 << 20))
        ws.cshift++;
    const size_t cwords = (size_t)(nl + 1) << ws.cshift;
    ws.fused = (size_t)nl * sizeof(uint32_t) <= kFusedSortMaxSmem;
    if (const char *e = std::getenv("BAOBZI_B200_FUSED_SORT"))
        ws.fused = ws.fused && std::atoi(e) != 0;
    ws.sort_ctas = 0;
    if (ws.fused) { // cooperative launch: the grid is what is co-resident
        const size_t hist = (size_t)nl * sizeof(uint32_t);
        cudaFuncSetAttribute(sort_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFusedSortMaxSmem);
        int want = 2;
        if (const char *e = std::getenv("BAOBZI_B200_SORT_CTAS"))
            if (std::atoi(e) > 0)
                want = std::atoi(e);
        // preferred: `want` CTAs per SM, each keeping the 16-bit keys of its range in shared memory
        ws.sort_skey = 0;
        ws.sort_sbin = 0;
        ws.sort_smem = hist;
        int per_sm = 0;
        if (nl < 0xFFFFu && !(std::getenv("BAOBZI_B200_SORT_SKEY") && std::atoi(std::getenv("BAOBZI_B200_SORT_SKEY")) == 0)) {
            const int64_t range = (chunk + (int64_t)dt.sm_count * want - 1) / ((int64_t)dt.sm_count * want);
            int64_t cap = range > kSortThreads * kSortPts ? range : kSortThreads * kSortPts; // small launches use fewer CTAs
            cap = (cap + 7) / 8 * 8;
            const size_t smem = hist + (size_t)cap * sizeof(uint16_t);
            if (smem <= kFusedSortMaxSmem) {
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sort_fn, kSortThreads, smem);
                if (per_sm >= want) {
                    per_sm = want;
                    ws.sort_skey = (uint32_t)cap;
                    ws.sort_smem = smem;
                }
            }
            // ... and the bin_leaf table as 16-bit ranks behind them, when `want` CTAs still fit
            const size_t smem_b = smem + ((size_t)dt.view.n_bins * sizeof(uint16_t) + 15) / 16 * 16;
            const char *sb = std::getenv("BAOBZI_B200_SORT_SBIN");
            if (ws.sort_skey && smem_b <= kFusedSortMaxSmem && !(sb && std::atoi(sb) == 0)) {
                int occ = 0;
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sort_fn, kSortThreads, smem_b);
                if (occ >= want) {
                    ws.sort_sbin = dt.view.n_bins;
                    ws.sort_smem = smem_b;
                }
            }
        }
        if (!ws.sort_skey) {
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sort_fn, kSortThreads, hist);
            per_sm = per_sm < want ? per_sm : want;
        }
        // optional: the coalescing scatter (three tables + the 16-bit local order of the CTA's range).  Measured on
        // B200 (C3): 3.82 ms per 2e8 points against 3.37 ms -- the second shared-memory pass costs more than the merged
        // stores save -- so it is opt-in (BAOBZI_B200_SORT_COAL=1).
        ws.sort_coal = 0;
        if (std::getenv("BAOBZI_B200_SORT_COAL") && std::atoi(std::getenv("BAOBZI_B200_SORT_COAL")) != 0) {
            const int64_t range = (chunk + (int64_t)dt.sm_count * want - 1) / ((int64_t)dt.sm_count * want);
            int64_t cap = range > kSortThreads * kSortPts ? range : kSortThreads * kSortPts;
            cap = (cap + 7) / 8 * 8;
            const size_t smem = 3 * hist + 8 + (size_t)cap * sizeof(uint16_t);
            int occ = 0;
            if (cap <= 0xFFFF && smem <= kFusedSortMaxSmem &&
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sort_fn, kSortThreads, smem) ==
                    cudaSuccess &&
                occ >= want) {
                per_sm = want;
                ws.sort_coal = (uint32_t)cap;
                ws.sort_skey = 0;
                ws.sort_sbin = 0;
                ws.sort_smem = smem;
            }
        }
        ws.sort_ctas = dt.sm_count * per_sm;
        ws.sort_threads = kSortThreads;
        ws.fused = per_sm >= 1;
        // Shared memory / L1 split: by default the driver sizes the carve-out for as many CTAs as shared memory could
        // hold (3 at 65 KB each => 200 or 233 KB), but only per_sm are ever resident (cooperative grid).  Ask for what
        // they need and leave the rest to L1, which serves the bin_leaf / node-record gathers of the lookup: a 2 KB
        // growth of the static shared memory once pushed the carve-out from 200 to 233 KB and the 2D sort from 3.1 to
        // 4.6 ms per 1e8 points (ncu: launch__shared_mem_config_size).
        cudaFuncAttributes fa;
        if (per_sm >= 1 && cudaFuncGetAttributes(&fa, sort_fn) == cudaSuccess) {
            const double need = (double)per_sm * (double)(ws.sort_smem + fa.sharedSizeBytes + 1024);
            int pct = (int)(100.0 * need / (228.0 * 1024.0)) + 1;
            pct = pct > 100 ? 100 : pct;
            if (const char *e = std::getenv("BAOBZI_B200_SORT_CARVEOUT"))
                pct = std::atoi(e);
            if (pct >= 0)
                cudaFuncSetAttribute(sort_fn, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        }
        // Overlap: can one sort CTA per SM live beside the evaluation kernel's CTA?  Then the sort of the next chunk
        // runs concurrently with the evaluation of this one (different streams); the sort is sized to what is left.
        ws.sort_overlap_ok = false;
        int k4_regs = 0;
        size_t k4_smem = 0;
        cudaFuncAttributes sa;
        const char *ov = std::getenv("BAOBZI_B200_OVERLAP");
        // Measured on B200 (C3): beside the evaluation kernel the sort is left 6 warps per SM and takes 3.3x as long, and
        // its L2 traffic slows the evaluation by a third -- 18.1 ms per 2e8 points against 13.6 ms back to back.
        // Off by default; BAOBZI_B200_OVERLAP=1 enables it.
        if (ws.fused && !is2d && ov && std::atoi(ov) != 0 && bin_eval_footprint(ws.variant, &k4_regs, &k4_smem) &&
            cudaFuncGetAttributes(&sa, sort_fn) == cudaSuccess) {
            const int per_warp = (sa.numRegs * 32 + 511) / 512 * 512;
            int warps = (65536 - k4_regs) / per_warp;
            warps = warps > kSortThreads / 32 ? kSortThreads / 32 : warps;
            warps = warps >= 8 ? warps / 4 * 4 : warps; // keep the sub-partitions even
            const size_t smem_left = 227 * 1024 > k4_smem ? 227 * 1024 - k4_smem : 0;
            if (warps >= 4 && hist + sa.sharedSizeBytes + 1024 <= smem_left) {
                ws.sort_overlap_ok = true;
                ws.sort_threads = warps * 32;
                ws.sort_ctas = dt.sm_count;
                ws.sort_skey = 0;
                ws.sort_sbin = 0;
                ws.sort_coal = 0;
                ws.sort_smem = hist;
            }
        }
        if (std::getenv("BAOBZI_B200_DEBUG"))
            std::fprintf(stderr, "baobzi: overlap %d: sort CTA %d threads beside %d K4 registers / %zu B smem\n",
                         (int)ws.sort_overlap_ok, ws.sort_threads, k4_regs, k4_smem);
        if (std::getenv("BAOBZI_B200_DEBUG"))
            std::fprintf(stderr, "baobzi: fused sort: %d CTAs/SM, %zu B smem, %u shared keys / %u local-sort slots per CTA, chunk %lld\n",
                         per_sm, ws.sort_smem, ws.sort_skey, ws.sort_coal, (long long)chunk);
    }
    if (is2d && !ws.fused)
        return (int)cudaErrorInvalidValue; // the four-kernel fallback sort exists for 3D only
    cudaError_t e = bz_malloc((void **)&ws.key, sizeof(uint32_t) * chunk);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.perm, sizeof(uint32_t) * ws.perm_capacity);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.count, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.cursor, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.cnt, sizeof(uint32_t) * (nl + 1));
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.start, sizeof(uint32_t) * (nl + 1));
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.tile_leaf, sizeof(uint32_t) * ws.tile_capacity);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.n_tiles, sizeof(uint32_t));
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.tile_info, sizeof(uint2) * ws.tile_capacity);
    if (e == cudaSuccess)
        e = cudaMemset(ws.count, 0, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = cudaMemset(ws.cursor, 0, sizeof(uint32_t) * cwords);
    // cudaMemset of device memory is asynchronous on the legacy default stream, and the kernels that read these
    // counters run on the caller's stream, which may be non-blocking (the host-pointer pipeline's are): without this
    // wait a first sort could histogram into uninitialised counters when the default stream had a backlog
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(nullptr);
    ws.overlap = ws.fused && ws.sort_overlap_ok;
    // Split sort: opt-in / opt-out through BAOBZI_B200_SORT_SPLIT (value = CTAs per SM and kernel, 0 = off)
    ws.split = false;
    if (e == cudaSuccess && ws.fused && !ws.overlap) {
        // Default: trees from kSortSplitMinLeaves leaves up.  With few leaves the scatter's stores share sectors and the
        // scan + scatter launch alone is quick (57 us per 8 Mi points at 512 leaves against 87 us at 4096): beside the
        // lookup launch both then take ~140 us, more than one after the other (measured, 3D order 12: 6.41e9 evals/s
        // with the single launch, 6.21e9 split; 4096 leaves at order 8: 1.53e10 -> 1.60e10).
        // 2D: from kSortSplitMinLeaves2D leaves.  scripts/split_threshold.py (1e8 points, single launch -> split): 217 leaves
        // 5.51 -> 4.75 ms, 256 leaves 3.98 -> 3.62, 1147 leaves 3.60 -> 3.18, 8860 leaves (C1) 5.15 -> 4.59; the 3D
        // trees below the threshold gain nothing measurable (64 leaves 15.46 -> 15.36 ms, 512 leaves 10.24 -> 10.15).
        const int64_t min_leaves = is2d ? kSortSplitMinLeaves2D : kSortSplitMinLeaves;
        int per = nl >= min_leaves ? kSortSplitDefault : 0;
        if (const char *sp = std::getenv("BAOBZI_B200_SORT_SPLIT"))
            per = std::atoi(sp);
        // Piggyback schedule (3D): the evaluation launch of chunk c carries phase B of chunk c+1 and phase A of chunk
        // c+2 (bin_eval.cu).  Needs the split sort's buffers (one CTA per SM and phase) whatever the leaf count, a
        // second set of perm / tile arrays, and the two shared-memory tables beside the evaluation's stages.
        // packed keys (sort_common.cuh): 15 bits of leaf, 17 bits of rank inside (CTA, leaf)
        const bool keys_fit = nl < (int64_t)kKeyNoLeaf && chunk / dt.sm_count + 16 < (int64_t(1) << (32 - kKeyLeafBits));
        bool want_piggy = !is2d && keys_fit && bin_eval_is_default(ws.variant) && bin_eval_piggy_P(dt.view.order) == P;
        // Opt-in (BAOBZI_B200_PIGGY=1): measured slower than the split sort (DESIGN.md 4.2d has the numbers and why)
        {
            const char *pe = std::getenv("BAOBZI_B200_PIGGY");
            want_piggy = want_piggy && pe && std::atoi(pe) != 0;
        }
        uint32_t piggy_sbin = 0;
        if (want_piggy) {
            piggy_sbin = dt.view.n_bins;
            if (bin_eval_piggy_smem(dt.view.order, (uint32_t)nl, piggy_sbin) > 227 * 1024)
                piggy_sbin = 0;
            if (bin_eval_piggy_smem(dt.view.order, (uint32_t)nl, 0) > 227 * 1024)
                want_piggy = false;
        }
        const bool split_wanted = per > 0;
        if (want_piggy)
            per = 1; // the standalone phase launches that start the pipeline use one CTA per SM, like the side jobs
        if (per > 0 && keys_fit) { // leaf + rank in one 32-bit key between the two launches
            const size_t hist = (size_t)nl * sizeof(uint32_t);
            const size_t tblb = ((size_t)dt.view.n_bins * sizeof(uint16_t) + 15) / 16 * 16;
            ws.split_sbin = nl < 0xFFFFu && hist + tblb <= kFusedSortMaxSmem ? dt.view.n_bins : 0u;
            ws.split_smem_a = hist + (ws.split_sbin ? tblb : 0);
            ws.split_smem_b = hist;
            // coalescing scatter of the scatter launch: [leaf] local starts + 16-bit local order of the CTA's range
            ws.split_coal = 0;
            {
                const int64_t range = (chunk + (int64_t)dt.sm_count * per - 1) / (dt.sm_count * per) + 16;
                const size_t need = hist + sizeof(uint32_t) * ((size_t)nl + 2) + sizeof(uint16_t) * (size_t)((range + 7) / 8 * 8);
                const char *ce = std::getenv("BAOBZI_B200_SPLIT_COAL");
                // 3D only by default: measured on B200 (profiles/r02_sweep_e.txt) C3, 14 points per (CTA, leaf) run: scatter
                // span 104 -> 65 us per chunk; the 2D benchmark tree (8860 leaves, 3 points per run: nothing to merge, one
                // more pass through shared memory) 2.09 -> 2.72 ms per 1e8 points, so 2D keeps the direct scatter
                const bool want_coal = ce ? std::atoi(ce) != 0 : !is2d;
                if (range <= 65536 && need <= kFusedSortMaxSmem && want_coal) {
                    ws.split_coal = (uint32_t)range;
                    ws.split_smem_b = need;
                }
            }
            ws.split_ctas = dt.sm_count * per;
            ws.split = split_wanted;
            ws.piggy = want_piggy;
            ws.piggy_sbin = piggy_sbin;
            cudaFuncAttributes fa;
            if (cudaFuncGetAttributes(&fa, sort_fn) == cudaSuccess) {
                // both phases' CTAs must be resident together; never below 32 KB (a carve-out sized for a tree with a
                // few hundred leaves left no room for the second kernel's CTA: the phases ran one after the other)
                double need = (double)per * (double)(ws.split_smem_a + ws.split_smem_b + 2 * fa.sharedSizeBytes + 2048);
                need = need < 32.0 * 1024.0 ? 32.0 * 1024.0 : need;
                int pct = (int)(100.0 * need / (228.0 * 1024.0)) + 1;
                pct = pct > 100 ? 100 : pct;
                if (const char *ce = std::getenv("BAOBZI_B200_SORT_CARVEOUT"))
                    pct = std::atoi(ce);
                if (pct >= 0) {
                    const void *fa1 = is2d ? (const void *)bin_sort_fused_kernel<2, false, 1> : (const void *)bin_sort_fused_kernel<3, false, 1>;
                    const void *fa2 = is2d ? (const void *)bin_sort_fused_kernel<2, false, 2> : (const void *)bin_sort_fused_kernel<3, false, 2>;
                    cudaFuncSetAttribute(fa1, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
                    cudaFuncSetAttribute(fa2, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
                    if (!is2d) { // experiment variants of the lookup launch (BAOBZI_B200_SPLIT_APTS)
                        cudaFuncSetAttribute(bin_sort_fused_kernel<3, false, 1, 8>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
                        cudaFuncSetAttribute(bin_sort_fused_kernel<3, false, 1, 2>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
                    }
                }
            }
            {   // trees above ~10 K leaves: the phase launches need more than the 48 KB a kernel gets without opting in
                const void *fa1 = is2d ? (const void *)bin_sort_fused_kernel<2, false, 1> : (const void *)bin_sort_fused_kernel<3, false, 1>;
                const void *fa2 = is2d ? (const void *)bin_sort_fused_kernel<2, false, 2> : (const void *)bin_sort_fused_kernel<3, false, 2>;
                cudaFuncSetAttribute(fa1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFusedSortMaxSmem);
                cudaFuncSetAttribute(fa2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFusedSortMaxSmem);
                if (!is2d) { // experiment variants of the lookup launch (BAOBZI_B200_SPLIT_APTS)
                    cudaFuncSetAttribute(bin_sort_fused_kernel<3, false, 1, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFusedSortMaxSmem);
                    cudaFuncSetAttribute(bin_sort_fused_kernel<3, false, 1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFusedSortMaxSmem);
                }
            }
            e = alloc_split_buffers(ws, cwords, chunk, (size_t)nl);
            if (e == cudaSuccess && ws.piggy) {
                e = bz_malloc((void **)&ws.alt_perm, sizeof(uint32_t) * ws.perm_capacity);
                if (e == cudaSuccess)
                    e = bz_malloc((void **)&ws.alt_tile_leaf, sizeof(uint32_t) * ws.tile_capacity);
                if (e == cudaSuccess)
                    e = bz_malloc((void **)&ws.alt_tile_info, sizeof(uint2) * ws.tile_capacity);
                if (e == cudaSuccess)
                    e = bz_malloc((void **)&ws.alt_n_tiles, sizeof(uint32_t));
            }
        }
    }
    if (e == cudaSuccess && ws.overlap) {
        e = bz_malloc((void **)&ws.alt_key, sizeof(uint32_t) * chunk);
        if (e == cudaSuccess)
            e = bz_malloc((void **)&ws.alt_perm, sizeof(uint32_t) * ws.perm_capacity);
        if (e == cudaSuccess)
            e = bz_malloc((void **)&ws.alt_tile_leaf, sizeof(uint32_t) * ws.tile_capacity);
        if (e == cudaSuccess)
            e = bz_malloc((void **)&ws.alt_tile_info, sizeof(uint2) * ws.tile_capacity);
        if (e == cudaSuccess)
            e = bz_malloc((void **)&ws.alt_n_tiles, sizeof(uint32_t));
        if (e == cudaSuccess)
            e = cudaStreamCreateWithFlags(&ws.sort_stream, cudaStreamNonBlocking);
        if (e == cudaSuccess)
            e = cudaEventCreateWithFlags(&ws.ev_begin, cudaEventDisableTiming);
        for (int h = 0; h < 2 && e == cudaSuccess; ++h) {
            e = cudaEventCreateWithFlags(&ws.ev_sorted[h], cudaEventDisableTiming);
            if (e == cudaSuccess)
                e = cudaEventCreateWithFlags(&ws.ev_evaluated[h], cudaEventDisableTiming);
        }
    }
    if (e != cudaSuccess) {
        free_workspace(ws);
        return (int)e;
    }
    ws.chunk_capacity = chunk;
    return 0;
}

void free_workspace(BinWorkspace &ws) {
    bz_free(ws.key);
    bz_free(ws.perm);
    bz_free(ws.count);
    bz_free(ws.cursor);
    bz_free(ws.cnt);
    bz_free(ws.start);
    bz_free(ws.tile_leaf);
    bz_free(ws.n_tiles);
    bz_free(ws.tile_info);
    bz_free(ws.count3);
    bz_free(ws.base_tbl[0]);
    bz_free(ws.base_tbl[1]);
    bz_free(ws.alt_key);
    bz_free(ws.alt_perm);
    bz_free(ws.alt_tile_leaf);
    bz_free(ws.alt_tile_info);
    bz_free(ws.alt_n_tiles);
    bz_free(ws.bucket_desc);
    bz_free(ws.alt_start);
    if (ws.sort_stream)
        cudaStreamDestroy(ws.sort_stream);
    if (ws.ev_begin)
        cudaEventDestroy(ws.ev_begin);
    for (int h = 0; h < 2; ++h) {
        if (ws.ev_sorted[h])
            cudaEventDestroy(ws.ev_sorted[h]);
        if (ws.ev_evaluated[h])
            cudaEventDestroy(ws.ev_evaluated[h]);
    }
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
    int rc_sort = 0;
    if (ws.fused && ws.overlap) {
        // sort(c) goes to the side stream into buffer set `half`; it waits only for the call's inputs (ev_begin) and
        // for the evaluation that last read this set (chunk c-2), so it runs beside the evaluation of chunk c-1.
        const int h = ws.half;
        cudaStream_t ss = ws.sort_stream;
        if (ws.evaluated_valid[h])
            cudaStreamWaitEvent(ss, ws.ev_evaluated[h], 0);
        if (pf) {
            cudaEventRecord(ev[0], ss);
        }
        uint32_t grid = (uint32_t)ws.sort_ctas;
        const uint32_t per_cta = (uint32_t)ws.sort_threads * kSortPts;
        const uint32_t useful = (n + per_cta - 1) / per_cta;
        grid = grid < useful ? grid : (useful ? useful : 1u);
        uint32_t *cur = ws.flip ? ws.cursor : ws.count, *next = ws.flip ? ws.count : ws.cursor;
        ws.flip = !ws.flip;
        TreeView tvv = tv;
        int cshift = ws.cshift;
        uint32_t tile = TILE, nn = n, skey = 0, nkeys = tv.n_leaves, coal = 0, sbin = 0, phase0 = 0;
        uint32_t *no_base = nullptr;
        int bshift = 0;
        void *args[] = {&tvv, (void *)&d_x, (void *)&d_res, &ws.key, &cur, &next, &cshift, &tile, &ws.perm,
                        &ws.tile_leaf, &ws.tile_info, &ws.n_tiles, &nn, &skey, &nkeys, &bshift, &ws.start, &coal, &sbin, &phase0, &no_base};
        rc_sort = (int)cudaLaunchCooperativeKernel(tv.dim == 2 ? (void *)bin_sort_fused_kernel<2, false> : (void *)bin_sort_fused_kernel<3, false>,
                                                   dim3(grid), dim3(ws.sort_threads), args,
                                                   ws.sort_smem, ss);
        if (pf) {
            cudaEventRecord(ev[1], ss);
            pf->spans.push_back({kStageLookup, ev[0], ev[1], (int64_t)n});
        }
        cudaEventRecord(ws.ev_sorted[h], ss);
        cudaStreamWaitEvent(st, ws.ev_sorted[h], 0);
        if (pf)
            cudaEventRecord(ev[3], st);
        int rc = launch_bin_eval3d(dt, ws, d_x, d_res, st);
        if (pf) {
            cudaEventRecord(ev[4], st);
            pf->spans.push_back({kStageEval, ev[3], ev[4], (int64_t)n});
            pf->pool.push_back(ev[2]);
        }
        cudaEventRecord(ws.ev_evaluated[h], st);
        ws.evaluated_valid[h] = true;
        // the next chunk uses the other buffer set
        std::swap(ws.key, ws.alt_key);
        std::swap(ws.perm, ws.alt_perm);
        std::swap(ws.tile_leaf, ws.alt_tile_leaf);
        std::swap(ws.tile_info, ws.alt_tile_info);
        std::swap(ws.n_tiles, ws.alt_n_tiles);
        ws.half ^= 1;
        return rc_sort ? rc_sort : rc ? rc : (int)cudaGetLastError();
    }
    if (ws.fused) {
        uint32_t grid = (uint32_t)ws.sort_ctas;
        const uint32_t per_cta = (uint32_t)ws.sort_threads * kSortPts;
        const uint32_t useful = (n + per_cta - 1) / per_cta;
        grid = grid < useful ? grid : (useful ? useful : 1u);
        // launch-sized calls: every CTA flushes and scans the whole histogram, so below ~8 K points per CTA fewer CTAs
        // win (C1 tree, 8860 leaves, 1 M points: 53.7 us with one CTA per SM against 59.4 with two; 4 M points: 132 / 118)
        if (n / 8192u < grid) {
            const uint32_t one_per_sm = (uint32_t)dt.sm_count < grid ? (uint32_t)dt.sm_count : grid;
            grid = n / 8192u > one_per_sm ? n / 8192u : one_per_sm;
        }
        uint32_t *cur = ws.flip ? ws.cursor : ws.count, *next = ws.flip ? ws.count : ws.cursor;
        ws.flip = !ws.flip;
        if (ws.count3) { // the split sort's three rotating histogram buffers (calls of both kinds may alternate)
            uint32_t *hist[3] = {ws.count, ws.cursor, ws.count3};
            cur = hist[ws.rot % 3];
            next = hist[(ws.rot + 2) % 3];
            ws.rot = (ws.rot + 1) % 3;
        }
        TreeView tvv = tv;
        int cshift = ws.cshift;
        uint32_t tile = TILE, nn = n, skey = ws.sort_skey, nkeys = tv.n_leaves, coal = ws.sort_coal, sbin = ws.sort_sbin, phase0 = 0;
        uint32_t *no_base = nullptr;
        int bshift = 0;
        void *args[] = {&tvv, (void *)&d_x, (void *)&d_res, &ws.key, &cur, &next, &cshift, &tile, &ws.perm,
                        &ws.tile_leaf, &ws.tile_info, &ws.n_tiles, &nn, &skey, &nkeys, &bshift, &ws.start, &coal, &sbin, &phase0, &no_base};
        rc_sort = (int)cudaLaunchCooperativeKernel(tv.dim == 2 ? (void *)bin_sort_fused_kernel<2, false> : (void *)bin_sort_fused_kernel<3, false>,
                                                   dim3(grid), dim3(ws.sort_threads), args,
                                                   ws.sort_smem, st);
        if (pf) // the three classic stages are one launch here: all of it is booked under "sort"
            cudaEventRecord(ev[1], st);
    } else {
        bin_lookup_kernel<3><<<g1, kBinThreads, 0, st>>>(tv, d_x, d_res, ws.key, ws.count, ws.cshift, n);
        if (pf)
            cudaEventRecord(ev[1], st);
        bin_scan_kernel<<<1, 1024, 0, st>>>(TILE, tv.n_leaves, ws.count, ws.cursor, ws.cshift, ws.cnt, ws.start, ws.n_tiles);
        bin_fill_kernel<<<(tv.n_leaves * 32 + kBinThreads - 1) / kBinThreads, kBinThreads, 0, st>>>(
            tv, TILE, tv.n_leaves, ws.cnt, ws.start, ws.tile_leaf, ws.tile_info, ws.perm);
        if (pf)
            cudaEventRecord(ev[2], st);
        bin_scatter_kernel<<<g1, kBinThreads, 0, st>>>(ws.key, ws.start, ws.cursor, ws.cshift, ws.perm, n);
        if (pf)
            cudaEventRecord(ev[3], st);
    }
    int rc = tv.dim == 2 ? launch_bin_eval2d(dt, ws, d_x, d_res, st) : launch_bin_eval3d(dt, ws, d_x, d_res, st);
    if (pf) {
        cudaEventRecord(ev[4], st);
        pf->spans.push_back({kStageLookup, ev[0], ev[1], (int64_t)n});
        if (ws.fused) {
            pf->spans.push_back({kStageEval, ev[1], ev[4], (int64_t)n});
            pf->pool.push_back(ev[2]);
            pf->pool.push_back(ev[3]);
        } else {
            pf->spans.push_back({kStageScan, ev[1], ev[2], (int64_t)n});
            pf->spans.push_back({kStageScatter, ev[2], ev[3], (int64_t)n});
            pf->spans.push_back({kStageEval, ev[3], ev[4], (int64_t)n});
        }
    }
    return rc_sort ? rc_sort : rc ? rc : (int)cudaGetLastError();
}

// One phase of the split sort for one chunk (ordinary launch: the kernel never reaches its grid barrier).
static int launch_sort_phase(const DeviceTree &dt, BinWorkspace &ws, uint32_t phase, const double *d_x, double *d_res,
                             uint32_t n, uint32_t *key, uint32_t *cur, uint32_t *next, uint32_t *base, cudaStream_t st,
                             bool full_grid = false, int half = 0) {
    const TreeView &tv = dt.view;
    // half = 1: the scatter fills the second set of output buffers (bucketed path with the sort beside the evaluation)
    uint32_t *const out_perm = half ? ws.alt_perm : ws.perm, *const out_start = half ? ws.alt_start : ws.start;
    uint32_t *const out_n_tiles = half ? ws.alt_n_tiles : ws.n_tiles;
    uint32_t grid = (uint32_t)ws.split_ctas;
    const uint32_t per_cta = (uint32_t)kSortThreads * kSortPts;
    const uint32_t useful = (n + per_cta - 1) / per_cta;
    if (!full_grid) // (the piggyback schedule pairs these launches with side jobs of a fixed grid: one CTA per SM)
        grid = grid < useful ? grid : (useful ? useful : 1u); // a function of n alone: the same for both phases of a chunk
    const bool bybin = ws.bucketed; // bucketed path: key = grid cell >> bshift, no tile descriptors, no table
    const uint32_t TILE = bybin ? 1u : 32u * (uint32_t)ws.P;
    const size_t smem = phase == 1 ? ws.split_smem_a : ws.split_smem_b;
    const uint32_t coal = phase == 2 && !ws.bucketed ? ws.split_coal : 0u;
    const uint32_t sbin = phase == 1 ? ws.split_sbin : 0u;
    // Threads of the two launches (one CTA of each per SM).  In 2D the lookup (grid cell + descent) is the longer of the
    // two, so it gets 768 threads and the scatter 256; in 3D 512 + 512.  Measured on B200, A+B threads -> ms per step:
    // C1 tree, 1e8 points: 512+512 4.95, 640+384 4.76, 640+256 4.73, 768+256 4.59;
    // C3 tree, 2e8 points: 512+512 12.48, 640+384 12.50, 640+256 12.65, 768+256 12.66.
    static const int a_env = std::getenv("BAOBZI_B200_SPLIT_ATHREADS") ? std::atoi(std::getenv("BAOBZI_B200_SPLIT_ATHREADS")) : 0;
    static const int b_env = std::getenv("BAOBZI_B200_SPLIT_BTHREADS") ? std::atoi(std::getenv("BAOBZI_B200_SPLIT_BTHREADS")) : 0;
    const bool long_lookup = tv.dim == 2 && !bybin;
    // 3D with the coalescing scatter: the lookup launch is the longer one (bound by its one shared atomic per point, 2 cycles
    // per lane = 58 us per 8.4 M points), so it gets 576 threads and the scatter 448: 512+512 12.77, 768+256 12.06,
    // 640+384 11.72-11.86, 640+448 11.62, 576+448 11.58 ms per 2e8 points (direct scatter, 512+512: 12.59)
    const bool coal3d = !long_lookup && !bybin && ws.split_coal > 0;
    const int a_threads = a_env >= 32 && a_env <= kSortThreadsA ? a_env / 32 * 32 : (long_lookup ? kSortThreadsA : (coal3d ? 576 : kSortThreads));
    const int b_threads = b_env >= 32 && b_env <= kSortThreads ? b_env / 32 * 32 : (long_lookup ? 256 : (coal3d ? 448 : kSortThreads));
    const int threads = phase == 2 ? b_threads : a_threads;
    uint32_t *tile_leaf = bybin ? nullptr : ws.tile_leaf;
    uint2 *tile_info = bybin ? nullptr : ws.tile_info;
    const uint32_t nkeys = bybin ? ws.n_keys : tv.n_leaves;
    const int bshift = bybin ? ws.bshift : 0;
#define BZ_SORT_PHASE(D, BB, PH)                                                                                     \
    bin_sort_fused_kernel<D, BB, PH><<<grid, threads, smem, st>>>(tv, d_x, d_res, key, cur, next, ws.cshift, TILE,  \
                                                                  out_perm, tile_leaf, tile_info, out_n_tiles, n, 0u, \
                                                                  nkeys, bshift, out_start, coal, sbin, phase, base)
#define BZ_SORT_PHASES(D, BB)                                                                                        \
    do {                                                                                                             \
        if (phase == 1)                                                                                              \
            BZ_SORT_PHASE(D, BB, 1);                                                                                 \
        else                                                                                                         \
            BZ_SORT_PHASE(D, BB, 2);                                                                                 \
    } while (0)
    if (bybin) {
        if (tv.dim == 1)
            BZ_SORT_PHASES(1, true);
        else if (tv.dim == 2)
            BZ_SORT_PHASES(2, true);
        else
            BZ_SORT_PHASES(3, true);
    } else if (tv.dim == 2) {
        BZ_SORT_PHASES(2, false);
    } else {
        static const int apts = std::getenv("BAOBZI_B200_SPLIT_APTS") ? std::atoi(std::getenv("BAOBZI_B200_SPLIT_APTS")) : kSortPts;
        if (phase == 1 && apts == 8)
            bin_sort_fused_kernel<3, false, 1, 8><<<grid, threads, smem, st>>>(tv, d_x, d_res, key, cur, next, ws.cshift, TILE, ws.perm,
                                                                              tile_leaf, tile_info, ws.n_tiles, n, 0u, nkeys, bshift,
                                                                              ws.start, coal, sbin, phase, base);
        else if (phase == 1 && apts == 2)
            bin_sort_fused_kernel<3, false, 1, 2><<<grid, threads, smem, st>>>(tv, d_x, d_res, key, cur, next, ws.cshift, TILE, ws.perm,
                                                                              tile_leaf, tile_info, ws.n_tiles, n, 0u, nkeys, bshift,
                                                                              ws.start, coal, sbin, phase, base);
        else
            BZ_SORT_PHASES(3, false);
    }
#undef BZ_SORT_PHASES
#undef BZ_SORT_PHASE
    return (int)cudaGetLastError();
}

static int launch_eval_split(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, int64_t n,
                             cudaStream_t stream, int64_t *launches) {
    const int dim = dt.view.dim;
    const int64_t step = ws.chunk_limit > 0 && ws.chunk_limit < ws.chunk_capacity ? ws.chunk_limit : ws.chunk_capacity;
    const int64_t nc = (n + step - 1) / step;
    uint32_t *hist[3] = {ws.count, ws.cursor, ws.count3};
    uint32_t *keys[2] = {ws.key, ws.alt_key};
    cudaStream_t ss = ws.sort_stream;
    StageProfile *pf = ws.prof;
    auto chunk_n = [&](int64_t c) { return (uint32_t)(n - c * step < step ? n - c * step : step); };
    cudaEventRecord(ws.ev_begin, stream); // fork: the side stream starts once everything queued before this call is done
    cudaStreamWaitEvent(ss, ws.ev_begin, 0);
    int rc = launch_sort_phase(dt, ws, 1, d_x, d_res, chunk_n(0), keys[0], hist[ws.rot % 3], nullptr, ws.base_tbl[0], ss);
    cudaEventRecord(ws.ev_sorted[0], ss);
    for (int64_t c = 0; c < nc && rc == 0; ++c) {
        const int h = (int)(c & 1);
        const uint32_t m = chunk_n(c);
        cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
        cudaStreamWaitEvent(stream, ws.ev_sorted[h], 0);
        if (pf) {
            for (auto &e : ev)
                e = pf->get();
            cudaEventRecord(ev[0], stream);
        }
        rc = launch_sort_phase(dt, ws, 2, d_x + dim * c * step, d_res + c * step, m, keys[h], hist[(ws.rot + c) % 3],
                               hist[(ws.rot + c + 2) % 3], ws.base_tbl[h], stream);
        if (pf)
            cudaEventRecord(ev[1], stream);
        if (c + 1 < nc && rc == 0) { // phase A of the next chunk: beside the scatter above, but after evaluation c-1
            if (c >= 1)
                cudaStreamWaitEvent(ss, ws.ev_evaluated[h ^ 1], 0);
            rc = launch_sort_phase(dt, ws, 1, d_x + dim * (c + 1) * step, d_res + (c + 1) * step, chunk_n(c + 1), keys[h ^ 1],
                                   hist[(ws.rot + c + 1) % 3], nullptr, ws.base_tbl[h ^ 1], ss);
            cudaEventRecord(ws.ev_sorted[h ^ 1], ss);
            // The evaluation below does NOT wait for this launch: its CTAs take the SMs as phase A's CTAs leave them.
            // Measured (C3, 2e8 points): 12.49 ms against 12.79 ms with a wait, at the price of a per-launch evaluation
            // time (bench.py's roofline) that includes phase A's tail: 411 us against 400 us per 8.4 M points.
        }
        if (rc == 0) {
            const double *xc = d_x + dim * c * step;
            double *rcp = d_res + c * step;
            rc = ws.bucketed ? (ws.stage_ok ? launch_eval_buckets_staged(dt, ws, xc, rcp, stream)
                                            : launch_eval_buckets(dt, xc, rcp, ws.perm, ws.start, ws.n_tiles + 1, ws.n_keys, ws.bshift, stream))
                 : dim == 2  ? launch_bin_eval2d(dt, ws, xc, rcp, stream)
                             : launch_bin_eval3d(dt, ws, xc, rcp, stream);
        }
        cudaEventRecord(ws.ev_evaluated[h], stream);
        if (pf) {
            cudaEventRecord(ev[2], stream);
            pf->spans.push_back({kStageLookup, ev[0], ev[1], (int64_t)m});
            pf->spans.push_back({kStageEval, ev[1], ev[2], (int64_t)m});
        }
        if (launches)
            *launches += 3;
    }
    ws.rot = (int)((ws.rot + nc) % 3);
    return rc ? rc : (int)cudaGetLastError();
}

// Bucketed path with the pipelined staged kernel: BOTH sort launches of chunk c+1 run on the side stream beside the
// evaluation of chunk c (the evaluation kernel is compiled for two CTAs' worth of registers per SM and uses two stages, so
// that one sort CTA fits an SM next to it); the scatter writes the second set of perm / key_start buffers.
static int launch_eval_bucket_overlap(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, int64_t n,
                                      cudaStream_t stream, int64_t *launches) {
    const int dim = dt.view.dim;
    const int64_t step = ws.chunk_limit > 0 && ws.chunk_limit < ws.chunk_capacity ? ws.chunk_limit : ws.chunk_capacity;
    const int64_t nc = (n + step - 1) / step;
    uint32_t *hist[3] = {ws.count, ws.cursor, ws.count3};
    uint32_t *keys[2] = {ws.key, ws.alt_key};
    cudaStream_t ss = ws.sort_stream;
    StageProfile *pf = ws.prof;
    auto chunk_n = [&](int64_t c) { return (uint32_t)(n - c * step < step ? n - c * step : step); };
    cudaEventRecord(ws.ev_begin, stream);
    cudaStreamWaitEvent(ss, ws.ev_begin, 0);
    int rc = 0;
    auto sort_chunk = [&](int64_t c) {
        const int h = (int)(c & 1);
        const uint32_t m = chunk_n(c);
        const double *xc = d_x + dim * c * step;
        double *rcp = d_res + c * step;
        if (c >= 2) // evaluation c - 2 read the buffers this scatter fills
            cudaStreamWaitEvent(ss, ws.ev_evaluated[h], 0);
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (pf) {
            e0 = pf->get();
            e1 = pf->get();
            cudaEventRecord(e0, ss);
        }
        rc = launch_sort_phase(dt, ws, 1, xc, rcp, m, keys[h], hist[(ws.rot + c) % 3], nullptr, ws.base_tbl[h], ss, false, h);
        if (rc == 0)
            rc = launch_sort_phase(dt, ws, 2, xc, rcp, m, keys[h], hist[(ws.rot + c) % 3], hist[(ws.rot + c + 2) % 3],
                                   ws.base_tbl[h], ss, false, h);
        if (pf) {
            cudaEventRecord(e1, ss);
            pf->spans.push_back({kStageLookup, e0, e1, (int64_t)m});
        }
        cudaEventRecord(ws.ev_sorted[h], ss);
    };
    sort_chunk(0);
    for (int64_t c = 0; c < nc && rc == 0; ++c) {
        const int h = (int)(c & 1);
        if (c + 1 < nc)
            sort_chunk(c + 1);
        if (rc)
            break;
        cudaStreamWaitEvent(stream, ws.ev_sorted[h], 0);
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (pf) {
            e0 = pf->get();
            e1 = pf->get();
            cudaEventRecord(e0, stream);
        }
        rc = launch_eval_buckets_staged(dt, ws, d_x + dim * c * step, d_res + c * step, stream, h);
        if (pf) {
            cudaEventRecord(e1, stream);
            pf->spans.push_back({kStageEval, e0, e1, (int64_t)chunk_n(c)});
        }
        cudaEventRecord(ws.ev_evaluated[h], stream);
        if (launches)
            *launches += 3;
    }
    ws.rot = (int)((ws.rot + nc) % 3);
    return rc ? rc : (int)cudaGetLastError();
}

// Piggyback schedule: after three standalone sort launches (phases A and B of chunk 0, phase A of chunk 1), launch c
// evaluates chunk c and, between its tiles, scatters chunk c+1 and looks up chunk c+2.  One stream, no events; the
// per-chunk cost is one launch.
static int launch_eval_piggy(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, int64_t n,
                             cudaStream_t stream, int64_t *launches) {
    const int dim = dt.view.dim;
    const int64_t step = ws.chunk_limit > 0 && ws.chunk_limit < ws.chunk_capacity ? ws.chunk_limit : ws.chunk_capacity;
    const int64_t nc = (n + step - 1) / step;
    uint32_t *hist[3] = {ws.count, ws.cursor, ws.count3};
    uint32_t *keys[2] = {ws.key, ws.alt_key};
    uint32_t *perms[2] = {ws.perm, ws.alt_perm}, *tleaf[2] = {ws.tile_leaf, ws.alt_tile_leaf},
             *ntl[2] = {ws.n_tiles, ws.alt_n_tiles};
    uint2 *tinfo[2] = {ws.tile_info, ws.alt_tile_info};
    auto use_set = [&](int k) {
        ws.perm = perms[k];
        ws.tile_leaf = tleaf[k];
        ws.tile_info = tinfo[k];
        ws.n_tiles = ntl[k];
    };
    auto chunk_n = [&](int64_t c) { return (uint32_t)(n - c * step < step ? n - c * step : step); };
    StageProfile *pf = ws.prof;
    cudaEvent_t ev[2] = {nullptr, nullptr};
    if (pf) {
        ev[0] = pf->get();
        ev[1] = pf->get();
        cudaEventRecord(ev[0], stream);
    }
    const int rot = ws.rot;
    use_set(0);
    int rc = launch_sort_phase(dt, ws, 1, d_x, d_res, chunk_n(0), keys[0], hist[rot % 3], nullptr, ws.base_tbl[0], stream, true);
    if (rc == 0)
        rc = launch_sort_phase(dt, ws, 2, d_x, d_res, chunk_n(0), keys[0], hist[rot % 3], hist[(rot + 2) % 3], ws.base_tbl[0],
                               stream, true);
    if (rc == 0 && nc > 1)
        rc = launch_sort_phase(dt, ws, 1, d_x + dim * step, d_res + step, chunk_n(1), keys[1], hist[(rot + 1) % 3], nullptr,
                               ws.base_tbl[1], stream, true);
    if (launches)
        *launches += nc > 1 ? 3 : 2;
    if (pf) {
        cudaEventRecord(ev[1], stream);
        pf->spans.push_back({kStageLookup, ev[0], ev[1], (int64_t)chunk_n(0) + (nc > 1 ? (int64_t)chunk_n(1) : 0)});
    }
    for (int64_t c = 0; c < nc && rc == 0; ++c) {
        PiggyArgs pg{};
        pg.cshift = ws.cshift;
        pg.sbin_n = ws.piggy_sbin;
        if (c + 1 < nc) {
            const int k = (int)((c + 1) & 1);
            pg.nb = chunk_n(c + 1);
            pg.key_b = keys[k];
            pg.count_b = hist[(rot + c + 1) % 3];
            pg.count_clear = hist[(rot + c + 3) % 3];
            pg.base_b = ws.base_tbl[k];
            pg.perm_b = perms[k];
            pg.tile_leaf_b = tleaf[k];
            pg.tile_info_b = tinfo[k];
            pg.n_tiles_b = ntl[k];
        }
        if (c + 2 < nc) {
            const int k = (int)(c & 1);
            pg.na = chunk_n(c + 2);
            pg.xa = d_x + dim * (c + 2) * step;
            pg.resa = d_res + (c + 2) * step;
            pg.key_a = keys[k];
            pg.count_a = hist[(rot + c + 2) % 3];
            pg.base_a = ws.base_tbl[k];
        }
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (pf) {
            e0 = pf->get();
            e1 = pf->get();
            cudaEventRecord(e0, stream);
        }
        use_set((int)(c & 1));
        rc = launch_bin_eval3d_piggy(dt, ws, d_x + dim * c * step, d_res + c * step, stream, pg);
        if (pf) {
            cudaEventRecord(e1, stream);
            pf->spans.push_back({kStageEval, e0, e1, (int64_t)chunk_n(c)});
        }
        if (launches)
            *launches += 1;
    }
    use_set(0);
    ws.rot = (int)((rot + nc) % 3);
    return rc ? rc : (int)cudaGetLastError();
}

int launch_eval_binned(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, int64_t n,
                       cudaStream_t stream, int64_t *launches) {
    if (n <= 0)
        return 0;
    if ((dt.view.dim != 3 && dt.view.dim != 2) || ws.bucketed || ws.chunk_capacity <= 0)
        return (int)cudaErrorInvalidValue;
    {   // the split sort pays when there is a next chunk to overlap with; a single chunk takes the one-launch sort
        const int64_t step = ws.chunk_limit > 0 && ws.chunk_limit < ws.chunk_capacity ? ws.chunk_limit : ws.chunk_capacity;
        if (ws.piggy && n > step && dt.view.dim == 3)
            return launch_eval_piggy(dt, ws, d_x, d_res, n, stream, launches);
        if (ws.split && n > step)
            return launch_eval_split(dt, ws, d_x, d_res, n, stream, launches);
        // experiment knob: the two ordinary launches also for a single chunk (no cooperative launch, no grid barrier)
        static const bool split_single = std::getenv("BAOBZI_B200_SPLIT_SINGLE") && std::atoi(std::getenv("BAOBZI_B200_SPLIT_SINGLE")) != 0;
        // ... and by default in 3D with the coalescing scatter, where they beat the cooperative launch even without a next
        // chunk to overlap with (C3 tree, 8 M points: 0.509 against 0.530 ms; 2D, 1 M / 4 M points: 0.089 / 0.211 against
        // 0.079 / 0.198 -- the fork / join of the second stream costs more than the scatter saves; profiles/r02_sweep_i.txt)
        const bool single3d = dt.view.dim == 3 && ws.split_coal > 0 && n >= (int64_t(1) << 21);
        if (ws.split && (split_single || single3d))
            return launch_eval_split(dt, ws, d_x, d_res, n, stream, launches);
    }
    if (ws.overlap) { // fork: the side stream may start once everything queued before this call is done
        cudaEventRecord(ws.ev_begin, stream);
        cudaStreamWaitEvent(ws.sort_stream, ws.ev_begin, 0);
    }
    int rc = 0;
    const int64_t step = ws.chunk_limit > 0 && ws.chunk_limit < ws.chunk_capacity ? ws.chunk_limit : ws.chunk_capacity;
    for (int64_t o = 0; o < n && rc == 0; o += step) {
        const int64_t m = n - o < step ? n - o : step;
        rc = launch_chunk_3d(dt, ws, d_x + dt.view.dim * o, d_res + o, (uint32_t)m, stream);
        if (launches)
            *launches += ws.fused ? 2 : 5;
    }
    return rc; // every side-stream launch is an ancestor of the last evaluation on `stream`: the join is implicit
}

// ------------------------------------------------------------------------------------------ bucketed path
template <int DIM>
static int launch_bucket_sort(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, uint32_t n,
                              cudaStream_t st) {
    uint32_t grid = (uint32_t)ws.sort_ctas;
    const uint32_t per_cta = (uint32_t)ws.sort_threads * kSortPts;
    const uint32_t useful = (n + per_cta - 1) / per_cta;
    grid = grid < useful ? grid : (useful ? useful : 1u);
    uint32_t *cur = ws.flip ? ws.cursor : ws.count, *next = ws.flip ? ws.count : ws.cursor;
    ws.flip = !ws.flip;
    if (ws.count3) { // the split sort's three rotating histogram buffers (calls of both kinds may alternate)
        uint32_t *hist[3] = {ws.count, ws.cursor, ws.count3};
        cur = hist[ws.rot % 3];
        next = hist[(ws.rot + 2) % 3];
        ws.rot = (ws.rot + 1) % 3;
    }
    TreeView tvv = dt.view;
    int cshift = ws.cshift, bshift = ws.bshift;
    uint32_t tile = 1, nn = n, skey = ws.sort_skey, nkeys = ws.n_keys, coal = ws.sort_coal, sbin = 0, phase0 = 0;
    uint32_t *no_base = nullptr;
    uint32_t *no_tile_leaf = nullptr;
    uint2 *no_tile_info = nullptr;
    void *args[] = {&tvv, (void *)&d_x, (void *)&d_res, &ws.key, &cur, &next, &cshift, &tile, &ws.perm,
                    &no_tile_leaf, &no_tile_info, &ws.n_tiles, &nn, &skey, &nkeys, &bshift, &ws.start, &coal, &sbin, &phase0, &no_base};
    return (int)cudaLaunchCooperativeKernel((void *)bin_sort_fused_kernel<DIM, true>, dim3(grid), dim3(ws.sort_threads),
                                            args, ws.sort_smem, st);
}

template <int DIM>
static int configure_bucket_sort(const DeviceTree &dt, BinWorkspace &ws, int64_t chunk) {
    const size_t hist = (size_t)ws.n_keys * sizeof(uint32_t);
    cudaFuncSetAttribute(bin_sort_fused_kernel<DIM, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFusedSortMaxSmem);
    const int want = 2;
    const int64_t range = (chunk + (int64_t)dt.sm_count * want - 1) / ((int64_t)dt.sm_count * want);
    int64_t cap = range > kSortThreads * kSortPts ? range : kSortThreads * kSortPts;
    cap = (cap + 7) / 8 * 8;
    int per_sm = 0;
    ws.sort_skey = 0;
    ws.sort_smem = hist;
    if (hist + (size_t)cap * sizeof(uint16_t) <= kFusedSortMaxSmem) {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bin_sort_fused_kernel<DIM, true>, kSortThreads,
                                                      hist + (size_t)cap * sizeof(uint16_t));
        if (per_sm >= want) {
            ws.sort_skey = (uint32_t)cap;
            ws.sort_smem = hist + (size_t)cap * sizeof(uint16_t);
        }
    }
    if (!ws.sort_skey)
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bin_sort_fused_kernel<DIM, true>, kSortThreads, hist);
    per_sm = per_sm < want ? per_sm : want;
    ws.sort_coal = 0;
    if (std::getenv("BAOBZI_B200_SORT_COAL") && std::atoi(std::getenv("BAOBZI_B200_SORT_COAL")) != 0) {
        const size_t smem = 3 * hist + 8 + (size_t)cap * sizeof(uint16_t);
        int occ = 0;
        if (cap <= 0xFFFF && smem <= kFusedSortMaxSmem &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, bin_sort_fused_kernel<DIM, true>, kSortThreads, smem) ==
                cudaSuccess &&
            occ >= want) {
            per_sm = want;
            ws.sort_coal = (uint32_t)cap;
            ws.sort_skey = 0;
            ws.sort_smem = smem;
        }
    }
    if (per_sm < 1)
        return (int)cudaErrorInvalidConfiguration;
    ws.sort_ctas = dt.sm_count * per_sm;
    ws.sort_threads = kSortThreads;
    return 0;
}

int alloc_bucket_workspace(const DeviceTree &dt, int64_t chunk, BinWorkspace &ws) {
    if (chunk < 1)
        chunk = 1;
    if (chunk > 0x7FFFFFFFll)
        chunk = 0x7FFFFFFFll;
    ws.bucketed = true;
    // at most kMaxBuckets buckets of 2^bshift consecutive grid cells (cells are x-fastest, nodes and coefficient
    // blocks are stored cell by cell, so a bucket is one contiguous piece of the tree, small enough for L1)
    // Bucket count: 8192 when the pipelined staged kernel (bucket_stage.cu) then serves the tree -- three stages of
    // the > L2 benchmark tree's 48 KB pieces fit an SM; measured 1.76e10 evals/s at 8192 buckets against 1.47e10 for the
    // one-stage kernel and 1.32e10 for the gathering kernel, whose own best is 4096 (profiles/r02_sweep_*) -- else 4096.
    // BAOBZI_B200_BUCKETS fixes it.
    uint64_t candidates[2] = {8192, 4096};
    int n_cand = 2;
    if (const char *e = std::getenv("BAOBZI_B200_BUCKETS"))
        if (std::atoll(e) > 0) {
            candidates[0] = (uint64_t)std::atoll(e);
            n_cand = 1;
        }
    for (int c = 0; c < n_cand; ++c) {
        const uint64_t max_buckets = candidates[c];
        ws.bshift = 0;
        while ((((uint64_t)dt.view.n_bins + (1ull << ws.bshift) - 1) >> ws.bshift) > max_buckets)
            ws.bshift++;
        ws.n_keys = (uint32_t)(((uint64_t)dt.view.n_bins + (1ull << ws.bshift) - 1) >> ws.bshift);
        const int prc = plan_bucket_stage(dt, ws);
        if (prc) {
            free_workspace(ws);
            return prc;
        }
        if ((ws.stage_ok && ws.stage_pipe) || c + 1 == n_cand)
            break;
        bz_free(ws.bucket_desc); // the next candidate describes its own buckets
        ws.bucket_desc = nullptr;
        ws.stage_ok = false;
    }
    ws.cshift = 0;
    ws.fused = true;
    int rc = dt.view.dim == 1 ? configure_bucket_sort<1>(dt, ws, chunk)
             : dt.view.dim == 2 ? configure_bucket_sort<2>(dt, ws, chunk)
                                : configure_bucket_sort<3>(dt, ws, chunk);
    if (rc) {
        free_workspace(ws);
        return rc;
    }
    const size_t cwords = ws.n_keys + 1;
    cudaError_t e = bz_malloc((void **)&ws.key, sizeof(uint32_t) * chunk);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.perm, sizeof(uint32_t) * chunk);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.count, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.cursor, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.n_tiles, 2 * sizeof(uint32_t));
    if (e == cudaSuccess)
        e = bz_malloc((void **)&ws.start, sizeof(uint32_t) * (cwords + 1));
    if (e == cudaSuccess)
        e = cudaMemset(ws.count, 0, sizeof(uint32_t) * cwords);
    if (e == cudaSuccess)
        e = cudaMemset(ws.cursor, 0, sizeof(uint32_t) * cwords);
    // cudaMemset of device memory is asynchronous on the legacy default stream, and the kernels that read these
    // counters run on the caller's stream, which may be non-blocking (the host-pointer pipeline's are): without this
    // wait a first sort could histogram into uninitialised counters when the default stream had a backlog
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(nullptr);
    // split sort for the bucket keys too (same leaf-count rule; BAOBZI_B200_SORT_SPLIT overrides)
    ws.split = false;
    if (e == cudaSuccess) {
        int per = (int64_t)ws.n_keys >= kSortSplitMinLeaves ? kSortSplitDefault : 0;
        if (const char *sp = std::getenv("BAOBZI_B200_SORT_SPLIT"))
            per = std::atoi(sp);
        if (per > 0 && ws.n_keys < kKeyNoLeaf && chunk / dt.sm_count + 16 < (int64_t(1) << (32 - kKeyLeafBits))) {
            const size_t hist = (size_t)ws.n_keys * sizeof(uint32_t);
            ws.split_sbin = 0;
            ws.split_smem_a = ws.split_smem_b = hist;
            ws.split_ctas = dt.sm_count * per;
            ws.split = hist <= 48 * 1024; // ordinary launches below the opt-in shared-memory size
            if (ws.split) {
                double need = (double)per * (double)(2 * hist + 2 * 2304 + 2048);
                need = need < 32.0 * 1024.0 ? 32.0 * 1024.0 : need;
                int pct = (int)(100.0 * need / (228.0 * 1024.0)) + 1;
                pct = pct > 100 ? 100 : pct;
                if (ws.bucket_overlap) // co-resident with the evaluation kernel, which takes the largest carve-out: an SM
                    pct = 100;         // only runs CTAs of kernels that agree on the shared-memory / L1 split
                const void *fa[6] = {(const void *)bin_sort_fused_kernel<1, true, 1>, (const void *)bin_sort_fused_kernel<1, true, 2>,
                                     (const void *)bin_sort_fused_kernel<2, true, 1>, (const void *)bin_sort_fused_kernel<2, true, 2>,
                                     (const void *)bin_sort_fused_kernel<3, true, 1>, (const void *)bin_sort_fused_kernel<3, true, 2>};
                const int d0 = (dt.view.dim - 1) * 2;
                cudaFuncSetAttribute(fa[d0], cudaFuncAttributePreferredSharedMemoryCarveout, pct);
                cudaFuncSetAttribute(fa[d0 + 1], cudaFuncAttributePreferredSharedMemoryCarveout, pct);
                e = alloc_split_buffers(ws, cwords, chunk, (size_t)ws.n_keys);
            }
        }
    }
    if (e == cudaSuccess && ws.bucket_overlap) {
        if (ws.split && ws.stage_ok && ws.stage_pipe) { // second set of scatter outputs
            e = bz_malloc((void **)&ws.alt_perm, sizeof(uint32_t) * chunk);
            if (e == cudaSuccess)
                e = bz_malloc((void **)&ws.alt_start, sizeof(uint32_t) * (cwords + 1));
            if (e == cudaSuccess)
                e = bz_malloc((void **)&ws.alt_n_tiles, 2 * sizeof(uint32_t));
        } else {
            ws.bucket_overlap = false;
        }
    }
    if (e != cudaSuccess) {
        free_workspace(ws);
        return (int)e;
    }
    ws.perm_capacity = chunk;
    ws.chunk_capacity = chunk;
    return 0;
}

int launch_eval_bucketed(const DeviceTree &dt, BinWorkspace &ws, const double *d_x, double *d_res, int64_t n,
                         cudaStream_t stream, int64_t *launches) {
    if (n <= 0)
        return 0;
    if (!ws.bucketed || ws.chunk_capacity <= 0)
        return (int)cudaErrorInvalidValue;
    const int dim = dt.view.dim;
    const int64_t step = ws.chunk_limit > 0 && ws.chunk_limit < ws.chunk_capacity ? ws.chunk_limit : ws.chunk_capacity;
    if (ws.split && ws.bucket_overlap && ws.stage_ok && n > step)
        return launch_eval_bucket_overlap(dt, ws, d_x, d_res, n, stream, launches);
    if (ws.split && n > step)
        return launch_eval_split(dt, ws, d_x, d_res, n, stream, launches);
    StageProfile *pf = ws.prof;
    int rc = 0;
    for (int64_t o = 0; o < n && rc == 0; o += step) {
        const uint32_t m = (uint32_t)(n - o < step ? n - o : step);
        const double *xc = d_x + dim * o;
        double *rcp = d_res + o;
        cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
        if (pf) {
            for (auto &e : ev)
                e = pf->get();
            cudaEventRecord(ev[0], stream);
        }
        rc = dim == 1 ? launch_bucket_sort<1>(dt, ws, xc, rcp, m, stream)
             : dim == 2 ? launch_bucket_sort<2>(dt, ws, xc, rcp, m, stream)
                        : launch_bucket_sort<3>(dt, ws, xc, rcp, m, stream);
        if (pf)
            cudaEventRecord(ev[1], stream);
        if (rc == 0)
            rc = ws.stage_ok ? launch_eval_buckets_staged(dt, ws, xc, rcp, stream)
                             : launch_eval_buckets(dt, xc, rcp, ws.perm, ws.start, ws.n_tiles + 1, ws.n_keys, ws.bshift, stream);
        if (pf) {
            cudaEventRecord(ev[2], stream);
            pf->spans.push_back({kStageLookup, ev[0], ev[1], (int64_t)m});
            pf->spans.push_back({kStageEval, ev[1], ev[2], (int64_t)m});
        }
        if (launches)
            *launches += 2;
    }
    return rc;
}

} // namespace bz
