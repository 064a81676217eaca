// bin_eval.cu -- K4 of the binned path: evaluation of leaf-grouped points (see binned.cu for the sort).
//
// One warp per tile of 32*P consecutive slots of ONE leaf; lane l owns slots l, l+32, ... (P points, register
// blocked).  Because the whole warp sits in one leaf, a coefficient pair is a warp-uniform 16-byte shared-memory
// load whose value feeds 2*P FMAs per lane.
//
// What bounds it (ncu, profiles/): the FP64 pipe issues one warp-wide DFMA per 2 cycles per SM sub-partition.
// A warp-uniform LDS costs one LSU wavefront per coefficient, so with P points per lane the four sub-partitions
// together ask for 2/P wavefronts per cycle at full FP64 rate: P = 2 would be LSU bound, P = 3 needs 2/3 of the
// LSU, P = 4 one half.  The other limit is latency hiding: P*O accumulators + P*O values of T(z) per lane put a
// P = 4 warp at ~250 registers, i.e. 2 warps per scheduler, and two warps leave the pipe idle whenever both wait
// (measured: 70 % busy).  Hence the knobs below: P, warps per CTA, T(z) either held in registers or advanced on
// the fly in the k loop (3 live values per point instead of O), T(y) rebuilt per slab instead of kept live.
//
// Each warp runs a 2-stage software pipeline over its tiles: while tile j is evaluated out of shared memory,
// tile j+1's gathered points and their indices are in flight as cp.async copies, and tile j+2's descriptor is being
// fetched into registers.  A warp walks a contiguous run of tiles, so the leaf's coefficient block and node record
// are fetched once per leaf (RUN; the strided variant refetches them with every tile).
//
// Arithmetic: cheb.cuh's order (bit-identical to the reference's AVX2+FMA build).  Multiplications by
// T_0 = 1.0 are not executed (x * 1.0 == x bit for bit): a row chain starts as fma(c_1, T_1, c_0), and the
// T recurrences run from T_{-1} = T_1 = x so that every slab executes the same code (fma(2x, 1, -x) == x exactly).
#include "baobzi/cheb.cuh"
#include "device_tree.hpp"
#include "baobzi/lookup.cuh"
#include "sort_common.cuh"

#include <cstdio>
#include <cstdlib>

namespace bz {

constexpr uint32_t kInvalidSlotE = 0xFFFFFFFFu;
constexpr int kMaxDevices = 64;

__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
// Experiment (BAOBZI_B200_K4_EVICT=1): the gathered points and the slot indices are read once per chunk and the chunk is
// larger than L2; marked evict-first they should leave the result sectors (four scattered 8-byte stores each) and the
// coefficient blocks alone.  The flag lives in device memory so that no kernel signature changes for it.
__device__ int g_k4_evict_first = 0;
__device__ __forceinline__ uint64_t gather_policy() {
    uint64_t pol;
    if (g_k4_evict_first)
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    else
        asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void cp_async8_hint(void *smem, const void *gmem, uint64_t pol) {
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 8, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)),
                 "l"(gmem), "l"(pol));
}
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int O, int P>
struct PipeStage { // per warp, per stage
    static constexpr int TILE = 32 * P;
    double2 coeff[O * O * O / 2];
    double pts[3 * TILE]; // [coordinate][slot]
    double2 node[4];      // the leaf's 64-byte NodeRec
    uint32_t idx[TILE];
};

// <Ty, acc> in Eigen's packet order (cheb.cuh dot_packets) with Ty[0] == 1.0 known: its product is acc[0] itself.
template <int START, int LEN, int STRIDE, int LANE, int O>
__device__ __forceinline__ double dot_tree_one(const double (&a)[O], const double (&b)[O]) {
    constexpr int i0 = START * STRIDE + LANE;
    if constexpr (i0 != 0) {
        return dot_tree<START, LEN, STRIDE, LANE, O>(a, b);
    } else if constexpr (LEN == 1) {
        return b[0];
    } else if constexpr (LEN == 2) {
        return __fma_rn(a[STRIDE], b[STRIDE], b[0]);
    } else if constexpr (LEN == 3) {
        return __dadd_rn(b[0], dot_tree<START + 1, 2, STRIDE, LANE, O>(a, b)); // fma(1, b0, s) == b0 + s
    } else {
        constexpr int H = LEN / 2;
        return __dadd_rn(dot_tree_one<START, H, STRIDE, LANE, O>(a, b), dot_tree<START + H, LEN - H, STRIDE, LANE, O>(a, b));
    }
}
template <int O>
__device__ __forceinline__ double dot_packets_one(const double (&a)[O], const double (&b)[O]) {
    if constexpr (O % 4 == 0) {
        const double l0 = dot_tree_one<0, O / 4, 4, 0, O>(a, b);
        const double l1 = dot_tree<0, O / 4, 4, 1, O>(a, b);
        const double l2 = dot_tree<0, O / 4, 4, 2, O>(a, b);
        const double l3 = dot_tree<0, O / 4, 4, 3, O>(a, b);
        return __dadd_rn(__dadd_rn(l0, l2), __dadd_rn(l1, l3));
    } else {
        const double l0 = dot_tree_one<0, O / 2, 2, 0, O>(a, b);
        const double l1 = dot_tree<0, O / 2, 2, 1, O>(a, b);
        return __dadd_rn(l0, l1);
    }
}

// One tile: 32 * P slots of ONE leaf whose coefficient block / node record sit in shared memory (coeff_s / node_s) and
// whose gathered points sit in pts_s ([coordinate][slot]) with their result indices in idx_s.  Shared by every 3D
// variant below (per-warp block copies, CTA-shared TMA-staged block), so the arithmetic -- and the bits -- are one.
template <int O, int P, bool TZFLY, bool TYKEEP>
__device__ __forceinline__ void eval_tile3d(const double2 *coeff_s, const double2 *node_s, const double *pts_s,
                                            const uint32_t *idx_s, uint32_t lane, double *__restrict__ res) {
    constexpr int TILE = 32 * P;
    constexpr int NTZ = TZFLY ? 1 : O;
    const double2 n0 = node_s[0], n1 = node_s[1], n2 = node_s[2], n3 = node_s[3];
    const double c0 = n0.x, c1 = n0.y, c2c = n1.x, ih0 = n2.x, ih1 = n2.y, ih2 = n3.x;
    double Tz[P][NTZ], Tyk[P][TYKEEP ? O : 1], acc[P][O];
    double v[P], w[P], x2[P], w2[P], tx[P], tx_prev[P], val[P];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        const bool ok = idx_s[p * 32 + lane] != kInvalidSlotE;
        const double px = ok ? pts_s[0 * TILE + p * 32 + lane] : c0; // padding lanes evaluate the leaf centre
        const double py = ok ? pts_s[1 * TILE + p * 32 + lane] : c1;
        const double pz = ok ? pts_s[2 * TILE + p * 32 + lane] : c2c;
        const double u = __dmul_rn(__dsub_rn(px, c0), ih0);
        v[p] = __dmul_rn(__dsub_rn(py, c1), ih1);
        w[p] = __dmul_rn(__dsub_rn(pz, c2c), ih2);
        if constexpr (TZFLY)
            w2[p] = __dadd_rn(w[p], w[p]);
        else
            cheb_T<O>(w[p], Tz[p]);
        if constexpr (TYKEEP)
            cheb_T<O>(v[p], Tyk[p]);
        x2[p] = __dadd_rn(u, u);
        tx_prev[p] = u; // T_{-1} = T_1
        tx[p] = 1.0;
        val[p] = 0.0;
    }
#pragma unroll 1
    for (int i = 0; i < O; ++i) {
        const double2 *cc = coeff_s + i * (O * O / 2);
        // acc[p][r] = sum_k C_i(r,k) Tz[p][k]: row chains in k order; the k = 0 product is C_i(r,0) itself
        double tk[P], tkm[P];
#pragma unroll
        for (int k = 1; k < O; ++k) {
#pragma unroll
            for (int p = 0; p < P; ++p) {
                if constexpr (TZFLY) {
                    if (k == 1) {
                        tkm[p] = 1.0;
                        tk[p] = w[p];
                    } else {
                        const double tn = __fma_rn(w2[p], tk[p], -tkm[p]);
                        tkm[p] = tk[p];
                        tk[p] = tn;
                    }
                } else {
                    tk[p] = Tz[p][k];
                }
            }
#pragma unroll
            for (int r = 0; r < O; r += 2) {
                const double2 c = cc[(k * O + r) / 2];
                if (k == 1) {
                    const double2 cz = cc[r / 2];
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        acc[p][r] = __fma_rn(c.x, tk[p], cz.x);
                        acc[p][r + 1] = __fma_rn(c.y, tk[p], cz.y);
                    }
                } else {
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        acc[p][r] = __fma_rn(c.x, tk[p], acc[p][r]);
                        acc[p][r + 1] = __fma_rn(c.y, tk[p], acc[p][r + 1]);
                    }
                }
            }
        }
#pragma unroll
        for (int p = 0; p < P; ++p) {
            double d;
            if constexpr (TYKEEP) {
                d = dot_packets_one<O>(Tyk[p], acc[p]);
            } else {
                double vv = v[p];
                asm volatile("" : "+d"(vv)); // opaque per slab: T(y) is rebuilt here, not hoisted and kept live
                double Ty[O];
                cheb_T<O>(vv, Ty);
                d = dot_packets_one<O>(Ty, acc[p]);
            }
            val[p] = __fma_rn(tx[p], d, val[p]);
            const double tn = __fma_rn(x2[p], tx[p], -tx_prev[p]);
            tx_prev[p] = tx[p];
            tx[p] = tn;
        }
    }
#pragma unroll
    for (int p = 0; p < P; ++p) {
        const uint32_t id = idx_s[p * 32 + lane];
        if (id != kInvalidSlotE)
            res[id] = val[p];
    }
}

// TZFLY: T_k(z) advanced inside the k loop (costs O-2 extra FMAs per point and slab, frees (O-3)*P doubles).
// TYKEEP: T(y) computed once per tile and kept live (O*P doubles) instead of rebuilt after every slab.
// RUN: every warp walks a contiguous run of tiles (tiles of one leaf are adjacent) and keeps ONE copy of the current
// leaf's coefficient block, refetched only when the leaf changes; the pipeline stages then hold points only.  Half
// the shared memory per warp, so high orders can run 2-3 times the warps.
template <int O, int P>
struct RunSmem { // per warp
    static constexpr int TILE = 32 * P;
    double2 coeff[O * O * O / 2];
    double2 node[4];
    double pts[2][3 * TILE];
    uint32_t idx[2][TILE];
};

template <int O, int P, int WARPS, bool TZFLY, bool TYKEEP, bool RUN>
__global__ void __launch_bounds__(WARPS * 32, 1)
bin_eval3d_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                  const uint32_t *__restrict__ perm, const uint2 *__restrict__ tile_info,
                  const uint32_t *__restrict__ n_tiles_p) {
    constexpr int TILE = 32 * P;
    constexpr int BLK2 = O * O * O / 2;
    using Stage = PipeStage<O, P>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t n_tiles = __ldg(n_tiles_p);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Stage *stages = reinterpret_cast<Stage *>(smem_raw) + 2 * warp;               // !RUN
    RunSmem<O, P> &rs = reinterpret_cast<RunSmem<O, P> *>(smem_raw)[warp];          // RUN
    const uint32_t nw = gridDim.x * WARPS, wid = blockIdx.x * WARPS + warp;
    const uint32_t stride = RUN ? 1u : nw;
    uint32_t t = RUN ? (uint32_t)(((uint64_t)n_tiles * wid) / nw) : wid;
    const uint32_t t_end = RUN ? (uint32_t)(((uint64_t)n_tiles * (wid + 1)) / nw) : n_tiles;
    auto pts_of = [&](int st) -> double * { return RUN ? rs.pts[st] : stages[st].pts; };
    auto idx_of = [&](int st) -> uint32_t * { return RUN ? rs.idx[st] : stages[st].idx; };

    // descriptor of a tile: {node, coeff16} + the P slot indices of this lane
    const uint64_t pol = gather_policy();
    auto load_meta = [&](uint32_t tt, uint2 &info, uint32_t (&id)[P]) {
        if (tt < t_end) {
            info = __ldg(tile_info + tt);
#pragma unroll
            for (int p = 0; p < P; ++p)
                id[p] = g_k4_evict_first ? __ldcs(perm + (size_t)tt * TILE + p * 32 + lane)
                                         : __ldg(perm + (size_t)tt * TILE + p * 32 + lane);
        } else {
            info = make_uint2(0, 0);
#pragma unroll
            for (int p = 0; p < P; ++p)
                id[p] = kInvalidSlotE;
        }
    };
    auto fetch_leaf = [&](const uint2 &info, double2 *coeff, double2 *node) {
        const double2 *Cg = reinterpret_cast<const double2 *>(tv.coeffs + 2ull * info.y);
#pragma unroll 4
        for (int j = lane; j < BLK2; j += 32)
            cp_async16(&coeff[j], Cg + j);
        if (lane < 4)
            cp_async16(&node[lane], reinterpret_cast<const double2 *>(tv.nodes + info.x) + lane);
    };
    auto issue = [&](uint32_t tt, const uint2 &info, const uint32_t (&id)[P], int sidx) {
        if (tt < t_end) {
            if (!RUN)
                fetch_leaf(info, stages[sidx].coeff, stages[sidx].node);
            double *pts = pts_of(sidx);
            uint32_t *sidxs = idx_of(sidx);
#pragma unroll
            for (int p = 0; p < P; ++p) {
                sidxs[p * 32 + lane] = id[p];
                if (id[p] != kInvalidSlotE) {
                    const double *q = x + 3ull * id[p];
                    cp_async8_hint(&pts[0 * TILE + p * 32 + lane], q, pol);
                    cp_async8_hint(&pts[1 * TILE + p * 32 + lane], q + 1, pol);
                    cp_async8_hint(&pts[2 * TILE + p * 32 + lane], q + 2, pol);
                }
            }
        }
        cp_async_commit();
    };

    uint2 info_n, info_cur;
    uint32_t idx_n[P];
    load_meta(t, info_n, idx_n);
    issue(t, info_n, idx_n, 0);
    info_cur = info_n;
    load_meta(t + stride, info_n, idx_n); // descriptor of the second tile
    uint32_t cur_node = 0xFFFFFFFFu;
    int cur = 0;
    for (; t < t_end; t += stride, cur ^= 1) {
        __syncwarp(); // every lane is done reading stage cur^1 (tile j-1) and, in RUN mode, the coefficient block
        issue(t + stride, info_n, idx_n, cur ^ 1);
        const uint2 info_next = info_n;
        load_meta(t + 2 * stride, info_n, idx_n); // lands while this tile computes
        if (RUN && info_cur.x != cur_node) {    // warp uniform: a new leaf starts here
            cur_node = info_cur.x;
            fetch_leaf(info_cur, rs.coeff, rs.node);
            cp_async_commit();
            cp_async_wait<0>();
        } else {
            cp_async_wait<1>();
        }
        __syncwarp();
        info_cur = info_next;
        const double2 *coeff_s = RUN ? rs.coeff : stages[cur].coeff;
        const double2 *node_s = RUN ? rs.node : stages[cur].node;
        const double *pts_s = pts_of(cur);
        const uint32_t *idx_s = idx_of(cur);

        eval_tile3d<O, P, TZFLY, TYKEEP>(coeff_s, node_s, pts_s, idx_s, lane, res);
    }
    cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------ K4 + sort side jobs
// The sort of a chunk is latency / request-rate bound and uses no FP64; the evaluation is FP64 bound and leaves most
// issue slots and nearly all of the memory system idle -- but its CTA holds ~62 K of the SM's 64 K registers, so no
// sort CTA can ever be resident beside it (two streams, round 1: 18.1 ms against 13.6 back to back).  Registers are
// only full DURING a tile, though: between two tiles a warp holds a handful of values.  So the evaluation warps carry
// the sort of the FOLLOWING chunks themselves, one slice per tile, in the gap between tiles:
//     launch c:  evaluation of chunk c   +   phase B (scan + scatter) of chunk c+1   +   phase A (lookup + histogram)
//                of chunk c+2
// with kernel boundaries as the only global synchronisation (no grid barrier, no spinning): phase A of a chunk ends
// with the launch that carried it (histogram flushed, per-CTA bases saved), phase B of that chunk runs in the next
// launch, its evaluation in the one after.  The data the side jobs need is requested BEFORE the tile's arithmetic
// (cp.async of the slice's points into a per-warp stage, the slice's 16-bit keys into two registers) and consumed
// after it, so their latency hides behind ~3.6 us of DFMAs.  Per tile and lane: ~60 extra FP64-pipe instructions (the
// floor transform and bounds test of 4 points) against ~2300, ~8 shared atomics, 8 stores.
// Every CTA owns the same contiguous point range of a chunk in both phases (as in the split sort of binned.cu, whose
// standalone kernels start the pipeline for chunks 0 and 1).
constexpr int kPiggySlice = 128; // points (phase A) / keys (phase B) per warp and tile

template <int O, int P, int WARPS, bool TZFLY, bool TYKEEP>
__global__ void __launch_bounds__(WARPS * 32, 1)
bin_eval3d_piggy_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                        const uint32_t *__restrict__ perm, const uint2 *__restrict__ tile_info,
                        const uint32_t *__restrict__ n_tiles_p, const PiggyArgs pg) {
    constexpr int TILE = 32 * P;
    constexpr int BLK2 = O * O * O / 2;
    constexpr int KS = kPiggySlice / 32; // per lane
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t warp_sums[32];
    __shared__ uint4 big[kBigLeafCap];
    __shared__ uint32_t n_big;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t nl = tv.n_leaves;
    RunSmem<O, P> &rs = reinterpret_cast<RunSmem<O, P> *>(smem_raw)[warp];
    uint32_t *tbl_b = reinterpret_cast<uint32_t *>(smem_raw + sizeof(RunSmem<O, P>) * WARPS);
    uint32_t *hist_a = tbl_b + nl;
    double *stage_a = reinterpret_cast<double *>(hist_a + nl + (nl & 1)) + (size_t)warp * 3 * kPiggySlice;
    uint16_t *sbin = reinterpret_cast<uint16_t *>(reinterpret_cast<double *>(hist_a + nl + (nl & 1)) + (size_t)WARPS * 3 * kPiggySlice);

    // ---- side-job setup (CTA wide)
    if (tid == 0)
        n_big = 0;
    for (uint32_t l = tid; l < nl; l += WARPS * 32) {
        hist_a[l] = 0;
        tbl_b[l] = pg.nb ? pg.base_b[(size_t)blockIdx.x * nl + l] : 0u;
    }
    for (uint32_t l = tid; l < pg.sbin_n; l += WARPS * 32)
        sbin[l] = (uint16_t)__ldg(tv.bin_leaf + l); // kNotLeaf -> 0xFFFF
    __syncthreads();
    if (pg.nb) // padded scan of chunk c+1's histogram: tbl_b[l] = first free slot of this CTA inside leaf l; descriptors
        sort_scan_describe<false>(tv, tbl_b, pg.count_b, pg.count_clear, pg.cshift, (uint32_t)TILE, pg.perm_b, pg.tile_leaf_b,
                                  pg.tile_info_b, pg.n_tiles_b, nl, nullptr, warp_sums, big, &n_big);
    __syncthreads();

    // this warp's slices: the CTA's range of the chunk (multiples of 8, as in the split sort), cut into WARPS pieces of
    // whole slices
    auto warp_range = [&](uint32_t n, uint32_t &lo, uint32_t &hi) {
        const uint32_t clo = (uint32_t)(((uint64_t)blockIdx.x * n) / gridDim.x) & ~7u;
        const uint32_t chi = blockIdx.x + 1 != gridDim.x ? (uint32_t)(((uint64_t)(blockIdx.x + 1) * n) / gridDim.x) & ~7u : n;
        const uint32_t per = ((chi - clo + WARPS - 1) / WARPS + kPiggySlice - 1) / kPiggySlice * kPiggySlice;
        lo = min(clo + warp * per, chi);
        hi = min(lo + per, chi);
    };
    uint32_t a_pos = 0, a_end = 0, b_pos = 0, b_end = 0;
    if (pg.na)
        warp_range(pg.na, a_pos, a_end);
    if (pg.nb)
        warp_range(pg.nb, b_pos, b_end);

    // phase A, first half: request the slice's points (raw copy of 3 * 128 contiguous doubles)
    auto a_request = [&]() {
        if (a_pos < a_end) {
            const uint32_t m3 = 3u * min((uint32_t)kPiggySlice, a_end - a_pos);
            const double *src = pg.xa + 3ull * a_pos;
#pragma unroll
            for (int j = 0; j < 3 * KS; ++j)
                if (j * 32 + lane < m3)
                    cp_async8(stage_a + j * 32 + lane, src + j * 32 + lane);
        }
        cp_async_commit();
    };
    // phase A, second half (the copy has landed): leaf of every point, key, histogram, NaN for points outside.
    // Written as batches over the slice's four points per lane -- all shared-memory loads, then all floor transforms,
    // then all table lookups, then all atomics, then all stores -- so that the four dependent chains overlap: the first
    // version walked the points one after the other and its ~6000 cycles of exposed shared-memory latency per tile
    // (ncu: short scoreboard 1.19 per issue against 0.07) cost more than the sort launches it replaced.
    auto a_consume = [&]() {
        if (a_pos >= a_end)
            return;
        const uint32_t m = min((uint32_t)kPiggySlice, a_end - a_pos);
        double p[KS][3];
        bool have[KS], in[KS];
        uint32_t lin[KS], leaf[KS], rank[KS];
#pragma unroll
        for (int j = 0; j < KS; ++j) {
            const uint32_t q = j * 32 + lane;
            have[j] = q < m;
            const uint32_t qq = have[j] ? q : 0u;
            p[j][0] = stage_a[3 * qq];
            p[j][1] = stage_a[3 * qq + 1];
            p[j][2] = stage_a[3 * qq + 2];
        }
#pragma unroll
        for (int j = 0; j < KS; ++j) {
            in[j] = have[j] && !outside<3>(tv, p[j]);
            lin[j] = linear_bin<3>(tv, p[j]); // clamped: harmless for points outside
        }
#pragma unroll
        for (int j = 0; j < KS; ++j) {
            uint32_t v;
            if (pg.sbin_n) { // kernel-uniform
                v = sbin[lin[j]];
                v = v == 0xFFFFu ? kNotLeaf : v;
            } else {
                v = __ldg(tv.bin_leaf + lin[j]);
            }
            leaf[j] = v;
        }
#pragma unroll
        for (int j = 0; j < KS; ++j) {
            if (!in[j])
                leaf[j] = kNotLeaf;
            else if (leaf[j] == kNotLeaf) // divided grid cell (none in trees whose fit stopped at the grid): descend
                leaf[j] = walk_to_leaf<3>(tv, lin[j], p[j]);
        }
        if (!keys_repeat(leaf[0], lane)) { // the usual case: four independent atomics in flight
#pragma unroll
            for (int j = 0; j < KS; ++j)
                rank[j] = leaf[j] != kNotLeaf ? atomicAdd(hist_a + leaf[j], 1u) : 0u;
        } else {
#pragma unroll
            for (int j = 0; j < KS; ++j)
                rank[j] = grouped_add(hist_a, leaf[j], true, lane);
        }
#pragma unroll
        for (int j = 0; j < KS; ++j) {
            const uint32_t q = j * 32 + lane;
            if (have[j]) {
                pg.key_a[a_pos + q] = leaf[j] == kNotLeaf ? kKeyNoLeaf : (leaf[j] | (rank[j] << kKeyLeafBits));
                if (leaf[j] == kNotLeaf)
                    pg.resa[a_pos + q] = __longlong_as_double(0x7ff8000000000000ll);
            }
        }
        a_pos += m;
    };
    // phase B: four packed keys per lane (one 16-byte load, requested before the tile); slot = tbl_b[leaf] + rank, no atomics
    auto b_request = [&]() -> uint4 {
        if (b_pos + 4 * lane < b_end) // ranges are multiples of 8 except the chunk's very end (the buffer is padded)
            return __ldcs(reinterpret_cast<const uint4 *>(pg.key_b + b_pos) + lane);
        return make_uint4(kKeyNoLeaf, kKeyNoLeaf, kKeyNoLeaf, kKeyNoLeaf);
    };
    auto b_consume = [&](const uint4 kk) {
        if (b_pos >= b_end)
            return;
        const uint32_t w[4] = {kk.x, kk.y, kk.z, kk.w};
        const uint32_t i0 = b_pos + 4 * lane;
        uint32_t slot[4];
        bool ok[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t leaf = w[j] & kKeyNoLeaf;
            ok[j] = leaf != kKeyNoLeaf && i0 + j < b_end;
            slot[j] = ok[j] ? tbl_b[leaf] + (w[j] >> kKeyLeafBits) : 0u;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (ok[j])
                pg.perm_b[slot[j]] = i0 + j;
        b_pos = min(b_pos + (uint32_t)kPiggySlice, b_end);
    };

    // ---- evaluation of chunk c: exactly the per-warp-run kernel above
    const uint32_t n_tiles = __ldg(n_tiles_p);
    const uint32_t nw = gridDim.x * WARPS, wid = blockIdx.x * WARPS + warp;
    uint32_t t = (uint32_t)(((uint64_t)n_tiles * wid) / nw);
    const uint32_t t_end = (uint32_t)(((uint64_t)n_tiles * (wid + 1)) / nw);
    auto load_meta = [&](uint32_t tt, uint2 &info, uint32_t (&id)[P]) {
        if (tt < t_end) {
            info = __ldg(tile_info + tt);
#pragma unroll
            for (int p = 0; p < P; ++p)
                id[p] = __ldg(perm + (size_t)tt * TILE + p * 32 + lane);
        } else {
            info = make_uint2(0, 0);
#pragma unroll
            for (int p = 0; p < P; ++p)
                id[p] = kInvalidSlotE;
        }
    };
    auto fetch_leaf = [&](const uint2 &info) {
        const double2 *Cg = reinterpret_cast<const double2 *>(tv.coeffs + 2ull * info.y);
#pragma unroll 4
        for (int j = lane; j < BLK2; j += 32)
            cp_async16(&rs.coeff[j], Cg + j);
        if (lane < 4)
            cp_async16(&rs.node[lane], reinterpret_cast<const double2 *>(tv.nodes + info.x) + lane);
    };
    auto issue = [&](uint32_t tt, const uint32_t (&id)[P], int sidx) {
        if (tt < t_end) {
#pragma unroll
            for (int p = 0; p < P; ++p) {
                rs.idx[sidx][p * 32 + lane] = id[p];
                if (id[p] != kInvalidSlotE) {
                    const double *q = x + 3ull * id[p];
                    cp_async8(&rs.pts[sidx][0 * TILE + p * 32 + lane], q);
                    cp_async8(&rs.pts[sidx][1 * TILE + p * 32 + lane], q + 1);
                    cp_async8(&rs.pts[sidx][2 * TILE + p * 32 + lane], q + 2);
                }
            }
        }
        cp_async_commit();
    };

    uint2 info_n, info_cur;
    uint32_t idx_n[P];
    load_meta(t, info_n, idx_n);
    issue(t, idx_n, 0);
    info_cur = info_n;
    load_meta(t + 1, info_n, idx_n);
    uint32_t cur_node = 0xFFFFFFFFu;
    int cur = 0;
    for (; t < t_end; ++t, cur ^= 1) {
        __syncwarp(); // every lane is done reading stage cur^1, the coefficient block and the side-job stage
        issue(t + 1, idx_n, cur ^ 1);
        const uint2 info_next = info_n;
        load_meta(t + 2, info_n, idx_n);
        a_request();                 // the slice's points: land while this tile computes
        const uint4 kk = b_request(); // ... and its keys
        if (info_cur.x != cur_node) { // warp uniform: a new leaf starts here
            cur_node = info_cur.x;
            fetch_leaf(info_cur);
            cp_async_commit();
            cp_async_wait<0>();
        } else {
            cp_async_wait<2>(); // this tile's points; the next tile's and the slice may still be in flight
        }
        __syncwarp();
        info_cur = info_next;
        eval_tile3d<O, P, TZFLY, TYKEEP>(rs.coeff, rs.node, rs.pts[cur], rs.idx[cur], lane, res);
        cp_async_wait<0>();
        __syncwarp();
        b_consume(kk);
        a_consume();
    }
    cp_async_wait<0>();
    // whatever the tiles did not cover (chunks with few tiles: most points outside the domain, a short last chunk)
    while (b_pos < b_end) {
        const uint4 kk = b_request();
        b_consume(kk);
    }
    while (a_pos < a_end) {
        __syncwarp();
        a_request();
        cp_async_wait<0>();
        __syncwarp();
        a_consume();
    }
    // ---- phase A epilogue: reserve this CTA's share of every leaf, save the bases for the launch that scatters
    if (pg.na) {
        __syncthreads();
        for (uint32_t lb = tid; lb < nl; lb += kSortBatch * WARPS * 32) {
            uint32_t h[kSortBatch], r[kSortBatch];
#pragma unroll
            for (int j = 0; j < kSortBatch; ++j) {
                const uint32_t l = lb + j * WARPS * 32;
                h[j] = l < nl ? hist_a[l] : 0u;
            }
#pragma unroll
            for (int j = 0; j < kSortBatch; ++j)
                r[j] = h[j] ? atomicAdd(pg.count_a + ((size_t)(lb + j * WARPS * 32) << pg.cshift), h[j]) : 0u;
#pragma unroll
            for (int j = 0; j < kSortBatch; ++j) {
                const uint32_t l = lb + j * WARPS * 32;
                if (l < nl)
                    pg.base_a[(size_t)blockIdx.x * nl + l] = r[j];
            }
        }
    }
}

// shared memory of the piggy kernel for a tree of nl leaves (sbin_n table entries)
template <int O, int P, int WARPS>
static size_t piggy_smem(uint32_t nl, uint32_t sbin_n) {
    return sizeof(RunSmem<O, P>) * WARPS + sizeof(uint32_t) * (2 * (size_t)nl + (nl & 1)) +
           sizeof(double) * 3 * kPiggySlice * WARPS + sizeof(uint16_t) * (((size_t)sbin_n + 7) / 8 * 8);
}

// ------------------------------------------------------------------------------------------ K4, CTA-shared block
// High orders (block = 13.8 KB at order 12, 22 KB at 14, 32 KB at 16; reference cheb_eval<3>, template.hpp:179-194):
// a private copy of the block per warp is what limited the kernel above to 4 warps per SM at order 16 -- one warp
// per scheduler.  A chunk holds hundreds to thousands of tiles per leaf at these orders, so here the whole CTA works
// in ONE leaf at a time: the CTA owns a contiguous run of tiles, its warps take them round-robin, and the leaf's block
// (plus its 64-byte node record) is staged ONCE per CTA by the TMA engine -- one elected thread arms an mbarrier with
// the byte count and issues two cp.async.bulk copies (SASS: UBLKCP), every thread then waits on the barrier's phase.
// A segment (= the CTA's tiles of one leaf) ends with one __syncthreads, after which the same buffer takes the next
// leaf; at <= 3 segments per CTA and chunk the exposed copy latency is noise, so there is no second buffer.
// Which tile starts the next segment is found by the warps themselves: each one reports its first tile outside the
// current leaf (shared atomicMin); tiles are ordered by leaf, so the minimum is the next segment's first tile.
template <int O, int P>
struct SharedWarpStage { // per warp: gathered points of two tiles in flight
    static constexpr int TILE = 32 * P;
    double pts[2][3 * TILE];
    uint32_t idx[2][TILE];
};
constexpr int kSharedHeader = 128; // node record (64 B) + mbarrier (8 B) + two segment cursors, behind the block

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
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

template <int O, int P, int WARPS, bool TZFLY, bool TYKEEP>
__global__ void __launch_bounds__(WARPS * 32, 1)
bin_eval3d_shared_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                         const uint32_t *__restrict__ perm, const uint2 *__restrict__ tile_info,
                         const uint32_t *__restrict__ n_tiles_p) {
    constexpr int TILE = 32 * P;
    constexpr uint32_t BLK_BYTES = O * O * O * 8;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2 *coeff_s = reinterpret_cast<double2 *>(smem_raw);
    double2 *node_s = reinterpret_cast<double2 *>(smem_raw + BLK_BYTES);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw + BLK_BYTES + 64);
    uint32_t *s_next = reinterpret_cast<uint32_t *>(smem_raw + BLK_BYTES + 72);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    SharedWarpStage<O, P> &st = reinterpret_cast<SharedWarpStage<O, P> *>(smem_raw + BLK_BYTES + kSharedHeader)[warp];
    const uint32_t n_tiles = __ldg(n_tiles_p);
    const uint32_t T0 = (uint32_t)(((uint64_t)n_tiles * blockIdx.x) / gridDim.x);
    const uint32_t T1 = (uint32_t)(((uint64_t)n_tiles * (blockIdx.x + 1)) / gridDim.x);
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        s_next[0] = T0;
        s_next[1] = 0xFFFFFFFFu;
    }
    __syncthreads();

    auto load_meta = [&](uint32_t tt, uint2 &info, uint32_t (&id)[P]) {
        if (tt < T1) {
            info = __ldg(tile_info + tt);
#pragma unroll
            for (int p = 0; p < P; ++p)
                id[p] = __ldg(perm + (size_t)tt * TILE + p * 32 + lane);
        } else {
            info = make_uint2(0, 0);
#pragma unroll
            for (int p = 0; p < P; ++p)
                id[p] = kInvalidSlotE;
        }
    };
    auto issue = [&](uint32_t tt, const uint32_t (&id)[P], int sidx) {
        if (tt < T1) {
#pragma unroll
            for (int p = 0; p < P; ++p) {
                st.idx[sidx][p * 32 + lane] = id[p];
                if (id[p] != kInvalidSlotE) {
                    const double *q = x + 3ull * id[p];
                    cp_async8(&st.pts[sidx][0 * TILE + p * 32 + lane], q);
                    cp_async8(&st.pts[sidx][1 * TILE + p * 32 + lane], q + 1);
                    cp_async8(&st.pts[sidx][2 * TILE + p * 32 + lane], q + 2);
                }
            }
        }
        cp_async_commit();
    };

    uint32_t t = T0 + warp; // this warp's tiles: T0 + warp, + WARPS, ...
    uint2 info_cur, info_n;
    uint32_t idx_n[P];
    load_meta(t, info_cur, idx_n);
    issue(t, idx_n, 0);
    load_meta(t + WARPS, info_n, idx_n);
    int cur = 0;
    for (uint32_t seg = 0;; ++seg) {
        const uint32_t first = s_next[seg & 1]; // CTA uniform: first tile of this segment
        if (first >= T1)
            break;
        const uint2 sinfo = __ldg(tile_info + first);
        if (threadIdx.x == 0) {
            s_next[(seg + 1) & 1] = 0xFFFFFFFFu; // nobody reports into it before the copy below has landed
            mbar_expect_tx(bar, BLK_BYTES + 64);
            tma_bulk_g2s(coeff_s, tv.coeffs + 2ull * sinfo.y, BLK_BYTES, bar);
            tma_bulk_g2s(node_s, tv.nodes + sinfo.x, 64, bar);
        }
        mbar_wait(bar, seg & 1);
        while (t < T1 && info_cur.x == sinfo.x) { // warp uniform
            __syncwarp(); // every lane is done reading stage cur^1
            issue(t + WARPS, idx_n, cur ^ 1);
            const uint2 info_next = info_n;
            load_meta(t + 2 * WARPS, info_n, idx_n); // lands while this tile computes
            cp_async_wait<1>();
            __syncwarp();
            eval_tile3d<O, P, TZFLY, TYKEEP>(coeff_s, node_s, st.pts[cur], st.idx[cur], lane, res);
            info_cur = info_next;
            t += WARPS;
            cur ^= 1;
        }
        if (lane == 0 && t < T1)
            atomicMin(&s_next[(seg + 1) & 1], t);
        __syncthreads(); // the block may be overwritten; the next segment's first tile is known
    }
    cp_async_wait<0>();
}

template <int O, int P, int WARPS, bool TZFLY, bool TYKEEP>
static int launch_k4_shared(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    auto kern = bin_eval3d_shared_kernel<O, P, WARPS, TZFLY, TYKEEP>;
    constexpr size_t smem = (size_t)O * O * O * 8 + kSharedHeader + sizeof(SharedWarpStage<O, P>) * WARPS;
    static_assert(smem <= 227 * 1024, "shared block + point stages exceed the shared memory of an SM");
    static bool configured[kMaxDevices] = {};
    if (dt.device < 0 || dt.device >= kMaxDevices || !configured[dt.device]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return (int)e;
        if (dt.device >= 0 && dt.device < kMaxDevices)
            configured[dt.device] = true;
    }
    kern<<<dt.sm_count, WARPS * 32, smem, st>>>(dt.view, d_x, d_res, ws.perm, ws.tile_info, ws.n_tiles);
    return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------ K4, 2D
// Same scheme for 2D trees: one warp per tile of 32*P slots of one leaf, the leaf's O x O block resident in the warp's
// shared memory for a whole run of tiles, every coefficient a warp-uniform load feeding P FMAs per lane.  Arithmetic of
// cheb_eval_2d (template.hpp:163-177): tmp = C * T(y) row by row in k order, value = <T(x), tmp> as the scalar halving
// tree; products with T_0 = 1 are not executed.  (Thread per point, every lane gathers its own block: 8e9 evals/s on the
// L2-resident C1 tree, bound by the L1 tag rate.)
template <int O, int P>
struct Run2Smem { // per warp
    static constexpr int TILE = 32 * P;
    double2 coeff[O * O / 2];
    double2 node[4];
    double pts[2][2 * TILE];
    uint32_t idx[2][TILE];
};

template <int O, int P, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
bin_eval2d_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                  const uint32_t *__restrict__ perm, const uint2 *__restrict__ tile_info,
                  const uint32_t *__restrict__ n_tiles_p) {
    constexpr int TILE = 32 * P;
    constexpr int BLK2 = O * O / 2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t n_tiles = __ldg(n_tiles_p);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Run2Smem<O, P> &rs = reinterpret_cast<Run2Smem<O, P> *>(smem_raw)[warp];
    const uint32_t nw = gridDim.x * WARPS, wid = blockIdx.x * WARPS + warp;
    uint32_t t = (uint32_t)(((uint64_t)n_tiles * wid) / nw);
    const uint32_t t_end = (uint32_t)(((uint64_t)n_tiles * (wid + 1)) / nw);

    auto load_meta = [&](uint32_t tt, uint2 &info, uint32_t (&id)[P]) {
        if (tt < t_end) {
            info = __ldg(tile_info + tt);
#pragma unroll
            for (int p = 0; p < P; ++p)
                id[p] = __ldg(perm + (size_t)tt * TILE + p * 32 + lane);
        } else {
            info = make_uint2(0, 0);
#pragma unroll
            for (int p = 0; p < P; ++p)
                id[p] = kInvalidSlotE;
        }
    };
    auto issue = [&](uint32_t tt, const uint32_t (&id)[P], int sidx) {
        if (tt < t_end) {
#pragma unroll
            for (int p = 0; p < P; ++p) {
                rs.idx[sidx][p * 32 + lane] = id[p];
                if (id[p] != kInvalidSlotE) // a 2D point is one aligned 16-byte item
                    cp_async16(&rs.pts[sidx][2 * (p * 32 + lane)], x + 2ull * id[p]);
            }
        }
        cp_async_commit();
    };

    uint2 info_n, info_cur;
    uint32_t idx_n[P];
    load_meta(t, info_n, idx_n);
    issue(t, idx_n, 0);
    info_cur = info_n;
    load_meta(t + 1, info_n, idx_n);
    uint32_t cur_node = 0xFFFFFFFFu;
    int cur = 0;
    for (; t < t_end; ++t, cur ^= 1) {
        __syncwarp();
        issue(t + 1, idx_n, cur ^ 1);
        const uint2 info_next = info_n;
        load_meta(t + 2, info_n, idx_n);
        if (info_cur.x != cur_node) { // warp uniform: a new leaf starts here
            cur_node = info_cur.x;
            const double2 *Cg = reinterpret_cast<const double2 *>(tv.coeffs + 2ull * info_cur.y);
            for (int j = lane; j < BLK2; j += 32)
                cp_async16(&rs.coeff[j], Cg + j);
            if (lane < 4)
                cp_async16(&rs.node[lane], reinterpret_cast<const double2 *>(tv.nodes + info_cur.x) + lane);
            cp_async_commit();
            cp_async_wait<0>();
        } else {
            cp_async_wait<1>();
        }
        __syncwarp();
        info_cur = info_next;
        const double c0 = rs.node[0].x, c1 = rs.node[0].y, ih0 = rs.node[2].x, ih1 = rs.node[2].y;
        double Ty[P][O], u[P], acc[P][O];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const bool ok = rs.idx[cur][p * 32 + lane] != kInvalidSlotE;
            const double2 q = *reinterpret_cast<const double2 *>(&rs.pts[cur][2 * (p * 32 + lane)]);
            const double px = ok ? q.x : c0, py = ok ? q.y : c1; // padding lanes evaluate the leaf centre
            u[p] = __dmul_rn(__dsub_rn(px, c0), ih0);
            cheb_T<O>(__dmul_rn(__dsub_rn(py, c1), ih1), Ty[p]);
        }
#pragma unroll
        for (int k = 1; k < O; ++k) {
#pragma unroll
            for (int r = 0; r < O; r += 2) {
                const double2 c = rs.coeff[(k * O + r) / 2];
                if (k == 1) {
                    const double2 cz = rs.coeff[r / 2];
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        acc[p][r] = __fma_rn(c.x, Ty[p][1], cz.x);
                        acc[p][r + 1] = __fma_rn(c.y, Ty[p][1], cz.y);
                    }
                } else {
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        acc[p][r] = __fma_rn(c.x, Ty[p][k], acc[p][r]);
                        acc[p][r + 1] = __fma_rn(c.y, Ty[p][k], acc[p][r + 1]);
                    }
                }
            }
        }
#pragma unroll
        for (int p = 0; p < P; ++p) {
            double Tx[O];
            cheb_T<O>(u[p], Tx);
            const double val = dot_tree_one<0, O, 1, 0, O>(Tx, acc[p]);
            const uint32_t id = rs.idx[cur][p * 32 + lane];
            if (id != kInvalidSlotE)
                res[id] = val;
        }
    }
    cp_async_wait<0>();
}

template <int O, int P, int WARPS>
static int launch_k4_2d(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    auto kern = bin_eval2d_kernel<O, P, WARPS>;
    constexpr size_t smem = sizeof(Run2Smem<O, P>) * WARPS;
    static bool configured[kMaxDevices] = {}; // function attributes are per device
    if (dt.device < 0 || dt.device >= kMaxDevices || !configured[dt.device]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return (int)e;
        if (dt.device >= 0 && dt.device < kMaxDevices)
            configured[dt.device] = true;
    }
    int per_sm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS * 32, smem);
    if (per_sm < 1)
        per_sm = 1;
    kern<<<dt.sm_count * per_sm, WARPS * 32, smem, st>>>(dt.view, d_x, d_res, ws.perm, ws.tile_info, ws.n_tiles);
    return (int)cudaGetLastError();
}

// points per lane of the 2D kernel for this order
int bin_eval2d_P(int order) { return order <= 10 ? 4 : 2; }

int launch_bin_eval2d(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    switch (dt.view.order) {
    case 6: return launch_k4_2d<6, 4, 12>(dt, ws, d_x, d_res, st);
    case 8: return launch_k4_2d<8, 4, 12>(dt, ws, d_x, d_res, st);
    case 10: return launch_k4_2d<10, 4, 12>(dt, ws, d_x, d_res, st);
    case 12: return launch_k4_2d<12, 2, 12>(dt, ws, d_x, d_res, st);
    case 14: return launch_k4_2d<14, 2, 12>(dt, ws, d_x, d_res, st);
    case 16: return launch_k4_2d<16, 2, 12>(dt, ws, d_x, d_res, st);
    default: return (int)cudaErrorInvalidValue;
    }
}

// ------------------------------------------------------------------------------------------ K4, FP64 tensor cores
// Order 8 on the DMMA path.  Measured on B200 (scripts/micro/): a DFMA stream with two fresh register operands per
// instruction (the register-blocked loop above) tops out at ~77 % of the FP64 peak once its coefficients come from
// shared memory; mma.sync.m8n8k4.f64 runs at the same peak as DFMA (36.9 TFLOP/s), reads a quarter of the registers
// per FMA and takes 1/8 of the issue slots -- and it rounds exactly like the chain it replaces: for every output,
//     d = c;  d = fma(a[m][k], b[k][n], d)  for k = 0, 1, 2, 3          (scripts/micro/dmma_order.cu: 0 of 12.8 M
// outputs differ, signed zeros and heavy cancellation included).  That is the reference's row chain
// acc[r] = fma(C_i(r,k), Tz[k], acc[r]), so the result stays bit-identical.
//
// Mapping (slab i of the leaf, a group of 8 points):  D[point][n] = sum_k Tz[point][k] * C_i(row(n), k), two chained
// MMAs (k = 0..3, 4..7) starting from C = -0.0 (fma(1, c, -0) == c: the reference's c * T_0).
//   A fragment  lane (g, t) holds Tz[point g][k = t (+4)]            -- registers, constant over the slabs
//   B fragment  lane (g, t) holds C_i(row(g), k = t (+4))            -- the WHOLE leaf block is 16 doubles per lane,
//               loaded once per leaf with coalesced 256-byte requests and kept in registers: a warp walks a
//               contiguous run of tiles, and tiles of one leaf are adjacent, so no shared-memory coefficient traffic
//   D fragment  lane (g, t) holds columns 2t, 2t+1; the column -> row map is row(2t) = t, row(2t+1) = t + 4, which
//               are exactly the two rows the reference's packet reduction pairs first:
//               l_t = fma(Ty[t+4], acc[t+4], Ty[t] * acc[t]) is lane local, then d = (l0 + l2) + (l1 + l3) takes two
//               quad shuffles.  Four groups (32 points) are reduced together so that every lane ends up owning one point.
struct MmaStage { // per warp, per stage
    double pts[3 * 128]; // [coordinate][slot]
    uint32_t idx[128];
};

__device__ __forceinline__ void dmma8(double &d0, double &d1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
// T[t] and T[t + 4] for t = lane & 3
__device__ __forceinline__ void pick_t(const double (&T)[8], uint32_t t, double &lo, double &hi) {
    const bool b0 = t & 1, b1 = t & 2;
    const double l01 = b0 ? T[1] : T[0], l23 = b0 ? T[3] : T[2];
    const double h01 = b0 ? T[5] : T[4], h23 = b0 ? T[7] : T[6];
    lo = b1 ? l23 : l01;
    hi = b1 ? h23 : h01;
}
__device__ __forceinline__ double shfl_xor_d(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }

template <int WARPS, bool DEFER>
__global__ void __launch_bounds__(WARPS * 32, 1)
bin_eval3d_mma8_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                       const uint32_t *__restrict__ perm, const uint2 *__restrict__ tile_info,
                       const uint32_t *__restrict__ n_tiles_p) {
    constexpr int TILE = 128;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t n_tiles = __ldg(n_tiles_p);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    MmaStage *stages = reinterpret_cast<MmaStage *>(smem_raw) + 2 * warp;
    // contiguous run of tiles per warp: a leaf's tiles are adjacent, so its coefficients are fetched once per run
    const uint32_t nw = gridDim.x * WARPS, wid = blockIdx.x * WARPS + warp;
    const uint32_t t_begin = (uint32_t)(((uint64_t)n_tiles * wid) / nw), t_end = (uint32_t)(((uint64_t)n_tiles * (wid + 1)) / nw);

    auto load_meta = [&](uint32_t tt, uint32_t (&id)[4]) {
#pragma unroll
        for (int p = 0; p < 4; ++p)
            id[p] = tt < t_end ? __ldg(perm + (size_t)tt * TILE + p * 32 + lane) : kInvalidSlotE;
    };
    auto issue = [&](const uint32_t (&id)[4], MmaStage &st) {
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            st.idx[p * 32 + lane] = id[p];
            if (id[p] != kInvalidSlotE) {
                const double *q = x + 3ull * id[p];
                cp_async8(&st.pts[0 * TILE + p * 32 + lane], q);
                cp_async8(&st.pts[1 * TILE + p * 32 + lane], q + 1);
                cp_async8(&st.pts[2 * TILE + p * 32 + lane], q + 2);
            }
        }
        cp_async_commit();
    };

    uint32_t idx_n[4];
    load_meta(t_begin, idx_n);
    issue(idx_n, stages[0]);
    load_meta(t_begin + 1, idx_n);
    uint32_t cur_node = 0xFFFFFFFFu;
    double breg[8][2]; // B fragments of the current leaf
    double c0 = 0, c1 = 0, c2 = 0, ih0 = 0, ih1 = 0, ih2 = 0;
    const uint32_t brow = (g >> 1) + 4 * (g & 1); // row(n = g)
    int cur = 0;
    for (uint32_t tt = t_begin; tt < t_end; ++tt, cur ^= 1) {
        __syncwarp(); // every lane is done reading stage cur^1
        issue(idx_n, stages[cur ^ 1]);
        load_meta(tt + 2, idx_n);
        const uint2 info = __ldg(tile_info + tt);
        if (info.x != cur_node) { // warp uniform
            cur_node = info.x;
            const double *C = tv.coeffs + 2ull * info.y;
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int kb = 0; kb < 2; ++kb)
                    breg[i][kb] = __ldg(C + i * 64 + (4 * kb + t) * 8 + brow);
            const NodeHead head = load_head(tv.nodes, info.x);
            const NodeTail tail = load_tail(tv.nodes, info.x);
            c0 = head.c0, c1 = head.c1, c2 = head.c2;
            ih0 = tail.ih0, ih1 = tail.ih1, ih2 = tail.ih2;
        }
        cp_async_wait<1>();
        __syncwarp();
        const MmaStage &st = stages[cur];
#pragma unroll 1
        for (int s = 0; s < 4; ++s) {
            const int base = s * 32;
            double a0[4], a1[4], ya[4], yb[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) { // the four points (group j, row g) this lane's fragments belong to
                const int q = base + 8 * j + g;
                const double v = __dmul_rn(__dsub_rn(st.pts[1 * TILE + q], c1), ih1);
                const double w = __dmul_rn(__dsub_rn(st.pts[2 * TILE + q], c2), ih2);
                double T[8];
                cheb_T<8>(w, T);
                pick_t(T, t, a0[j], a1[j]);
                cheb_T<8>(v, T);
                pick_t(T, t, ya[j], yb[j]);
            }
            const int own = base + 8 * t + g; // this lane's own point: group t
            const double u = __dmul_rn(__dsub_rn(st.pts[0 * TILE + own], c0), ih0);
            const double x2 = __dadd_rn(u, u);
            double tx = 1.0, tx_prev = u, val = 0.0;
            const bool low = t < 2, odd = t & 1;
            // DEFER: the slab loop only produces the lane-local pair sums l (independent two-deep chains that trail
            // the DMMAs); the quad reduction of all 8 slabs is done afterwards in wide, independent batches, so that
            // no short dependent FP64 chain sits between the DMMAs of consecutive slabs
            double keep0[DEFER ? 8 : 1], keep1[DEFER ? 8 : 1], send0[DEFER ? 8 : 1], send1[DEFER ? 8 : 1];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                double l[4], d0[4], d1[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    d0[j] = d1[j] = -0.0;
                    dmma8(d0[j], d1[j], a0[j], breg[i][0]);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    dmma8(d0[j], d1[j], a1[j], breg[i][1]);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    l[j] = __dmul_rn(ya[j], d0[j]);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    l[j] = __fma_rn(yb[j], d1[j], l[j]); // rows t and t + 4 of point (j, g)
                // (l0 + l2) + (l1 + l3) over the quad; groups 0,1 are finished on lanes t < 2, groups 2,3 on t >= 2
                if constexpr (DEFER) {
                    keep0[i] = low ? l[0] : l[2], keep1[i] = low ? l[1] : l[3];
                    send0[i] = low ? l[2] : l[0], send1[i] = low ? l[3] : l[1];
                } else {
                    const double r0 = shfl_xor_d(low ? l[2] : l[0], 2), r1 = shfl_xor_d(low ? l[3] : l[1], 2);
                    const double s0 = __dadd_rn(low ? l[0] : l[2], r0), s1 = __dadd_rn(low ? l[1] : l[3], r1);
                    const double d = __dadd_rn(odd ? s1 : s0, shfl_xor_d(odd ? s0 : s1, 1)); // lane t: point (t, g)
                    val = __fma_rn(tx, d, val);
                    const double tn = __fma_rn(x2, tx, -tx_prev);
                    tx_prev = tx;
                    tx = tn;
                }
            }
            if constexpr (DEFER) {
                double s0[8], s1[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    s0[i] = __dadd_rn(keep0[i], shfl_xor_d(send0[i], 2));
                    s1[i] = __dadd_rn(keep1[i], shfl_xor_d(send1[i], 2));
                }
                double d[8];
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    d[i] = __dadd_rn(odd ? s1[i] : s0[i], shfl_xor_d(odd ? s0[i] : s1[i], 1));
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    val = __fma_rn(tx, d[i], val);
                    const double tn = __fma_rn(x2, tx, -tx_prev);
                    tx_prev = tx;
                    tx = tn;
                }
            }
            const uint32_t id = st.idx[own];
            if (id != kInvalidSlotE)
                res[id] = val;
        }
    }
    cp_async_wait<0>();
}

template <int WARPS, bool DEFER>
static int launch_k4_mma8(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    auto kern = bin_eval3d_mma8_kernel<WARPS, DEFER>;
    constexpr size_t smem = sizeof(MmaStage) * 2 * WARPS;
    static bool configured[kMaxDevices] = {}; // function attributes are per device
    if (dt.device < 0 || dt.device >= kMaxDevices || !configured[dt.device]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return (int)e;
        if (dt.device >= 0 && dt.device < kMaxDevices)
            configured[dt.device] = true;
    }
    kern<<<dt.sm_count, WARPS * 32, smem, st>>>(dt.view, d_x, d_res, ws.perm, ws.tile_info, ws.n_tiles);
    return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------ host side
template <int O, int P, int WARPS, bool TZFLY, bool TYKEEP>
static int launch_k4_piggy(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st,
                           const PiggyArgs &pg) {
    auto kern = bin_eval3d_piggy_kernel<O, P, WARPS, TZFLY, TYKEEP>;
    const size_t smem = piggy_smem<O, P, WARPS>(dt.view.n_leaves, pg.sbin_n);
    if (smem > 227 * 1024)
        return (int)cudaErrorInvalidConfiguration;
    static size_t configured[kMaxDevices] = {}; // largest size asked for so far, per device
    if (dt.device < 0 || dt.device >= kMaxDevices || configured[dt.device] < smem) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return (int)e;
        if (dt.device >= 0 && dt.device < kMaxDevices)
            configured[dt.device] = smem;
    }
    kern<<<dt.sm_count, WARPS * 32, smem, st>>>(dt.view, d_x, d_res, ws.perm, ws.tile_info, ws.n_tiles, pg);
    return (int)cudaGetLastError();
}

// The piggy kernel exists for the default variant of every order (what bin_eval_select returns without an override).
#define BZ_PIGGY_CASES(X) X(6, 4, 8, false, true) X(8, 4, 8, false, true) X(10, 3, 8, false, false) X(12, 3, 8, false, false) \
    X(14, 2, 8, false, true) X(16, 2, 4, false, true)

int bin_eval_piggy_P(int order) { // points per lane of the piggy kernel: the workspace's tile size must be 32 * this
    switch (order) {
#define X(O_, P_, W_, F_, Y_) case O_: return P_;
        BZ_PIGGY_CASES(X)
#undef X
    default: return -1;
    }
}

size_t bin_eval_piggy_smem(int order, uint32_t n_leaves, uint32_t sbin_n) {
    switch (order) {
#define X(O_, P_, W_, F_, Y_) case O_: return piggy_smem<O_, P_, W_>(n_leaves, sbin_n);
        BZ_PIGGY_CASES(X)
#undef X
    default: return ~size_t(0);
    }
}

int launch_bin_eval3d_piggy(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st,
                            const PiggyArgs &pg) {
    switch (dt.view.order) {
#define X(O_, P_, W_, F_, Y_) case O_: return launch_k4_piggy<O_, P_, W_, F_, Y_>(dt, ws, d_x, d_res, st, pg);
        BZ_PIGGY_CASES(X)
#undef X
    default: return (int)cudaErrorInvalidValue;
    }
}

struct K4Variant {
    int order, P, warps;
    bool tzfly, tykeep;
    int run; // 0: tiles strided over the grid, block per tile; 1: per-warp runs, block per warp; 2: CTA-shared block (TMA)
    int (*launch)(const DeviceTree &, const BinWorkspace &, const double *, double *, cudaStream_t);
    const void *func; // for cudaFuncGetAttributes
    size_t smem;
};

template <int O, int P, int WARPS, bool TZFLY, bool TYKEEP, bool RUN>
static int launch_k4(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    auto kern = bin_eval3d_kernel<O, P, WARPS, TZFLY, TYKEEP, RUN>;
    constexpr size_t smem = RUN ? sizeof(RunSmem<O, P>) * WARPS : sizeof(PipeStage<O, P>) * 2 * WARPS;
    static_assert(smem <= 227 * 1024, "pipeline stages exceed the shared memory of an SM");
    static bool configured[kMaxDevices] = {}; // per instantiation and device (function attributes are per device)
    if (dt.device < 0 || dt.device >= kMaxDevices || !configured[dt.device]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return (int)e;
        if (dt.device >= 0 && dt.device < kMaxDevices)
            configured[dt.device] = true;
    }
    kern<<<dt.sm_count, WARPS * 32, smem, st>>>(dt.view, d_x, d_res, ws.perm, ws.tile_info, ws.n_tiles);
    return (int)cudaGetLastError();
}

#define BZ_K4(O_, P_, W_, F_, Y_, R_)                                                                                  \
    {O_, P_, W_, F_, Y_, R_, launch_k4<O_, P_, W_, F_, Y_, R_>, (const void *)bin_eval3d_kernel<O_, P_, W_, F_, Y_, R_>, \
     R_ ? sizeof(RunSmem<O_, P_>) * W_ : sizeof(PipeStage<O_, P_>) * 2 * W_}
#define BZ_K4S(O_, P_, W_, F_, Y_)                                                                                     \
    {O_, P_, W_, F_, Y_, 2, launch_k4_shared<O_, P_, W_, F_, Y_>, (const void *)bin_eval3d_shared_kernel<O_, P_, W_, F_, Y_>, \
     (size_t)O_ * O_ * O_ * 8 + kSharedHeader + sizeof(SharedWarpStage<O_, P_>) * W_}
#define BZ_K4M(W_, D_)                                                                                                 \
    {8, D_ ? 5 : 4, W_, true, true, 0, launch_k4_mma8<W_, D_>, (const void *)bin_eval3d_mma8_kernel<W_, D_>,       \
     sizeof(MmaStage) * 2 * W_}
// first entry of an order = its default
static const K4Variant kVariants[] = {
    BZ_K4(6, 4, 8, false, true, true),    BZ_K4(6, 4, 8, false, true, false),   BZ_K4(6, 4, 12, false, true, true),
    BZ_K4(8, 4, 8, false, true, true),    BZ_K4(8, 4, 8, false, true, false),   BZ_K4(8, 4, 8, false, false, false),
    BZ_K4(8, 3, 12, false, false, false), BZ_K4(8, 3, 12, false, false, true),  BZ_K4(8, 2, 16, false, true, false),
    BZ_K4(8, 4, 8, true, true, true),     BZ_K4(8, 4, 12, true, false, true),   BZ_K4(8, 3, 12, true, true, true),
    // tzfly && tykeep together = the DMMA kernel (P = 5 in the selector: deferred quad reduction)
    BZ_K4M(16, false), BZ_K4M(12, false), BZ_K4M(16, true),
    BZ_K4(10, 3, 8, false, false, true),  BZ_K4(10, 2, 8, false, true, false),  BZ_K4(10, 3, 8, false, false, false),
    BZ_K4(10, 3, 12, false, false, true),
    BZ_K4(12, 3, 8, false, false, true),  BZ_K4(12, 2, 4, false, true, false),  BZ_K4(12, 3, 4, false, false, false),
    BZ_K4(12, 2, 8, false, true, true),   BZ_K4(12, 2, 12, false, false, true),
    // orders 14 and 16 default to the CTA-shared, TMA-staged block (run = 2): measured on B200 with the order-12 benchmark
    // function at these orders, 1e8 points (profiles/r02_sweep.txt): order 14 0.727 of the FP64 peak with P = 3, T(z) on the
    // fly (4.22e9 evals/s; per-warp copies, 8 warps: 3.9e9), order 16 0.727 with P = 2, 8 warps (2.95e9; per-warp copies
    // allow 4 warps: 2.70e9).  Order 12 gains nothing (0.723 against 0.727) and keeps its per-warp copies.
    BZ_K4S(14, 3, 8, true, false),   BZ_K4S(14, 2, 8, false, true),   BZ_K4S(14, 2, 12, true, false),
    BZ_K4(14, 2, 8, false, true, true),   BZ_K4(14, 2, 4, false, true, false),  BZ_K4(14, 3, 8, true, false, true),
    BZ_K4S(16, 2, 8, false, true),   BZ_K4S(16, 2, 8, false, false),  BZ_K4S(16, 2, 12, true, false),
    BZ_K4(16, 2, 4, false, true, true),   BZ_K4(16, 2, 2, false, true, false),  BZ_K4(16, 2, 6, false, false, true),
    BZ_K4S(12, 3, 8, false, false),  BZ_K4S(12, 3, 12, true, false),  BZ_K4S(12, 2, 12, false, true),
};
#undef BZ_K4S
#undef BZ_K4
#undef BZ_K4M

static const K4Variant *find_variant(int order, int P, int warps, int tzfly, int tykeep, int run) {
    for (const K4Variant &v : kVariants)
        if (v.order == order && (P < 0 || (v.P == P && v.warps == warps && v.tzfly == (tzfly != 0) &&
                                           v.tykeep == (tykeep != 0) && v.run == run)))
            return &v;
    return nullptr;
}

// Variant for a 3D tree of this order: the default, or BAOBZI_B200_K4="P,warps,tzfly,tykeep[,run]" (benchmarks).
// Returns the points per lane (tile = 32 * P slots) and an opaque index for launch_bin_eval3d; -1 if unsupported.
bool bin_eval_is_default(int variant) {
    if (variant < 0 || variant >= (int)(sizeof(kVariants) / sizeof(kVariants[0])))
        return false;
    return &kVariants[variant] == find_variant(kVariants[variant].order, -1, 0, 0, 0, 0);
}

int bin_eval_select(int order, int *P) {
    const K4Variant *v = find_variant(order, -1, 0, 0, 0, 0);
    if (const char *e = std::getenv("BAOBZI_B200_K4")) {
        int p = 0, w = 0, f = 0, y = 0, r = 0;
        if (std::sscanf(e, "%d,%d,%d,%d,%d", &p, &w, &f, &y, &r) >= 4) {
            const K4Variant *o = find_variant(order, p, w, f, y, r);
            if (o)
                v = o;
            else
                std::fprintf(stderr, "baobzi: BAOBZI_B200_K4=%s is not built for order %d, using the default\n", e, order);
        }
    }
    if (!v)
        return -1;
    *P = v->P == 5 ? 4 : v->P;
    return (int)(v - kVariants);
}

// Registers (whole CTA) and shared memory the selected kernel holds on an SM: what is left is the room a
// concurrently running sort CTA may use.  Returns false when unknown.
bool bin_eval_footprint(int variant, int *cta_regs, size_t *cta_smem) {
    if (variant < 0 || variant >= (int)(sizeof(kVariants) / sizeof(kVariants[0])) || !kVariants[variant].func)
        return false;
    cudaFuncAttributes a;
    if (cudaFuncGetAttributes(&a, kVariants[variant].func) != cudaSuccess)
        return false;
    const int per_warp = (a.numRegs * 32 + 511) / 512 * 512; // allocation granularity: 512 registers per warp
    *cta_regs = per_warp * kVariants[variant].warps;
    *cta_smem = kVariants[variant].smem + a.sharedSizeBytes + 1024;
    return true;
}

int launch_bin_eval3d(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    { // experiment knob, read per call (tests flip it); the device copy of the flag is updated when it changes
        static int on_device[64] = {};
        const char *ev = std::getenv("BAOBZI_B200_K4_EVICT");
        const int evict = ev && std::atoi(ev) != 0 ? 1 : 0;
        const int d = dt.device >= 0 && dt.device < 64 ? dt.device : 0;
        if (evict != on_device[d]) {
            static int values[2] = {0, 1};
            cudaError_t e = cudaMemcpyToSymbolAsync(g_k4_evict_first, &values[evict], sizeof(int), 0, cudaMemcpyHostToDevice, st);
            if (e != cudaSuccess)
                return (int)e;
            on_device[d] = evict;
        }
    }
    if (ws.variant < 0 || ws.variant >= (int)(sizeof(kVariants) / sizeof(kVariants[0])) ||
        kVariants[ws.variant].order != dt.view.order)
        return (int)cudaErrorInvalidValue;
    return kVariants[ws.variant].launch(dt, ws, d_x, d_res, st);
}

} // namespace bz
