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
// tile j+1's inputs (the leaf's coefficient block, the node record, the gathered points and their indices) are
// in flight as cp.async copies, and tile j+2's descriptor is being fetched into registers.
//
// Arithmetic: cheb.cuh's order (bit-identical to the reference's AVX2+FMA build).  Multiplications by
// T_0 = 1.0 are not executed (x * 1.0 == x bit for bit): a row chain starts as fma(c_1, T_1, c_0), and the
// T recurrences run from T_{-1} = T_1 = x so that every slab executes the same code (fma(2x, 1, -x) == x exactly).
#include "cheb.cuh"
#include "device_tree.hpp"
#include "lookup.cuh"

#include <cstdio>
#include <cstdlib>

namespace bz {

constexpr uint32_t kInvalidSlotE = 0xFFFFFFFFu;

__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
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

// TZFLY: T_k(z) advanced inside the k loop (costs O-2 extra FMAs per point and slab, frees (O-3)*P doubles).
// TYKEEP: T(y) computed once per tile and kept live (O*P doubles) instead of rebuilt after every slab.
template <int O, int P, int WARPS, bool TZFLY, bool TYKEEP, int EXP = 0>
__global__ void __launch_bounds__(WARPS * 32, 1)
bin_eval3d_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                  const uint32_t *__restrict__ perm, const uint2 *__restrict__ tile_info,
                  const uint32_t *__restrict__ n_tiles_p) {
    constexpr int TILE = 32 * P;
    constexpr int BLK2 = O * O * O / 2;
    constexpr int NTZ = TZFLY ? 1 : O;
    using Stage = PipeStage<O, P>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t n_tiles = __ldg(n_tiles_p);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Stage *stages = reinterpret_cast<Stage *>(smem_raw) + 2 * warp;
    const uint32_t stride = gridDim.x * WARPS;
    uint32_t t = blockIdx.x * WARPS + warp;

    // descriptor of a tile: {node, coeff16} + the P slot indices of this lane
    auto load_meta = [&](uint32_t tt, uint2 &info, uint32_t (&id)[P]) {
        if (tt < n_tiles) {
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
    auto issue = [&](uint32_t tt, const uint2 &info, const uint32_t (&id)[P], Stage &st) {
        if (tt < n_tiles && !(EXP == 2 && tt >= blockIdx.x * WARPS + warp + 2 * stride)) {
            const double2 *Cg = reinterpret_cast<const double2 *>(tv.coeffs + 2ull * info.y);
#pragma unroll 4
            for (int j = lane; j < BLK2; j += 32)
                cp_async16(&st.coeff[j], Cg + j);
            if (lane < 4)
                cp_async16(&st.node[lane], reinterpret_cast<const double2 *>(tv.nodes + info.x) + lane);
#pragma unroll
            for (int p = 0; p < P; ++p) {
                st.idx[p * 32 + lane] = id[p];
                if (id[p] != kInvalidSlotE) {
                    const double *q = x + 3ull * id[p];
                    cp_async8(&st.pts[0 * TILE + p * 32 + lane], q);
                    cp_async8(&st.pts[1 * TILE + p * 32 + lane], q + 1);
                    cp_async8(&st.pts[2 * TILE + p * 32 + lane], q + 2);
                }
            }
        }
        cp_async_commit();
    };

    uint2 info_n;
    uint32_t idx_n[P];
    load_meta(t, info_n, idx_n);
    issue(t, info_n, idx_n, stages[0]);
    load_meta(t + stride, info_n, idx_n); // descriptor of the second tile
    int cur = 0;
    for (; t < n_tiles; t += stride, cur ^= 1) {
        __syncwarp(); // every lane is done reading stage cur^1 (tile j-1)
        issue(t + stride, info_n, idx_n, stages[cur ^ 1]);
        load_meta(t + 2 * stride, info_n, idx_n); // lands while this tile computes
        cp_async_wait<1>();
        __syncwarp();
        const Stage &st = stages[cur];

        const double2 n0 = st.node[0], n1 = st.node[1], n2 = st.node[2], n3 = st.node[3];
        const double c0 = n0.x, c1 = n0.y, c2c = n1.x, ih0 = n2.x, ih1 = n2.y, ih2 = n3.x;
        double Tz[P][NTZ], Tyk[P][TYKEEP ? O : 1], acc[P][O];
        double v[P], w[P], x2[P], w2[P], tx[P], tx_prev[P], val[P];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const bool ok = st.idx[p * 32 + lane] != kInvalidSlotE;
            const double px = ok ? st.pts[0 * TILE + p * 32 + lane] : c0; // padding lanes evaluate the leaf centre
            const double py = ok ? st.pts[1 * TILE + p * 32 + lane] : c1;
            const double pz = ok ? st.pts[2 * TILE + p * 32 + lane] : c2c;
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
            const double2 *cc = st.coeff + i * (O * O / 2);
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
                    double2 c = cc[(k * O + r) / 2];
                    if (EXP == 1) { // timing experiment: no shared-memory traffic in the slab loop
                        c = make_double2(x2[0], x2[1]);
                        asm volatile("" : "+d"(c.x), "+d"(c.y));
                    }
                    if (k == 1) {
                        double2 cz = cc[r / 2];
                        if (EXP == 1) {
                            cz = make_double2(x2[1], x2[0]);
                            asm volatile("" : "+d"(cz.x), "+d"(cz.y));
                        }
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
            const uint32_t id = st.idx[p * 32 + lane];
            if (id != kInvalidSlotE)
                res[id] = val[p];
        }
    }
    cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------ host side
struct K4Variant {
    int order, P, warps;
    bool tzfly, tykeep;
    int (*launch)(const DeviceTree &, const BinWorkspace &, const double *, double *, cudaStream_t);
};

template <int O, int P, int WARPS, bool TZFLY, bool TYKEEP, int EXP = 0>
static int launch_k4(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    auto kern = bin_eval3d_kernel<O, P, WARPS, TZFLY, TYKEEP, EXP>;
    constexpr size_t smem = sizeof(PipeStage<O, P>) * 2 * WARPS;
    static_assert(smem <= 227 * 1024, "pipeline stages exceed the shared memory of an SM");
    static bool configured = false; // per instantiation
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return (int)e;
        configured = true;
    }
    kern<<<dt.sm_count, WARPS * 32, smem, st>>>(dt.view, d_x, d_res, ws.perm, ws.tile_info, ws.n_tiles);
    return (int)cudaGetLastError();
}

#define BZ_K4(O_, P_, W_, F_, Y_) {O_, P_, W_, F_, Y_, launch_k4<O_, P_, W_, F_, Y_>}
// first entry of an order = its default
static const K4Variant kVariants[] = {
    BZ_K4(6, 4, 8, false, true),   BZ_K4(6, 4, 12, false, true),  BZ_K4(6, 4, 16, false, false),
    BZ_K4(8, 4, 8, false, true),   BZ_K4(8, 4, 8, false, false),  BZ_K4(8, 3, 12, false, false),
    BZ_K4(8, 3, 12, false, true),  BZ_K4(8, 4, 12, true, false),  BZ_K4(8, 2, 16, false, false),
    BZ_K4(8, 2, 16, false, true),  BZ_K4(8, 3, 8, false, true),
    BZ_K4(10, 2, 8, false, true),  BZ_K4(10, 3, 8, false, false), BZ_K4(10, 4, 8, true, false),
    BZ_K4(12, 2, 4, false, true),  BZ_K4(12, 3, 4, false, false), BZ_K4(12, 4, 4, true, false),
    BZ_K4(14, 2, 4, false, true),  BZ_K4(14, 3, 4, true, false),
    BZ_K4(16, 2, 2, false, true),  BZ_K4(16, 2, 3, false, false),
#ifdef BZ_EXPERIMENTS // timing experiments with WRONG results: P = 91 / 92 select them
    {8, 91, 8, false, true, launch_k4<8, 4, 8, false, true, 1>}, {8, 92, 8, false, true, launch_k4<8, 4, 8, false, true, 2>},
    {8, 91, 12, false, false, launch_k4<8, 3, 12, false, false, 1>}, {8, 92, 12, false, false, launch_k4<8, 3, 12, false, false, 2>},
#endif
};
#undef BZ_K4

static const K4Variant *find_variant(int order, int P, int warps, int tzfly, int tykeep) {
    for (const K4Variant &v : kVariants)
        if (v.order == order && (P < 0 || (v.P == P && v.warps == warps && v.tzfly == (tzfly != 0) && v.tykeep == (tykeep != 0))))
            return &v;
    return nullptr;
}

// Variant for a 3D tree of this order: the default, or BAOBZI_B200_K4="P,warps,tzfly,tykeep" (benchmarks).
// Returns the points per lane (tile = 32 * P slots) and an opaque index for launch_bin_eval3d; -1 if unsupported.
int bin_eval_select(int order, int *P) {
    const K4Variant *v = find_variant(order, -1, 0, 0, 0);
    if (const char *e = std::getenv("BAOBZI_B200_K4")) {
        int p = 0, w = 0, f = 0, y = 0;
        if (std::sscanf(e, "%d,%d,%d,%d", &p, &w, &f, &y) == 4) {
            const K4Variant *o = find_variant(order, p, w, f, y);
            if (o)
                v = o;
            else
                std::fprintf(stderr, "baobzi: BAOBZI_B200_K4=%s is not built for order %d, using the default\n", e, order);
        }
    }
    if (!v)
        return -1;
    *P = v->P > 90 ? (v->warps == 12 ? 3 : 4) : v->P;
    return (int)(v - kVariants);
}

int launch_bin_eval3d(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    if (ws.variant < 0 || ws.variant >= (int)(sizeof(kVariants) / sizeof(kVariants[0])) ||
        kVariants[ws.variant].order != dt.view.order)
        return (int)cudaErrorInvalidValue;
    return kVariants[ws.variant].launch(dt, ws, d_x, d_res, st);
}

} // namespace bz
