// bin_eval.cu -- K4 of the binned path: evaluation of leaf-grouped points (see binned.cu for K1-K3).
//
// One warp per tile of 32*P consecutive slots of ONE leaf; lane l owns slots l, l+32, ... (P points, register
// blocked).  Because the whole warp sits in one leaf, a coefficient pair is a warp-uniform 16-byte load whose
// value feeds 2*P FMAs per lane.
//
// What bounds it (ncu, profiles/): the FP64 pipe issues one warp-wide DFMA per 2 cycles per SM sub-partition,
// but a uniform 128-bit load still occupies the LSU data pipe for 4 wavefronts (512 B written back to the
// register file at 128 B/clk per SM).  With P points per lane the ratio is 4 LSU cycles (shared by the 4
// sub-partitions) per 2*P DFMAs = 4*P FP64 cycles of one sub-partition, so P = 2 is LSU-bound (4 x 4 > 8) and
// P = 4 is balanced (16 = 16).  Hence P = 4 where the registers allow it (order <= 8), the T-vector of the
// middle axis is rebuilt at the end of every slab instead of being kept live, and the coefficient block can be
// staged per warp in shared memory (SMEM variant) so that loads have a short fixed latency and never miss.
//
// Arithmetic: exactly cheb.cuh's order (bit-identical to the reference's AVX2+FMA build); only the interleaving
// of independent row chains across the P points differs, which does not change any rounding.
#include "cheb.cuh"
#include "device_tree.hpp"
#include "lookup.cuh"

namespace bz {

constexpr uint32_t kInvalidSlotE = 0xFFFFFFFFu;

template <int O, int P, bool SMEM, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
bin_eval3d_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                  const uint32_t *__restrict__ perm, const uint32_t *__restrict__ tile_leaf,
                  const uint32_t *__restrict__ n_tiles_p) {
    constexpr int TILE = 32 * P;
    constexpr int BLK2 = O * O * O / 2; // double2 per leaf block
    extern __shared__ double2 smem2[];
    const uint32_t n_tiles = __ldg(n_tiles_p);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double2 *wbuf = smem2 + (SMEM ? warp * BLK2 : 0);

    for (uint32_t t = blockIdx.x * WARPS + warp; t < n_tiles; t += gridDim.x * WARPS) {
        const uint32_t leaf = __ldg(tile_leaf + t);
        const uint32_t node = __ldg(tv.leaf_node + leaf);
        const NodeHead head = load_head(tv.nodes, node);
        const NodeTail tail = load_tail(tv.nodes, node);
        const double2 *Cg = reinterpret_cast<const double2 *>(tv.coeffs + 2ull * head.coeff16);

        uint32_t idx[P];
#pragma unroll
        for (int p = 0; p < P; ++p)
            idx[p] = __ldg(perm + (size_t)t * TILE + p * 32 + lane);
        if (SMEM) { // stage this leaf's block: coalesced 16-byte loads, no reuse across warps needed
            __syncwarp();
#pragma unroll 4
            for (int j = lane; j < BLK2; j += 32)
                wbuf[j] = __ldg(Cg + j);
        }
        double Tz[P][O], acc[P][O];
        double v[P], x2[P], tx[P], tx_prev[P], val[P];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            double px = head.c0, py = head.c1, pz = head.c2; // padding lanes evaluate the leaf centre
            if (idx[p] != kInvalidSlotE) {
                const double *q = x + 3ull * idx[p];
                px = __ldg(q);
                py = __ldg(q + 1);
                pz = __ldg(q + 2);
            }
            const double u = __dmul_rn(__dsub_rn(px, head.c0), tail.ih0);
            v[p] = __dmul_rn(__dsub_rn(py, head.c1), tail.ih1);
            const double w = __dmul_rn(__dsub_rn(pz, head.c2), tail.ih2);
            cheb_T<O>(w, Tz[p]);
            x2[p] = __dadd_rn(u, u);
            tx_prev[p] = 1.0;
            tx[p] = u;
            val[p] = 0.0;
        }
        if (SMEM)
            __syncwarp();
        const double2 *Cb = SMEM ? wbuf : Cg;

#pragma unroll 1
        for (int i = 0; i < O; ++i) {
            const double2 *c2 = Cb + i * (O * O / 2);
            // acc[p][r] = sum_k C_i(r,k) Tz[p][k]: row chains in k order, first term a plain product
#pragma unroll
            for (int k = 0; k < O; ++k) {
#pragma unroll
                for (int r = 0; r < O; r += 2) {
                    const double2 c = SMEM ? c2[(k * O + r) / 2] : __ldg(c2 + (k * O + r) / 2);
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        if (k == 0) {
                            acc[p][r] = __dmul_rn(c.x, Tz[p][0]);
                            acc[p][r + 1] = __dmul_rn(c.y, Tz[p][0]);
                        } else {
                            acc[p][r] = __fma_rn(c.x, Tz[p][k], acc[p][r]);
                            acc[p][r + 1] = __fma_rn(c.y, Tz[p][k], acc[p][r + 1]);
                        }
                    }
                }
            }
#pragma unroll
            for (int p = 0; p < P; ++p) {
                double Ty[O];
                cheb_T<O>(v[p], Ty); // rebuilt per slab: O-2 FMAs instead of O live registers per point
                const double d = dot_packets<O>(Ty, acc[p]);
                double ti;
                if (i == 0) {
                    ti = 1.0;
                } else if (i == 1) {
                    ti = tx[p];
                } else {
                    ti = __fma_rn(x2[p], tx[p], -tx_prev[p]);
                    tx_prev[p] = tx[p];
                    tx[p] = ti;
                }
                val[p] = __fma_rn(ti, d, val[p]);
            }
        }
#pragma unroll
        for (int p = 0; p < P; ++p)
            if (idx[p] != kInvalidSlotE)
                res[idx[p]] = val[p];
    }
}

// ------------------------------------------------------------------------------------------ pipelined K4
// Same arithmetic, but each warp runs a 2-stage software pipeline over its tiles: while tile j is evaluated
// out of shared memory, tile j+1's inputs (the leaf's coefficient block, the node record, the gathered points
// and their indices) are already in flight as cp.async copies, and tile j+2's descriptor is being fetched into
// registers.  The prologue latency (dependent L2/HBM loads: slot -> point) that left the FP64 pipe idle a third
// of the time in the unpipelined kernel (profiles/r01_k4_P4_smem_ncu_summary.json: long_scoreboard) is hidden.
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

template <int O, int P, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
bin_eval3d_pipe_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res,
                       const uint32_t *__restrict__ perm, const uint2 *__restrict__ tile_info,
                       const uint32_t *__restrict__ n_tiles_p) {
    constexpr int TILE = 32 * P;
    constexpr int BLK2 = O * O * O / 2;
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
        if (tt < n_tiles) {
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
        double Tz[P][O], acc[P][O];
        double v[P], x2[P], tx[P], tx_prev[P], val[P];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const bool ok = st.idx[p * 32 + lane] != kInvalidSlotE;
            const double px = ok ? st.pts[0 * TILE + p * 32 + lane] : c0; // padding lanes evaluate the leaf centre
            const double py = ok ? st.pts[1 * TILE + p * 32 + lane] : c1;
            const double pz = ok ? st.pts[2 * TILE + p * 32 + lane] : c2c;
            const double u = __dmul_rn(__dsub_rn(px, c0), ih0);
            v[p] = __dmul_rn(__dsub_rn(py, c1), ih1);
            const double w = __dmul_rn(__dsub_rn(pz, c2c), ih2);
            cheb_T<O>(w, Tz[p]);
            x2[p] = __dadd_rn(u, u);
            tx_prev[p] = 1.0;
            tx[p] = u;
            val[p] = 0.0;
        }
#pragma unroll 1
        for (int i = 0; i < O; ++i) {
            const double2 *cc = st.coeff + i * (O * O / 2);
#pragma unroll
            for (int k = 0; k < O; ++k) {
#pragma unroll
                for (int r = 0; r < O; r += 2) {
                    const double2 c = cc[(k * O + r) / 2];
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        if (k == 0) {
                            acc[p][r] = __dmul_rn(c.x, Tz[p][0]);
                            acc[p][r + 1] = __dmul_rn(c.y, Tz[p][0]);
                        } else {
                            acc[p][r] = __fma_rn(c.x, Tz[p][k], acc[p][r]);
                            acc[p][r + 1] = __fma_rn(c.y, Tz[p][k], acc[p][r + 1]);
                        }
                    }
                }
            }
#pragma unroll
            for (int p = 0; p < P; ++p) {
                double Ty[O];
                cheb_T<O>(v[p], Ty);
                const double d = dot_packets<O>(Ty, acc[p]);
                double ti;
                if (i == 0) {
                    ti = 1.0;
                } else if (i == 1) {
                    ti = tx[p];
                } else {
                    ti = __fma_rn(x2[p], tx[p], -tx_prev[p]);
                    tx_prev[p] = tx[p];
                    tx[p] = ti;
                }
                val[p] = __fma_rn(ti, d, val[p]);
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

template <int O, int P, int WARPS>
static int launch_pipe(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    auto kern = bin_eval3d_pipe_kernel<O, P, WARPS>;
    const size_t smem = sizeof(PipeStage<O, P>) * 2 * WARPS;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return (int)e;
        configured = true;
    }
    kern<<<dt.sm_count, WARPS * 32, smem, st>>>(dt.view, d_x, d_res, ws.perm, ws.tile_info, ws.n_tiles);
    return (int)cudaGetLastError();
}

template <int O, int P, bool SMEM, int WARPS>
static int launch_variant(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res,
                          cudaStream_t st) {
    auto kern = bin_eval3d_kernel<O, P, SMEM, WARPS>;
    const size_t smem = SMEM ? (size_t)WARPS * O * O * O * sizeof(double) : 0;
    static bool configured = false; // per instantiation
    if (!configured) {
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess)
                return (int)e;
        }
        configured = true;
    }
    int per_sm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS * 32, smem);
    if (per_sm < 1)
        per_sm = 1;
    kern<<<dt.sm_count * per_sm, WARPS * 32, smem, st>>>(dt.view, d_x, d_res, ws.perm, ws.tile_leaf, ws.n_tiles);
    return (int)cudaGetLastError();
}

// Default points per lane for a 3D tree of this order (what the registers allow, see header comment).
int bin_eval_default_P(int order) { return order <= 8 ? 4 : 2; }
bool bin_eval_default_smem(int order) { return order <= 10; }

// WARPS per CTA: the SMEM variant needs WARPS * 8 * O^3 bytes of shared memory (<= 227 KB)
#define BZ_K4(O_, P_, S_, W_)                                                                                          \
    if (order == O_ && P == P_ && smem == S_)                                                                          \
        return launch_variant<O_, P_, S_, W_>(dt, ws, d_x, d_res, st);

int launch_bin_eval3d(const DeviceTree &dt, const BinWorkspace &ws, const double *d_x, double *d_res, cudaStream_t st) {
    const int order = dt.view.order, P = ws.P;
    const bool smem = ws.smem_coeffs;
    if (ws.pipelined) {
        if (order == 6 && P == 4) return launch_pipe<6, 4, 8>(dt, ws, d_x, d_res, st);
        if (order == 8 && P == 4) return launch_pipe<8, 4, 8>(dt, ws, d_x, d_res, st);
        if (order == 8 && P == 2) return launch_pipe<8, 2, 8>(dt, ws, d_x, d_res, st);
        if (order == 10 && P == 2) return launch_pipe<10, 2, 8>(dt, ws, d_x, d_res, st);
        if (order == 12 && P == 2) return launch_pipe<12, 2, 4>(dt, ws, d_x, d_res, st);
        if (order == 14 && P == 2) return launch_pipe<14, 2, 4>(dt, ws, d_x, d_res, st);
        if (order == 16 && P == 2) return launch_pipe<16, 2, 2>(dt, ws, d_x, d_res, st);
    }
    BZ_K4(6, 4, true, 8) BZ_K4(6, 4, false, 8) BZ_K4(6, 2, true, 8) BZ_K4(6, 2, false, 8)
    BZ_K4(8, 4, true, 8) BZ_K4(8, 4, false, 8) BZ_K4(8, 2, true, 8) BZ_K4(8, 2, false, 8)
    BZ_K4(10, 2, true, 8) BZ_K4(10, 2, false, 8) BZ_K4(10, 4, false, 4)
    BZ_K4(12, 2, true, 8) BZ_K4(12, 2, false, 8) BZ_K4(12, 1, false, 8)
    BZ_K4(14, 2, true, 8) BZ_K4(14, 2, false, 8) BZ_K4(14, 1, false, 8)
    BZ_K4(16, 2, true, 4) BZ_K4(16, 2, false, 8) BZ_K4(16, 1, false, 8)
    return (int)cudaErrorInvalidValue;
}

bool bin_eval_pipe_exists(int order, int P) {
    return (order == 6 && P == 4) || (order == 8 && (P == 4 || P == 2)) || (order >= 10 && order <= 16 && P == 2);
}

bool bin_eval_variant_exists(int order, int P, bool smem) {
    static const int table[][3] = {{6, 4, 1}, {6, 4, 0}, {6, 2, 1}, {6, 2, 0}, {8, 4, 1}, {8, 4, 0}, {8, 2, 1}, {8, 2, 0},
                                   {10, 2, 1}, {10, 2, 0}, {10, 4, 0}, {12, 2, 1}, {12, 2, 0}, {12, 1, 0}, {14, 2, 1},
                                   {14, 2, 0}, {14, 1, 0}, {16, 2, 1}, {16, 2, 0}, {16, 1, 0}};
    for (auto &e : table)
        if (e[0] == order && e[1] == P && (e[2] != 0) == smem)
            return true;
    return false;
}

} // namespace bz
