// eval1d.cu -- 1D trees: the whole tree staged in shared memory, several points per thread in flight.
//
// Reference path (include/baobzi_template.hpp): Function::eval(xp, res, n) 866-889 -> bounds test 872-873 ->
// get_linear_bin 799-820 -> get_node_index 482-493 -> Node::eval 366-369 -> cheb_eval<1> 146-161.
//
// Why a kernel of its own: a 1D evaluation moves 16 bytes through HBM and costs ~3*ORDER flops, so the roofline
// is the memory system.  The generic direct kernel (kernels.cu) keeps one point per thread in flight and fetches
// the node records and the ORDER coefficients of every point through L1 (8 LDG.128 with up to 32 distinct lines
// per warp): it is bound by load latency and the L1 tag rate at ~0.29 of HBM bandwidth.  A fitted 1D tree is tiny
// (tens to hundreds of leaves), so here every CTA copies it into shared memory once:
//
//   s_centre[n_nodes] f64   Box::center                               (template.hpp:93)
//   s_ih    [n_nodes] f64   Box::inv_half_length                      (template.hpp:94)
//   s_coef  [ORDER/2][L] 2xf64  coefficients {k, k+1} of leaf l, TRANSPOSED: one LDS.128 per pair, lanes in
//                           different leaves read different banks, lanes in the same leaf are a broadcast
//   s_child [n_nodes] u32   internal node: global index of child 0; leaf: 0x80000000 | dense leaf rank
//   s_root  [n_bins]  u32   root node of each grid cell               (template.hpp:553)
//
// and then streams tiles of blockDim * P points: P coalesced 8-byte loads per thread are issued before any is
// used (memory-level parallelism without extra warps), the P descents and the P Clenshaw recurrences are
// interleaved (the recurrence is a dependent chain of ORDER-2 DADD + DFMA pairs per point), and one thread per CTA
// asks L2 for the tile the CTA will read two iterations later (cp.async.bulk.prefetch.L2).  The arithmetic is
// cheb_eval_1d of cheb.cuh operation for operation, so results stay bit-identical to the direct kernel's.
//
// What bounds it (ncu, order 16, 1e8 points, profiles/r01_eval1d_ncu_summary.json): the LSU data pipe at 95 % --
// every point needs its own ORDER coefficients in registers, 128 bytes per point at order 16, and shared-memory
// (like L1) write-back is 128 B/clk/SM whether or not lanes share an address: 32 clk per warp of points for the
// coefficients + ~10 for the point, the node table and the result.  That is 1.95e11 evals/s = 0.48 of HBM
// bandwidth (generic kernel: 1.17e11); orders <= 8 need half the bytes and come out HBM bound.
#include "baobzi/cheb.cuh"
#include "device_tree.hpp"
#include "baobzi/lookup.cuh"

#include <cstdio>
#include <cstdlib>

namespace bz {

constexpr int kT1MaxThreads = 1024;
constexpr uint32_t kT1LeafBit = 0x80000000u;
constexpr size_t kT1MaxSmem = 200 * 1024;

__device__ __forceinline__ void prefetch_l2_bulk(const void *p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

template <int ORDER, int P>
__global__ void __launch_bounds__(kT1MaxThreads, 1)
eval_table_1d_kernel(const TreeView tv, const double *__restrict__ x, double *__restrict__ res, int64_t n,
                     int64_t n_tiles) {
    extern __shared__ __align__(16) unsigned char t1_smem[];
    const uint32_t L = tv.n_leaves;
    double *s_centre = reinterpret_cast<double *>(t1_smem);
    double *s_ih = s_centre + tv.n_nodes;
    double *s_coef = s_ih + tv.n_nodes;
    uint32_t *s_child = reinterpret_cast<uint32_t *>(s_coef + (size_t)ORDER * L);
    uint32_t *s_root = s_child + tv.n_nodes;

    for (uint32_t node = threadIdx.x; node < tv.n_nodes; node += blockDim.x) {
        const NodeHead h = load_head(tv.nodes, node);
        const NodeTail t = load_tail(tv.nodes, node);
        s_centre[node] = h.c0;
        s_ih[node] = t.ih0;
        s_child[node] = h.child0 == kLeafMark ? (kT1LeafBit | t.leaf_rank) : h.child0;
    }
    for (uint32_t i = threadIdx.x; i < L * ORDER; i += blockDim.x) {
        const uint32_t leaf = i / ORDER, k = i - leaf * ORDER;
        const NodeHead h = load_head(tv.nodes, __ldg(tv.leaf_node + leaf));
        const double c = __ldg(tv.coeffs + 2ull * h.coeff16 + k); // first output component (template.hpp:366-369)
        s_coef[((k >> 1) * L + leaf) * 2 + (k & 1)] = c; // pair {c_k, c_k+1} of one leaf side by side
    }
    for (uint32_t b = threadIdx.x; b < tv.n_bins; b += blockDim.x)
        s_root[b] = __ldg(tv.bin_root + b);
    __syncthreads();

    const double lo = tv.lower_left[0], hi = tv.upper_right[0], inv_bin = tv.inv_bin_size[0];
    const int last_bin = tv.n_sub[0] - 1;
    const int64_t tile_pts = (int64_t)blockDim.x * P;
    const double qnan = __longlong_as_double(0x7ff8000000000000ll);

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t base = tile * tile_pts;
        if (threadIdx.x == 0) { // the tile of two iterations ahead: whole tiles only, 16-byte aligned start
            const int64_t ahead = tile + 2 * (int64_t)gridDim.x;
            if (ahead + 1 < n_tiles) {
                const uintptr_t a = (reinterpret_cast<uintptr_t>(x + ahead * tile_pts) + 15) & ~uintptr_t(15);
                prefetch_l2_bulk(reinterpret_cast<const void *>(a), (uint32_t)(tile_pts * 8));
            }
        }
        double xv[P];
#pragma unroll
        for (int j = 0; j < P; ++j) {
            const int64_t i = base + (int64_t)j * blockDim.x + threadIdx.x;
            xv[j] = i < n ? __ldcs(x + i) : lo;
        }
        uint32_t leaf[P];
        double u[P];
        bool out[P];
#pragma unroll
        for (int j = 0; j < P; ++j) {
            out[j] = (xv[j] < lo) || (xv[j] >= hi); // template.hpp:872-873; NaN compares false and stays "inside"
            int b = __double2int_rz(__dmul_rn(__dsub_rn(xv[j], lo), inv_bin));
            b = min(max(b, 0), last_bin);
            uint32_t node = s_root[b];
            double c = s_centre[node];
            uint32_t ch = s_child[node];
            while (!(ch & kT1LeafBit)) { // template.hpp:482-493
                node = ch + (uint32_t)(xv[j] > c);
                c = s_centre[node];
                ch = s_child[node];
            }
            leaf[j] = ch & ~kT1LeafBit;
            u[j] = __dmul_rn(__dsub_rn(xv[j], c), s_ih[node]);
        }
        // Clenshaw, coefficients highest degree first (template.hpp:146-161), P independent chains
        double c0[P], c1[P], u2[P];
        const double2 *s_pair = reinterpret_cast<const double2 *>(s_coef);
#pragma unroll
        for (int j = 0; j < P; ++j) {
            u2[j] = __dadd_rn(u[j], u[j]);
            const double2 c = s_pair[leaf[j]];
            c0[j] = c.x;
            c1[j] = c.y;
        }
#pragma unroll
        for (int k = 2; k < ORDER; k += 2) {
#pragma unroll
            for (int j = 0; j < P; ++j) {
                const double2 c = s_pair[(k >> 1) * L + leaf[j]];
                double tmp = c1[j];
                c1[j] = __dsub_rn(c.x, c0[j]);
                c0[j] = __fma_rn(c0[j], u2[j], tmp);
                tmp = c1[j];
                c1[j] = __dsub_rn(c.y, c0[j]);
                c0[j] = __fma_rn(c0[j], u2[j], tmp);
            }
        }
#pragma unroll
        for (int j = 0; j < P; ++j) {
            const int64_t i = base + (int64_t)j * blockDim.x + threadIdx.x;
            if (i < n)
                __stcs(res + i, out[j] ? qnan : __fma_rn(c0[j], u[j], c1[j]));
        }
    }
}

static size_t table_1d_bytes(const TreeView &tv) {
    return 8 * (2 * (size_t)tv.n_nodes + (size_t)tv.order * tv.n_leaves) + 4 * ((size_t)tv.n_nodes + tv.n_bins);
}

bool table_1d_fits(const DeviceTree &dt) {
    const TreeView &tv = dt.view;
    return tv.dim == 1 && tv.n_nodes < kT1LeafBit && table_1d_bytes(tv) <= kT1MaxSmem;
}

int64_t table_1d_min_points() {
    if (const char *s = std::getenv("BAOBZI_B200_TABLE1D_MIN")) // tests lower it; "-1" disables the kernel
        return std::atoll(s);
    return int64_t(1) << 18;
}

template <int ORDER>
static int launch_table_order(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream) {
    constexpr int P = 4;
    auto kern = eval_table_1d_kernel<ORDER, P>;
    const size_t smem = table_1d_bytes(dt.view);
    // as many resident threads as the table leaves room for: 4 x 256, 2 x 512 or 1 x 1024 per SM
    const int threads = smem <= 54 * 1024 ? 256 : (smem <= 110 * 1024 ? 512 : 1024);
    static bool configured[64] = {};
    if (dt.device >= 0 && dt.device < 64 && !configured[dt.device]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kT1MaxSmem);
        if (e != cudaSuccess)
            return (int)e;
        configured[dt.device] = true;
    }
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem);
    if (e != cudaSuccess)
        return (int)e;
    if (per_sm < 1)
        return (int)cudaErrorLaunchOutOfResources;
    const int64_t tile_pts = (int64_t)threads * P;
    const int64_t n_tiles = (n + tile_pts - 1) / tile_pts;
    int64_t grid = (int64_t)dt.sm_count * per_sm;
    if (grid > n_tiles)
        grid = n_tiles;
    kern<<<(unsigned)grid, threads, smem, stream>>>(dt.view, d_x, d_res, n, n_tiles);
    return (int)cudaGetLastError();
}

int launch_eval_table_1d(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream) {
    switch (dt.view.order) {
    case 6: return launch_table_order<6>(dt, d_x, d_res, n, stream);
    case 8: return launch_table_order<8>(dt, d_x, d_res, n, stream);
    case 10: return launch_table_order<10>(dt, d_x, d_res, n, stream);
    case 12: return launch_table_order<12>(dt, d_x, d_res, n, stream);
    case 14: return launch_table_order<14>(dt, d_x, d_res, n, stream);
    case 16: return launch_table_order<16>(dt, d_x, d_res, n, stream);
    default: return (int)cudaErrorInvalidValue;
    }
}

} // namespace bz
