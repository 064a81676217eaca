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

// ------------------------------------------------------------------------------------------ few-leaf 1D trees
// A smooth 1D function fits in a handful of leaves (BASELINE.json configs[1], 1.5 log x on [1,3] at order 16: TWO).
// The table kernel above still pays ORDER * 8 bytes of shared-memory write-back per point for coefficients that the
// whole warp could share -- if the warp sat in one leaf.  This kernel makes it so: per tile of blockDim * P points a
// CTA-local counting sort by leaf (the key has <= kU1MaxLeaves + 1 values), then one pass per leaf in which the
// Clenshaw coefficients are COMPILE-TIME offsets into the kernel's parameter block, i.e. constant-bank operands of the
// DADD / DFMA instructions themselves: no coefficient load of any kind.  Shared-memory traffic per point drops from
// 8 x LDS.128 + the node walk to three 8-byte and two 2-byte accesses, and the FP64 work is exactly cheb_eval_1d's
// (cheb.cuh; same operations in the same order, so the bits are those of the generic kernel).
//   A  P coalesced point loads per thread, bounds test, grid cell, descent in the shared-memory node table,
//      u = (x - centre) * inv_half_length (template.hpp:366-369); key = dense leaf rank, n_leaves when outside
//   B  histogram of the tile (one shared atomic per distinct key and warp: __match_any_sync), exclusive scan over
//      the <= 17 keys, placement: u and the origin (16-bit index in the tile) go to the key's run
//   C  for every leaf: its run, two points per thread in flight, coefficients from the constant bank; values go to
//      s_val[origin] (NaN for the outside run), then one coalesced streaming store of the tile
constexpr int kU1MaxLeaves = 16;
constexpr int kU1Threads = 256, kU1P = 4, kU1Tile = kU1Threads * kU1P;
constexpr int kU1MaxNodes = 64, kU1MaxBins = 64;
constexpr int kU1MinCtas = 6; // resident CTAs per SM asked of ptxas (register budget 42)

struct U1Coef { // by value: lives in the constant bank of the launch
    double c[kU1MaxLeaves][16];
};

template <int ORDER, int LEAF>
__device__ __forceinline__ void u1_run(const U1Coef &cf, const double *s_u, const uint16_t *s_org, double *s_val,
                                       uint32_t begin, uint32_t end) {
    for (uint32_t s = begin + threadIdx.x; s < end; s += 2 * kU1Threads) {
        const bool two = s + kU1Threads < end;
        const double ua = s_u[s], ub = two ? s_u[s + kU1Threads] : 0.0;
        const double a2 = __dadd_rn(ua, ua), b2 = __dadd_rn(ub, ub);
        double a0 = cf.c[LEAF][0], a1 = cf.c[LEAF][1], b0 = a0, b1 = a1;
#pragma unroll
        for (int i = 2; i < ORDER; ++i) { // template.hpp:146-161, two independent chains
            const double c = cf.c[LEAF][i];
            const double ta = a1, tb = b1;
            a1 = __dsub_rn(c, a0);
            b1 = __dsub_rn(c, b0);
            a0 = __fma_rn(a0, a2, ta);
            b0 = __fma_rn(b0, b2, tb);
        }
        // both chains run unconditionally (the compiler otherwise sinks the second one into a branch and the two no
        // longer overlap); a missing second point lands in the spare slot behind the tile
        s_val[s_org[s]] = __fma_rn(a0, ua, a1);
        s_val[two ? (uint32_t)s_org[s + kU1Threads] : (uint32_t)kU1Tile] = __fma_rn(b0, ub, b1);
    }
}

// the runs of the leaves one after the other; `run` = first slot of leaf LEAF (CTA uniform, carried along)
template <int ORDER, int LEAF>
__device__ __forceinline__ uint32_t u1_leaves(const U1Coef &cf, uint32_t n_leaves, const uint32_t *s_tot, uint32_t run,
                                              const double *s_u, const uint16_t *s_org, double *s_val) {
    if constexpr (LEAF < kU1MaxLeaves) {
        if (LEAF < n_leaves) { // CTA uniform
            const uint32_t end = run + s_tot[LEAF];
            u1_run<ORDER, LEAF>(cf, s_u, s_org, s_val, run, end);
            return u1_leaves<ORDER, LEAF + 1>(cf, n_leaves, s_tot, end, s_u, s_org, s_val);
        }
    }
    return run;
}

template <int ORDER, int MINCTAS>
__global__ void __launch_bounds__(kU1Threads, MINCTAS)
eval_sorted_1d_kernel(const TreeView tv, const U1Coef cf, const double *__restrict__ x, double *__restrict__ res,
                      int64_t n, int64_t n_tiles) {
    __shared__ double s_centre[kU1MaxNodes], s_ih[kU1MaxNodes];
    __shared__ uint32_t s_child[kU1MaxNodes], s_root[kU1MaxBins];
    __shared__ double s_u[kU1Tile], s_val[kU1Tile + 1];
    __shared__ uint16_t s_org[kU1Tile];
    __shared__ uint32_t s_tot[2][kU1MaxLeaves + 2]; // per-key counts of the tile, double buffered by tile parity
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    const uint32_t L = tv.n_leaves;
    for (uint32_t node = tid; node < tv.n_nodes; node += kU1Threads) {
        const NodeHead h = load_head(tv.nodes, node);
        const NodeTail t = load_tail(tv.nodes, node);
        s_centre[node] = h.c0;
        s_ih[node] = t.ih0;
        s_child[node] = h.child0 == kLeafMark ? (kT1LeafBit | t.leaf_rank) : h.child0;
    }
    for (uint32_t b = tid; b < tv.n_bins; b += kU1Threads)
        s_root[b] = __ldg(tv.bin_root + b);
    if (tid < 2 * (kU1MaxLeaves + 2))
        (&s_tot[0][0])[tid] = 0;
    __syncthreads();
    const double lo = tv.lower_left[0], hi = tv.upper_right[0], inv_bin = tv.inv_bin_size[0];
    const int last_bin = tv.n_sub[0] - 1;
    const double qnan = __longlong_as_double(0x7ff8000000000000ll);

    auto load_tile = [&](int64_t tile, double (&v)[kU1P]) {
        const int64_t base = tile * kU1Tile;
#pragma unroll
        for (int j = 0; j < kU1P; ++j) {
            const int64_t i = base + j * kU1Threads + tid;
            v[j] = tile < n_tiles && i < n ? __ldcs(x + i) : lo;
        }
    };
    double xv[kU1P], xn[kU1P];
    load_tile(blockIdx.x, xv);
    int par = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, par ^= 1) {
        const int64_t base = tile * kU1Tile;
        const uint32_t m = (uint32_t)(n - base < kU1Tile ? n - base : kU1Tile); // points in this tile
        load_tile(tile + gridDim.x, xn); // the next tile's points travel while this one is sorted and evaluated
        if (tid == 0 && tile + 2 * (int64_t)gridDim.x + 1 < n_tiles) { // ... and the one after is asked into L2
            const uintptr_t a = (reinterpret_cast<uintptr_t>(x + (tile + 2 * (int64_t)gridDim.x) * kU1Tile) + 15) & ~uintptr_t(15);
            prefetch_l2_bulk(reinterpret_cast<const void *>(a), (uint32_t)(kU1Tile * 8));
        }
        uint32_t *tot = s_tot[par];
        // ---- A: leaf and u of every point
        double u[kU1P];
        uint32_t key[kU1P], at[kU1P];
#pragma unroll
        for (int j = 0; j < kU1P; ++j) {
            const bool out = (xv[j] < lo) || (xv[j] >= hi); // template.hpp:872-873; NaN compares false: "inside"
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
            u[j] = __dmul_rn(__dsub_rn(xv[j], c), s_ih[node]);
            key[j] = out ? L : (ch & ~kT1LeafBit);
            if (j * kU1Threads + tid >= m)
                key[j] = 0xFFFFFFFFu; // past the end of the batch: belongs to no run
        }
        // ---- B: one shared atomic per distinct key and warp; what it returns is the group's offset inside the key's run
#pragma unroll
        for (int j = 0; j < kU1P; ++j) {
            const uint32_t grp = __match_any_sync(0xffffffffu, key[j]);
            const int leader = __ffs(grp) - 1;
            uint32_t base_g = 0;
            if (key[j] != 0xFFFFFFFFu && (uint32_t)leader == lane)
                base_g = atomicAdd(&tot[key[j]], (uint32_t)__popc(grp));
            at[j] = __shfl_sync(0xffffffffu, base_g, leader) + (uint32_t)__popc(grp & ((1u << lane) - 1u));
        }
        __syncthreads(); // the counts of the tile are complete
        {   // every thread walks the <= 17 counts itself: first slot of each of its keys
            uint32_t run = 0, first[kU1P];
#pragma unroll
            for (int j = 0; j < kU1P; ++j)
                first[j] = 0;
            for (uint32_t l = 0; l <= L; ++l) {
                const uint32_t t = tot[l];
#pragma unroll
                for (int j = 0; j < kU1P; ++j)
                    first[j] = key[j] == l ? run : first[j];
                run += t;
            }
#pragma unroll
            for (int j = 0; j < kU1P; ++j)
                if (key[j] != 0xFFFFFFFFu) {
                    s_u[first[j] + at[j]] = u[j];
                    s_org[first[j] + at[j]] = (uint16_t)(j * kU1Threads + tid);
                }
        }
        if (tid < kU1MaxLeaves + 2)
            s_tot[par ^ 1][tid] = 0; // the next tile's counters (last read before the previous tile's final barrier)
        __syncthreads();
        // ---- C: one pass per leaf, coefficients from the constant bank; then the points outside the domain
        const uint32_t in_end = u1_leaves<ORDER, 0>(cf, L, tot, 0u, s_u, s_org, s_val);
        for (uint32_t s = in_end + tid; s < m; s += kU1Threads)
            s_val[s_org[s]] = qnan; // outside [lower_left, upper_right): template.hpp:880-881
        __syncthreads();
#pragma unroll
        for (int j = 0; j < kU1P; ++j) {
            const uint32_t t = j * kU1Threads + tid;
            if (t < m)
                __stcs(res + base + t, s_val[t]);
            xv[j] = xn[j];
        }
    }
}

// ------------------------------------------------------------------------------------------ one or two leaves
// The smallest trees (BASELINE.json configs[1]: 1.5 log x on [1,3] at order 16 is TWO leaves) need neither a table walk per
// coefficient nor a sort: the leaf of a point is one predicate, and a coefficient is SEL(c_leaf0[k], c_leaf1[k], predicate)
// with both candidates coming from the launch's constant bank -- no load instruction at all in the recurrence.  Per
// coefficient and point: two 32-bit selects (ALU pipe) + DADD + DFMA (FP64 pipe), the two pipes evenly loaded, against
// 8 bytes of shared-memory write-back in the table kernel.  Same operations in the same order as cheb_eval_1d, so the
// same bits.  Streaming structure of the table kernel: P loads per thread in flight, L2 bulk prefetch two tiles ahead.
// PF: the next tile's points are loaded into registers before the current tile's recurrences start (the four warps of a
// scheduler otherwise all sit in the load at the top of a tile at times: long-scoreboard 1.26 per issue in
// profiles/r02_twoleaf1d_ncu_summary.json).
template <int ORDER, int P, bool PF>
__global__ void __launch_bounds__(256, 2)
eval_two_leaf_1d_kernel(const TreeView tv, const U1Coef cf, const double *__restrict__ x, double *__restrict__ res,
                        int64_t n, int64_t n_tiles) {
    __shared__ double s_centre[kU1MaxNodes], s_ih[kU1MaxNodes];
    __shared__ uint32_t s_child[kU1MaxNodes], s_root[kU1MaxBins];
    for (uint32_t node = threadIdx.x; node < tv.n_nodes; node += blockDim.x) {
        const NodeHead h = load_head(tv.nodes, node);
        const NodeTail t = load_tail(tv.nodes, node);
        s_centre[node] = h.c0;
        s_ih[node] = t.ih0;
        s_child[node] = h.child0 == kLeafMark ? (kT1LeafBit | t.leaf_rank) : h.child0;
    }
    for (uint32_t b = threadIdx.x; b < tv.n_bins; b += blockDim.x)
        s_root[b] = __ldg(tv.bin_root + b);
    __syncthreads();
    const double lo = tv.lower_left[0], hi = tv.upper_right[0], inv_bin = tv.inv_bin_size[0];
    const int last_bin = tv.n_sub[0] - 1;
    const int64_t tile_pts = (int64_t)blockDim.x * P;
    const double qnan = __longlong_as_double(0x7ff8000000000000ll);
    // both candidate blocks live in registers for the whole kernel (opaque to the compiler, which would otherwise
    // re-read them from the constant bank with predicated loads at every use: three LDC per coefficient and point)
    double ca[ORDER], cb[ORDER];
#pragma unroll
    for (int k = 0; k < ORDER; ++k) {
        ca[k] = cf.c[0][k];
        cb[k] = cf.c[1][k];
        asm volatile("" : "+d"(ca[k]), "+d"(cb[k]));
    }
    auto load_tile = [&](int64_t t, double (&dst)[P]) {
        const int64_t b0 = t * tile_pts;
        const bool whole = b0 + tile_pts <= n;
#pragma unroll
        for (int j = 0; j < P; ++j) {
            const int64_t i = b0 + (int64_t)j * blockDim.x + threadIdx.x;
            dst[j] = whole || i < n ? __ldcs(x + i) : lo;
        }
    };
    double xn[P];
    if (PF && (int64_t)blockIdx.x < n_tiles)
        load_tile(blockIdx.x, xn);
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t base = tile * tile_pts;
        if (threadIdx.x == 0) {
            const int64_t ahead = tile + (PF ? 3 : 2) * (int64_t)gridDim.x;
            if (ahead + 1 < n_tiles) {
                const uintptr_t a = (reinterpret_cast<uintptr_t>(x + ahead * tile_pts) + 15) & ~uintptr_t(15);
                prefetch_l2_bulk(reinterpret_cast<const void *>(a), (uint32_t)(tile_pts * 8));
            }
        }
        const bool full = base + tile_pts <= n; // CTA uniform: whole tiles carry no per-point guards, so that the P
                                                // recurrences stay interleaved instead of sinking into P branches
        double xv[P];
        if (PF) {
#pragma unroll
            for (int j = 0; j < P; ++j)
                xv[j] = xn[j];
            if (tile + gridDim.x < n_tiles)
                load_tile(tile + gridDim.x, xn);
        } else {
            load_tile(tile, xv);
        }
        double u[P], u2[P], c0[P], c1[P];
        bool out[P], second[P];
#pragma unroll
        for (int j = 0; j < P; ++j) {
            out[j] = (xv[j] < lo) || (xv[j] >= hi); // template.hpp:872-873; NaN compares false and stays "inside"
            int b = __double2int_rz(__dmul_rn(__dsub_rn(xv[j], lo), inv_bin));
            b = min(max(b, 0), last_bin);
            uint32_t node = s_root[b];
            double c = s_centre[node];
            uint32_t ch = s_child[node];
            while (!(ch & kT1LeafBit)) { // template.hpp:482-493 (at most one step in a two-leaf tree)
                node = ch + (uint32_t)(xv[j] > c);
                c = s_centre[node];
                ch = s_child[node];
            }
            second[j] = (ch & ~kT1LeafBit) != 0;
            u[j] = __dmul_rn(__dsub_rn(xv[j], c), s_ih[node]);
            u2[j] = __dadd_rn(u[j], u[j]);
            c0[j] = second[j] ? cb[0] : ca[0];
            c1[j] = second[j] ? cb[1] : ca[1];
        }
#pragma unroll
        for (int k = 2; k < ORDER; ++k) { // template.hpp:146-161, P independent chains
#pragma unroll
            for (int j = 0; j < P; ++j) {
                const double ck = second[j] ? cb[k] : ca[k];
                const double tmp = c1[j];
                c1[j] = __dsub_rn(ck, c0[j]);
                c0[j] = __fma_rn(c0[j], u2[j], tmp);
            }
        }
        if (full) {
#pragma unroll
            for (int j = 0; j < P; ++j)
                __stcs(res + base + (int64_t)j * blockDim.x + threadIdx.x, out[j] ? qnan : __fma_rn(c0[j], u[j], c1[j]));
        } else {
#pragma unroll
            for (int j = 0; j < P; ++j) {
                const int64_t i = base + (int64_t)j * blockDim.x + threadIdx.x;
                if (i < n)
                    __stcs(res + i, out[j] ? qnan : __fma_rn(c0[j], u[j], c1[j]));
            }
        }
    }
}

constexpr int kTwoLeafDefault = 41; // P = 4 with the register prefetch: 2.28e11 evals/s on the two-leaf order-16 tree against
                                    // 1.99e11 without, 2.18e11 for P = 5 or 3 with it (profiles/r02_sweep_o.txt)

static bool two_leaf_1d_fits(const DeviceTree &dt) {
    if (const char *e = std::getenv("BAOBZI_B200_TWOLEAF1D"))
        if (std::atoi(e) == 0)
            return false;
    const TreeView &tv = dt.view;
    return tv.dim == 1 && tv.n_leaves >= 1 && tv.n_leaves <= 2 && tv.n_nodes <= (uint32_t)kU1MaxNodes &&
           tv.n_bins <= (uint32_t)kU1MaxBins && dt.leaf_coef_1d.size() == (size_t)tv.n_leaves * tv.order;
}

template <int ORDER, int P, bool PF>
static int launch_two_leaf_variant(const DeviceTree &dt, const U1Coef &cf, const double *d_x, double *d_res, int64_t n,
                                   cudaStream_t stream) {
    auto kern = eval_two_leaf_1d_kernel<ORDER, P, PF>;
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, 0);
    if (e != cudaSuccess)
        return (int)e;
    if (per_sm < 1)
        return (int)cudaErrorLaunchOutOfResources;
    const int64_t tile_pts = 256 * P;
    const int64_t n_tiles = (n + tile_pts - 1) / tile_pts;
    int64_t grid = (int64_t)dt.sm_count * per_sm;
    if (grid > n_tiles)
        grid = n_tiles;
    kern<<<(unsigned)grid, 256, 0, stream>>>(dt.view, cf, d_x, d_res, n, n_tiles);
    return (int)cudaGetLastError();
}

template <int ORDER>
static int launch_two_leaf_order(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream) {
    constexpr int P = 4;
    U1Coef cf;
    for (int l = 0; l < kU1MaxLeaves; ++l)
        for (int k = 0; k < 16; ++k) // a one-leaf tree: both candidates are the same block
            cf.c[l][k] = k < ORDER && l < 2 ? dt.leaf_coef_1d[(size_t)((uint32_t)l < dt.view.n_leaves ? l : 0) * ORDER + k] : 0.0;
    // BAOBZI_B200_TWOLEAF1D_VARIANT = 10 * P + PF (experiments; the default is the measured best)
    static const int variant = std::getenv("BAOBZI_B200_TWOLEAF1D_VARIANT") ? std::atoi(std::getenv("BAOBZI_B200_TWOLEAF1D_VARIANT")) : kTwoLeafDefault;
    switch (variant) {
    case 41: return launch_two_leaf_variant<ORDER, 4, true>(dt, cf, d_x, d_res, n, stream);
    case 50: return launch_two_leaf_variant<ORDER, 5, false>(dt, cf, d_x, d_res, n, stream);
    case 51: return launch_two_leaf_variant<ORDER, 5, true>(dt, cf, d_x, d_res, n, stream);
    case 31: return launch_two_leaf_variant<ORDER, 3, true>(dt, cf, d_x, d_res, n, stream);
    default: return launch_two_leaf_variant<ORDER, P, false>(dt, cf, d_x, d_res, n, stream);
    }
}

static int launch_eval_two_leaf_1d(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream) {
    switch (dt.view.order) {
    case 6: return launch_two_leaf_order<6>(dt, d_x, d_res, n, stream);
    case 8: return launch_two_leaf_order<8>(dt, d_x, d_res, n, stream);
    case 10: return launch_two_leaf_order<10>(dt, d_x, d_res, n, stream);
    case 12: return launch_two_leaf_order<12>(dt, d_x, d_res, n, stream);
    case 14: return launch_two_leaf_order<14>(dt, d_x, d_res, n, stream);
    case 16: return launch_two_leaf_order<16>(dt, d_x, d_res, n, stream);
    default: return (int)cudaErrorInvalidValue;
    }
}

bool sorted_1d_fits(const DeviceTree &dt) {
    const TreeView &tv = dt.view;
    // Opt-in (BAOBZI_B200_SORTED1D=1): measured on B200 at 1.6e11 evals/s on the two-leaf benchmark tree against 1.95e11
    // for the table kernel -- the CTA-local sort costs ~140 instructions per 32 points and the kernel is issue bound
    // (ncu: issue slots 75 % busy, FP64 pipe 32 %; profiles/r02_sorted1d_ncu_summary.json).  Kept for the record.
    const char *e = std::getenv("BAOBZI_B200_SORTED1D"); // read per call: tests flip it
    if (!e || std::atoi(e) == 0)
        return false;
    return tv.dim == 1 && tv.n_leaves >= 1 && tv.n_leaves <= (uint32_t)kU1MaxLeaves && tv.n_nodes <= (uint32_t)kU1MaxNodes &&
           tv.n_bins <= (uint32_t)kU1MaxBins && dt.leaf_coef_1d.size() == (size_t)tv.n_leaves * tv.order;
}

template <int ORDER>
static int launch_sorted_order(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream) {
    U1Coef cf;
    for (int l = 0; l < kU1MaxLeaves; ++l)
        for (int k = 0; k < 16; ++k)
            cf.c[l][k] = (uint32_t)l < dt.view.n_leaves && k < ORDER ? dt.leaf_coef_1d[(size_t)l * ORDER + k] : 0.0;
    // resident CTAs per SM asked of ptxas: 6 (40 registers) or 5 (48); BAOBZI_B200_SORTED1D_CTAS picks
    static const int want = std::getenv("BAOBZI_B200_SORTED1D_CTAS") ? std::atoi(std::getenv("BAOBZI_B200_SORTED1D_CTAS")) : kU1MinCtas;
    auto kern = want == 5 ? eval_sorted_1d_kernel<ORDER, 5> : eval_sorted_1d_kernel<ORDER, 6>;
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kU1Threads, 0);
    if (e != cudaSuccess)
        return (int)e;
    if (per_sm < 1)
        return (int)cudaErrorLaunchOutOfResources;
    const int64_t n_tiles = (n + kU1Tile - 1) / kU1Tile;
    int64_t grid = (int64_t)dt.sm_count * per_sm;
    if (grid > n_tiles)
        grid = n_tiles;
    kern<<<(unsigned)grid, kU1Threads, 0, stream>>>(dt.view, cf, d_x, d_res, n, n_tiles);
    return (int)cudaGetLastError();
}

int launch_eval_sorted_1d(const DeviceTree &dt, const double *d_x, double *d_res, int64_t n, cudaStream_t stream) {
    switch (dt.view.order) {
    case 6: return launch_sorted_order<6>(dt, d_x, d_res, n, stream);
    case 8: return launch_sorted_order<8>(dt, d_x, d_res, n, stream);
    case 10: return launch_sorted_order<10>(dt, d_x, d_res, n, stream);
    case 12: return launch_sorted_order<12>(dt, d_x, d_res, n, stream);
    case 14: return launch_sorted_order<14>(dt, d_x, d_res, n, stream);
    case 16: return launch_sorted_order<16>(dt, d_x, d_res, n, stream);
    default: return (int)cudaErrorInvalidValue;
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
    if (two_leaf_1d_fits(dt)) // one or two leaves: coefficients selected between two constant-bank candidates
        return launch_eval_two_leaf_1d(dt, d_x, d_res, n, stream);
    if (sorted_1d_fits(dt)) // a handful of leaves: coefficients as constant-bank operands
        return launch_eval_sorted_1d(dt, d_x, d_res, n, stream);
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
