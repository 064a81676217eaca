// lookup.cuh -- device code shared by the direct and binned kernels: node record loads, the bounds
// test, the grid-cell floor transform and the subtree descent.
#pragma once

#include "device_layout.hpp"

#include <cuda_runtime.h>

namespace bz {

// ------------------------------------------------------------------------------------------ lookup
struct NodeHead { // first 32-byte sector of a NodeRec
    double c0, c1, c2;
    uint32_t child0, coeff16;
};

__device__ __forceinline__ NodeHead load_head(const NodeRec *nodes, uint32_t node) {
    const double2 *p = reinterpret_cast<const double2 *>(nodes + node);
    const double2 a = __ldg(p), b = __ldg(p + 1);
    NodeHead h;
    h.c0 = a.x;
    h.c1 = a.y;
    h.c2 = b.x;
    const unsigned long long w = (unsigned long long)__double_as_longlong(b.y);
    h.child0 = (uint32_t)w;
    h.coeff16 = (uint32_t)(w >> 32);
    return h;
}

struct NodeTail { // second sector
    double ih0, ih1, ih2;
    uint32_t out_dim, leaf_rank;
};

__device__ __forceinline__ NodeTail load_tail(const NodeRec *nodes, uint32_t node) {
    const double2 *p = reinterpret_cast<const double2 *>(nodes + node) + 2;
    const double2 a = __ldg(p), b = __ldg(p + 1);
    NodeTail t;
    t.ih0 = a.x;
    t.ih1 = a.y;
    t.ih2 = b.x;
    const unsigned long long w = (unsigned long long)__double_as_longlong(b.y);
    t.out_dim = (uint32_t)w;
    t.leaf_rank = (uint32_t)(w >> 32);
    return t;
}

// Bounds test of template.hpp:872-873: outside if any x < lower_left or any x >= upper_right
// (comparisons with NaN are false, so NaN is "inside", as in the reference, and propagates to the result).
template <int DIM>
__device__ __forceinline__ bool outside(const TreeView &tv, const double (&x)[DIM]) {
    bool out = false;
#pragma unroll
    for (int j = 0; j < DIM; ++j)
        out = out || (x[j] < tv.lower_left[j]) || (x[j] >= tv.upper_right[j]);
    return out;
}

// get_linear_bin (template.hpp:799-820): (int)((x - ll) * inv_bin), x fastest; then the subtree descent
// (template.hpp:482-493): child bit j = x[j] > centre[j].  The bin is clamped to the grid (the reference
// reads out of bounds when rounding pushes a point just under upper_right into bin n).
template <int DIM>
__device__ __forceinline__ uint32_t locate(const TreeView &tv, const double (&x)[DIM], NodeHead &head) {
    int lin = 0, stride = 1;
#pragma unroll
    for (int j = 0; j < DIM; ++j) {
        const double xb = __dsub_rn(x[j], tv.lower_left[j]);
        int b = __double2int_rz(__dmul_rn(xb, tv.inv_bin_size[j]));
        b = min(max(b, 0), tv.n_sub[j] - 1);
        lin += stride * b;
        stride *= tv.n_sub[j];
    }
    uint32_t node = __ldg(tv.bin_root + lin);
    head = load_head(tv.nodes, node);
    while (head.child0 != kLeafMark) {
        uint32_t bits = (uint32_t)(x[0] > head.c0);
        if (DIM > 1)
            bits |= (uint32_t)(x[1] > head.c1) << 1;
        if (DIM > 2)
            bits |= (uint32_t)(x[2] > head.c2) << 2;
        node = head.child0 + bits;
        head = load_head(tv.nodes, node);
    }
    return node;
}

// get_linear_bin alone (template.hpp:799-820), clamped like locate(): the key of the bucketed path, which needs
// no tree access at all.
template <int DIM>
__device__ __forceinline__ uint32_t linear_bin(const TreeView &tv, const double (&x)[DIM]) {
    int lin = 0, stride = 1;
#pragma unroll
    for (int j = 0; j < DIM; ++j) {
        const double xb = __dsub_rn(x[j], tv.lower_left[j]);
        int b = __double2int_rz(__dmul_rn(xb, tv.inv_bin_size[j]));
        b = min(max(b, 0), tv.n_sub[j] - 1);
        lin += stride * b;
        stride *= tv.n_sub[j];
    }
    return (uint32_t)lin;
}

// Same walk, but only the dense leaf rank is returned.  Undivided grid cells (every cell of a tree whose fit
// never refined past the grid, e.g. the 3D benchmark trees) are answered by the bin_leaf table alone; otherwise
// every step fetches the whole 64-byte record (both sectors of one line in a single round trip): one dependent
// load fewer than locate() + load_tail().
// descent from the root of grid cell `lin` to the leaf holding x, returning the dense leaf rank
template <int DIM>
__device__ __forceinline__ uint32_t walk_to_leaf(const TreeView &tv, uint32_t lin, const double (&x)[DIM]) {
    uint32_t node = __ldg(tv.bin_root + lin);
    for (;;) {
        const double2 *p = reinterpret_cast<const double2 *>(tv.nodes + node);
        const double2 a = __ldg(p), b = __ldg(p + 1), d = __ldg(p + 3);
        const uint32_t child0 = (uint32_t)(unsigned long long)__double_as_longlong(b.y);
        if (child0 == kLeafMark)
            return (uint32_t)((unsigned long long)__double_as_longlong(d.y) >> 32);
        uint32_t bits = (uint32_t)(x[0] > a.x);
        if (DIM > 1)
            bits |= (uint32_t)(x[1] > a.y) << 1;
        if (DIM > 2)
            bits |= (uint32_t)(x[2] > b.x) << 2;
        node = child0 + bits;
    }
}

template <int DIM>
__device__ __forceinline__ uint32_t locate_leaf(const TreeView &tv, const double (&x)[DIM]) {
    const uint32_t lin = linear_bin<DIM>(tv, x);
    const uint32_t direct = __ldg(tv.bin_leaf + lin); // undivided cell: done after one 4-byte load
    return direct != kNotLeaf ? direct : walk_to_leaf<DIM>(tv, lin, x);
}

template <int DIM>
__device__ __forceinline__ void load_point(const double *__restrict__ x, int64_t i, double (&p)[DIM]) {
    if constexpr (DIM == 2) {
        // one 16-byte load when the caller's array allows it (uniform test on the base pointer: the reference API does
        // not promise more than 8-byte alignment), else two 8-byte ones
        if ((reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
            const double2 v = __ldcs(reinterpret_cast<const double2 *>(x) + i); // streamed once
            p[0] = v.x;
            p[1] = v.y;
        } else {
            p[0] = __ldcs(x + 2 * i);
            p[1] = __ldcs(x + 2 * i + 1);
        }
    } else {
#pragma unroll
        for (int j = 0; j < DIM; ++j)
            p[j] = __ldcs(x + i * DIM + j);
    }
}


// same, but with the default cache policy: the binned path reads each point twice (lookup, then gather)
template <int DIM>
__device__ __forceinline__ void load_point_keep(const double *__restrict__ x, size_t i, double (&p)[DIM]) {
#pragma unroll
    for (int j = 0; j < DIM; ++j)
        p[j] = __ldg(x + i * DIM + j);
}

} // namespace bz
