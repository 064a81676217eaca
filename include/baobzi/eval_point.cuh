// eval_point.cuh -- one point through the whole path, callable from any kernel: bounds test -> grid cell ->
// subtree descent -> affine map -> Chebyshev sum (reference Function::eval(x), include/baobzi_template.hpp:851-860
// and 866-889).  kernels.cu builds the direct kernels on it; include/baobzi_device.cuh exposes it to user kernels
// (fused producer / consumer hooks).
#pragma once

#include "cheb.cuh"
#include "device_layout.hpp"
#include "lookup.cuh"

namespace bz {

// Node::eval (template.hpp:366-369) for the first output component.
template <int DIM, int ORDER>
__device__ __forceinline__ double eval_leaf(const TreeView &tv, uint32_t node, const NodeHead &head,
                                            const double (&x)[DIM]) {
    const NodeTail tail = load_tail(tv.nodes, node);
    const double *C = tv.coeffs + 2ull * head.coeff16;
    const double u = __dmul_rn(__dsub_rn(x[0], head.c0), tail.ih0);
    if constexpr (DIM == 1) {
        return cheb_eval_1d<ORDER>(u, C);
    } else if constexpr (DIM == 2) {
        const double v = __dmul_rn(__dsub_rn(x[1], head.c1), tail.ih1);
        return cheb_eval_2d<ORDER>(u, v, C, LdgLoad());
    } else {
        const double v = __dmul_rn(__dsub_rn(x[1], head.c1), tail.ih1);
        const double w = __dmul_rn(__dsub_rn(x[2], head.c2), tail.ih2);
        return cheb_eval_3d<ORDER>(u, v, w, C, LdgLoad());
    }
}

// Function::eval for one point: NaN outside [lower_left, upper_right) (template.hpp:872-873, 880-881), else the first
// output component of the leaf holding x.  `node` (optional) receives the reference's global node index, -1 outside.
template <int DIM, int ORDER>
__device__ __forceinline__ double eval_point(const TreeView &tv, const double (&x)[DIM], long long *node = nullptr) {
    if (outside<DIM>(tv, x)) {
        if (node)
            *node = -1;
        return __longlong_as_double(0x7ff8000000000000ll);
    }
    NodeHead head;
    const uint32_t n = locate<DIM>(tv, x, head);
    if (node)
        *node = (long long)n;
    return eval_leaf<DIM, ORDER>(tv, n, head, x);
}

} // namespace bz
