// baobzi_device.cuh -- device-side entry points of the B200 library (CUDA C++, header only).
//
// The reference has no counterpart: its evaluation is host code (include/baobzi_template.hpp:851-889).  This is the
// "fused producer / consumer hook" of SURVEY.md 8f-4: a kernel of yours that PRODUCES points can evaluate the fitted
// function in registers and CONSUME the value (accumulate, test, scatter ...) without the 8 * (dim + 1) bytes per
// evaluation that baobzi_eval_multi_device moves through HBM.
//
//   #include <baobzi_device.cuh>                     // nvcc -I<repo>/include, link libbaobzi.so
//   baobzi_device::View view;
//   baobzi_device_view(func, device, &view, sizeof view);            // C ABI, include/baobzi.h
//   my_kernel<<<...>>>(view, ...);
//   __global__ void my_kernel(baobzi_device::View view, ...) {
//       double x[2] = {...};
//       double y = baobzi_device::eval<2, 8>(view, x);               // DIM, ORDER of the fitted function
//   }
//
// The view stays valid until baobzi_free(func); it refers to the replica of the tree on `device`.  DIM and ORDER are
// template parameters because the arithmetic is fully unrolled; they must equal func->DIM / func->ORDER
// (baobzi_device::matches checks).  Values are bit-identical to baobzi_eval_multi's (same code: the library's own
// direct kernels are built on these functions), NaN outside [lower_left, upper_right).  Whatever flags the including
// translation unit is compiled with: every floating-point operation is an explicit round-to-nearest intrinsic, so
// nvcc's default -fmad=true cannot change the result.
#pragma once

#include "baobzi.h"

#include "baobzi/eval_point.cuh" // self-contained: everything it needs lives under include/baobzi/

namespace baobzi_device {

using View = bz::TreeView; // passed to kernels by value (160 bytes of launch parameters)

// true when the view was made from a function of this dimension and order
inline bool matches(const View &v, int dim, int order) { return v.dim == dim && v.order == order; }

// f(x), first output component; NaN outside the domain
template <int DIM, int ORDER>
__device__ __forceinline__ double eval(const View &v, const double (&x)[DIM]) {
    return bz::eval_point<DIM, ORDER>(v, x);
}

// same, also returning the reference's global node index of the leaf (-1 outside):
// Function::get_global_node_index, baobzi_template.hpp:846-849
template <int DIM, int ORDER>
__device__ __forceinline__ double eval(const View &v, const double (&x)[DIM], long long &node) {
    return bz::eval_point<DIM, ORDER>(v, x, &node);
}

// Every output component of a 1D multi-output fit: y[k], k < v.out_dim (reference Function::eval(x, res),
// baobzi_template.hpp:851-860 with Node::eval 371-375: component k uses the coefficient block at + k * ORDER).
template <int ORDER>
__device__ __forceinline__ void eval_all_1d(const View &v, double x, double *y) {
    const double p[1] = {x};
    if (bz::outside<1>(v, p)) {
        for (uint32_t k = 0; k < v.out_dim; ++k)
            y[k] = __longlong_as_double(0x7ff8000000000000ll);
        return;
    }
    bz::NodeHead head;
    const uint32_t node = bz::locate<1>(v, p, head);
    const bz::NodeTail tail = bz::load_tail(v.nodes, node);
    const double u = __dmul_rn(__dsub_rn(x, head.c0), tail.ih0);
    const double *C = v.coeffs + 2ull * head.coeff16;
    for (uint32_t k = 0; k < v.out_dim; ++k)
        y[k] = k < tail.out_dim ? bz::cheb_eval_1d<ORDER>(u, C + (size_t)k * ORDER)
                                : __longlong_as_double(0x7ff8000000000000ll);
}

} // namespace baobzi_device
