// cheb.cuh -- FP64 tensor-product Chebyshev evaluation on the device, operation for operation the
// reference's cheb_eval (include/baobzi_template.hpp:146-194) as its -mavx2 -mfma build executes it.
//
// Why the order is fixed: the parity contract is "same saved tree, same point => same double".  The
// reference's result depends on where its compiler fused multiply-adds and on the reduction tree of
// Eigen's unrolled dot product; both were read off the compiled reference and are restated (and
// checked bit for bit against all four of the reference's instruction-set builds) in
// oracle/baobzi_oracle.c.  The kernels here implement the AVX2+FMA flavour with explicit
// __fma_rn / __dmul_rn / __dadd_rn intrinsics, which nvcc neither contracts nor re-associates, so
// GPU results are bit-identical to the reference run with BAOBZI_ARCH=avx2, not merely within 4 ulp.
//
// Shape of the arithmetic (ORDER = O):
//   T_0 = 1, T_1 = x, T_n = fma(2x, T_{n-1}, -T_{n-2})                      (template.hpp:169-172, 182-187)
//   1D  Clenshaw on coefficients stored highest degree first                  (template.hpp:146-161)
//   2D  tmp = C * Ty  (each row: one multiply then O-1 dependent FMAs over the columns, Eigen GEMV),
//       value = <Tx, tmp> as a halving tree of O products                     (template.hpp:163-177)
//   3D  for each slab i: tmp = C_i * Tz (as above), d = <Ty, tmp> as Eigen's packet reduction
//       (P = 4 lanes when O % 4 == 0 else 2), value = fma(Tx[i], d, value)    (template.hpp:179-194)
// All O (2D) or O (3D, per slab) row chains are independent: that is the instruction-level
// parallelism the FP64 pipe is fed with.
#pragma once

#include <cuda_runtime.h>

namespace bz {

template <int O>
__device__ __forceinline__ void cheb_T(double x, double (&T)[O]) {
    const double x2 = __dadd_rn(x, x);
    T[0] = 1.0;
    T[1] = x;
#pragma unroll
    for (int j = 2; j < O; ++j)
        T[j] = __fma_rn(x2, T[j - 1], -T[j - 2]);
}

// Halving reduction tree over LEN products a[(START+m)*STRIDE+LANE] * b[...], m = 0..LEN-1
// (Eigen Redux.h unrollers).  Fusion pattern of the reference's FMA builds:
//   e0 + e1        -> fma(a1, b1, a0*b0)
//   e0 + (subtree) -> fma(a0, b0, subtree)
template <int START, int LEN, int STRIDE, int LANE, int O>
__device__ __forceinline__ double dot_tree(const double (&a)[O], const double (&b)[O]) {
    constexpr int i0 = START * STRIDE + LANE;
    if constexpr (LEN == 1) {
        return __dmul_rn(a[i0], b[i0]);
    } else if constexpr (LEN == 2) {
        constexpr int i1 = (START + 1) * STRIDE + LANE;
        return __fma_rn(a[i1], b[i1], __dmul_rn(a[i0], b[i0]));
    } else if constexpr (LEN == 3) {
        return __fma_rn(a[i0], b[i0], dot_tree<START + 1, 2, STRIDE, LANE, O>(a, b));
    } else {
        constexpr int H = LEN / 2;
        return __dadd_rn(dot_tree<START, H, STRIDE, LANE, O>(a, b), dot_tree<START + H, LEN - H, STRIDE, LANE, O>(a, b));
    }
}

// Eigen's vectorised dot of two contiguous O-vectors followed by its horizontal add
template <int O>
__device__ __forceinline__ double dot_packets(const double (&a)[O], const double (&b)[O]) {
    if constexpr (O % 4 == 0) {
        const double l0 = dot_tree<0, O / 4, 4, 0, O>(a, b);
        const double l1 = dot_tree<0, O / 4, 4, 1, O>(a, b);
        const double l2 = dot_tree<0, O / 4, 4, 2, O>(a, b);
        const double l3 = dot_tree<0, O / 4, 4, 3, O>(a, b);
        return __dadd_rn(__dadd_rn(l0, l2), __dadd_rn(l1, l3));
    } else {
        const double l0 = dot_tree<0, O / 2, 2, 0, O>(a, b);
        const double l1 = dot_tree<0, O / 2, 2, 1, O>(a, b);
        return __dadd_rn(l0, l1);
    }
}

// acc = C * v for a column-major O x O block at C (16-byte aligned).  LOAD(p) fetches a double2.
template <int O, class Load>
__device__ __forceinline__ void gemv_block(const double *__restrict__ C, const double (&v)[O], double (&acc)[O],
                                           Load load) {
    const double2 *c2 = reinterpret_cast<const double2 *>(C);
#pragma unroll
    for (int r = 0; r < O; r += 2) {
        const double2 c = load(c2 + r / 2);
        acc[r] = __dmul_rn(c.x, v[0]);
        acc[r + 1] = __dmul_rn(c.y, v[0]);
    }
#pragma unroll
    for (int k = 1; k < O; ++k) {
#pragma unroll
        for (int r = 0; r < O; r += 2) {
            const double2 c = load(c2 + (k * O + r) / 2);
            acc[r] = __fma_rn(c.x, v[k], acc[r]);
            acc[r + 1] = __fma_rn(c.y, v[k], acc[r + 1]);
        }
    }
}

struct LdgLoad {
    __device__ __forceinline__ double2 operator()(const double2 *p) const { return __ldg(p); }
};

template <int O>
__device__ __forceinline__ double cheb_eval_1d(double x, const double *__restrict__ c) {
    const double x2 = __dadd_rn(x, x);
    const double2 *c2 = reinterpret_cast<const double2 *>(c);
    double2 p = __ldg(c2);
    double c0 = p.x, c1 = p.y;
#pragma unroll
    for (int i = 2; i < O; i += 2) {
        p = __ldg(c2 + i / 2);
        double tmp = c1;
        c1 = __dsub_rn(p.x, c0);
        c0 = __fma_rn(c0, x2, tmp);
        tmp = c1;
        c1 = __dsub_rn(p.y, c0);
        c0 = __fma_rn(c0, x2, tmp);
    }
    return __fma_rn(c0, x, c1);
}

template <int O, class Load>
__device__ __forceinline__ double cheb_eval_2d(double x, double y, const double *__restrict__ C, Load load) {
    double Tx[O], Ty[O], acc[O];
    cheb_T<O>(x, Tx);
    cheb_T<O>(y, Ty);
    gemv_block<O>(C, Ty, acc, load);
    return dot_tree<0, O, 1, 0, O>(Tx, acc);
}

template <int O, class Load>
__device__ __forceinline__ double cheb_eval_3d(double x, double y, double z, const double *__restrict__ C, Load load) {
    double Ty[O], Tz[O], acc[O];
    cheb_T<O>(y, Ty);
    cheb_T<O>(z, Tz);
    const double x2 = __dadd_rn(x, x);
    double tx_prev = 1.0, tx = 1.0; // T_{i-1}, T_i of x, advanced as the slabs go by
    double res = 0.0;
#pragma unroll 1
    for (int i = 0; i < O; ++i) {
        if (i == 1) {
            tx_prev = 1.0;
            tx = x;
        } else if (i >= 2) {
            const double t = __fma_rn(x2, tx, -tx_prev);
            tx_prev = tx;
            tx = t;
        }
        gemv_block<O>(C + i * O * O, Tz, acc, load);
        const double d = dot_packets<O>(Ty, acc);
        res = __fma_rn(tx, d, res);
    }
    return res;
}

} // namespace bz
