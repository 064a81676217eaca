// reduce.cu -- evaluation fused with a consumer: sum_i w_i * f(x_i) without an n-sized result array.
//
// No reference counterpart (the reference only writes res[i], include/baobzi_template.hpp:866-889); the values being
// summed are the bit-exact ones of the evaluation path.  Two shapes:
//   direct trees   eval_dot_kernel: the library's own per-point device function (eval_point.cuh, the same one
//                  include/baobzi_device.cuh hands to user kernels) + a per-thread accumulator: 8*dim (+8 for w)
//                  bytes per evaluation instead of 8*(dim+1) (+8)
//   binned/bucketed trees  the chunk's values go to a chunk-sized scratch array (it lives in L2), dot_kernel consumes it
// The summation order is fixed by (n, grid), never by timing: thread t adds its grid-stride elements in sequence, lanes
// combine by a shuffle tree, warps and CTAs in index order -- two calls on the same device give the same bits.
#include "device_tree.hpp"
#include "baobzi/eval_point.cuh"

namespace bz {

constexpr int kDotThreads = 256;

// sum over the CTA in a fixed order; result valid in thread 0
__device__ __forceinline__ double block_sum(double v) {
    __shared__ double s_warp[32];
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1)
        v = __dadd_rn(v, __shfl_xor_sync(0xffffffffu, v, d));
    const int warp = threadIdx.x >> 5, n_warps = (blockDim.x + 31) >> 5;
    if ((threadIdx.x & 31) == 0)
        s_warp[warp] = v;
    __syncthreads();
    double total = 0.0;
    if (threadIdx.x == 0)
        for (int k = 0; k < n_warps; ++k)
            total = __dadd_rn(total, s_warp[k]);
    return total;
}

template <int DIM, int ORDER>
__global__ void __launch_bounds__(kDotThreads)
eval_dot_kernel(const TreeView tv, const double *__restrict__ x, const double *__restrict__ w, int64_t n,
                double *__restrict__ partial) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        double p[DIM];
        load_point<DIM>(x, i, p);
        const double y = eval_point<DIM, ORDER>(tv, p);
        acc = w ? __fma_rn(__ldcs(w + i), y, acc) : __dadd_rn(acc, y);
    }
    const double total = block_sum(acc);
    if (threadIdx.x == 0)
        partial[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kDotThreads)
dot_kernel(const double *__restrict__ y, const double *__restrict__ w, int64_t n, double *__restrict__ partial) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        acc = w ? __fma_rn(__ldcs(w + i), y[i], acc) : __dadd_rn(acc, y[i]);
    const double total = block_sum(acc);
    if (threadIdx.x == 0)
        partial[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) sum_partials_kernel(const double *__restrict__ partial, int64_t m,
                                                            double *__restrict__ out) {
    double acc = 0.0;
    for (int64_t i = threadIdx.x; i < m; i += blockDim.x)
        acc = __dadd_rn(acc, partial[i]);
    const double total = block_sum(acc);
    if (threadIdx.x == 0)
        *out = total;
}

int dot_grid(const DeviceTree &dt) { return dt.sm_count * 4; }

#define BZ_DOT_ORDER(DIM, order, CALL)                                                                                 \
    switch (order) {                                                                                                   \
    case 6: { constexpr int O = 6; constexpr int D = DIM; CALL; } break;                                               \
    case 8: { constexpr int O = 8; constexpr int D = DIM; CALL; } break;                                               \
    case 10: { constexpr int O = 10; constexpr int D = DIM; CALL; } break;                                             \
    case 12: { constexpr int O = 12; constexpr int D = DIM; CALL; } break;                                             \
    case 14: { constexpr int O = 14; constexpr int D = DIM; CALL; } break;                                             \
    case 16: { constexpr int O = 16; constexpr int D = DIM; CALL; } break;                                             \
    default: return (int)cudaErrorInvalidValue;                                                                        \
    }

// partial[0 .. dot_grid) = per-CTA sums of w_i * f(x_i) (w may be null: plain sum)
int launch_eval_dot(const DeviceTree &dt, const double *d_x, const double *d_w, int64_t n, double *d_partial,
                    cudaStream_t stream, int64_t *launches) {
    const unsigned grid = (unsigned)dot_grid(dt);
    switch (dt.view.dim) {
    case 1: BZ_DOT_ORDER(1, dt.view.order, (eval_dot_kernel<D, O><<<grid, kDotThreads, 0, stream>>>(dt.view, d_x, d_w, n, d_partial))) break;
    case 2: BZ_DOT_ORDER(2, dt.view.order, (eval_dot_kernel<D, O><<<grid, kDotThreads, 0, stream>>>(dt.view, d_x, d_w, n, d_partial))) break;
    case 3: BZ_DOT_ORDER(3, dt.view.order, (eval_dot_kernel<D, O><<<grid, kDotThreads, 0, stream>>>(dt.view, d_x, d_w, n, d_partial))) break;
    default: return (int)cudaErrorInvalidValue;
    }
    if (launches)
        *launches += 1;
    return (int)cudaGetLastError();
}

int launch_dot(const DeviceTree &dt, const double *d_y, const double *d_w, int64_t n, double *d_partial,
               cudaStream_t stream, int64_t *launches) {
    dot_kernel<<<(unsigned)dot_grid(dt), kDotThreads, 0, stream>>>(d_y, d_w, n, d_partial);
    if (launches)
        *launches += 1;
    return (int)cudaGetLastError();
}

int launch_sum_partials(const double *d_partial, int64_t m, double *d_out, cudaStream_t stream, int64_t *launches) {
    sum_partials_kernel<<<1, 1024, 0, stream>>>(d_partial, m, d_out);
    if (launches)
        *launches += 1;
    return (int)cudaGetLastError();
}

} // namespace bz
