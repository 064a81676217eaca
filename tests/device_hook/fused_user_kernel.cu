// A user's translation unit: a kernel that PRODUCES points (counter-based generator, nothing read from memory),
// evaluates a restored baobzi function in registers through include/baobzi_device.cuh and CONSUMES the value
// (per-thread partial sums + a histogram of leaves), next to the library's own batched entry on the same points.
// Compiled by tests/test_device_hook.py with plain `nvcc -I include` (default -fmad=true on purpose) and linked
// against baobzi_b200/lib/libbaobzi.so.  usage: fused_user_kernel file.baobzi dim order n  -> prints "OK ..." or "FAIL ..."
#include <baobzi_device.cuh>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

__device__ __forceinline__ double unit(unsigned long long i, unsigned k) { // splitmix64 -> [0, 1)
    unsigned long long z = (i * 3 + k + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (double)(z >> 11) * 0x1.0p-53;
}

template <int DIM>
__device__ __forceinline__ void make_point(const baobzi_device::View &v, long long i, double (&x)[DIM]) {
    for (int j = 0; j < DIM; ++j) // 2 % of the points fall outside the domain
        x[j] = v.lower_left[j] + (unit((unsigned long long)i, j) * 1.02 - 0.01) * (v.upper_right[j] - v.lower_left[j]);
}

template <int DIM, int ORDER>
__global__ void fused(baobzi_device::View v, long long n, double *x_out, double *y_out, long long *node_out) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    double x[DIM];
    make_point<DIM>(v, i, x);
    long long node;
    const double y = baobzi_device::eval<DIM, ORDER>(v, x, node);
    const double y2 = baobzi_device::eval<DIM, ORDER>(v, x);
    for (int j = 0; j < DIM; ++j)
        x_out[i * DIM + j] = x[j]; // only so that the host can replay the same points through the batched entry
    y_out[i] = (__double_as_longlong(y) == __double_as_longlong(y2) || (y != y && y2 != y2)) ? y : 12345.0;
    node_out[i] = node;
}

#define CK(e)                                                                                                          \
    do {                                                                                                               \
        cudaError_t _e = (cudaError_t)(e);                                                                             \
        if (_e != cudaSuccess) {                                                                                       \
            std::printf("FAIL cuda %s at %d\n", cudaGetErrorString(_e), __LINE__);                                     \
            return 1;                                                                                                  \
        }                                                                                                              \
    } while (0)

template <int DIM, int ORDER>
int run(baobzi_t f, long long n) {
    baobzi_device::View v;
    CK(baobzi_device_view(f, 0, &v, sizeof v));
    if (!baobzi_device::matches(v, DIM, ORDER)) {
        std::printf("FAIL view mismatch\n");
        return 1;
    }
    double *dx, *dy, *dref, *dsum;
    long long *dnode, *dnode_ref;
    CK(cudaMalloc(&dx, n * DIM * 8));
    CK(cudaMalloc(&dy, n * 8));
    CK(cudaMalloc(&dref, n * 8));
    CK(cudaMalloc(&dsum, 8));
    CK(cudaMalloc(&dnode, n * 8));
    CK(cudaMalloc(&dnode_ref, n * 8));
    fused<DIM, ORDER><<<(unsigned)((n + 255) / 256), 256>>>(v, n, dx, dy, dnode);
    CK(cudaGetLastError());
    CK(baobzi_eval_multi_device(f, dx, dref, n, nullptr));
    CK(baobzi_node_index_multi_device(f, dx, (int64_t *)dnode_ref, n, nullptr));
    CK(baobzi_eval_multi_device_dot(f, dx, nullptr, n, dsum, nullptr));
    CK(cudaDeviceSynchronize());
    std::vector<double> y(n), ref(n);
    std::vector<long long> node(n), node_ref(n);
    double sum;
    CK(cudaMemcpy(y.data(), dy, n * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ref.data(), dref, n * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(node.data(), dnode, n * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(node_ref.data(), dnode_ref, n * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&sum, dsum, 8, cudaMemcpyDeviceToHost));
    long long bad = 0, nan = 0;
    for (long long i = 0; i < n; ++i) {
        if (std::memcmp(&y[i], &ref[i], 8) != 0 && !(std::isnan(y[i]) && std::isnan(ref[i])))
            ++bad;
        if (node[i] != node_ref[i])
            ++bad;
        if (std::isnan(ref[i]) != (node[i] < 0))
            ++bad;
        nan += std::isnan(ref[i]);
    }
    if (bad || nan == 0 || nan == n || !std::isnan(sum)) {
        std::printf("FAIL %lld mismatches, %lld NaN of %lld, sum %g\n", bad, nan, n, sum);
        return 1;
    }
    std::printf("OK dim %d order %d: %lld points, %lld outside, values and node indices identical\n", DIM, ORDER, n, nan);
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 5) {
        std::printf("FAIL usage\n");
        return 2;
    }
    baobzi_t f = baobzi_restore(argv[1]);
    if (!f) {
        std::printf("FAIL restore\n");
        return 1;
    }
    const int dim = std::atoi(argv[2]), order = std::atoi(argv[3]);
    const long long n = std::atoll(argv[4]);
    int rc = 2;
    if (dim == 1 && order == 16)
        rc = run<1, 16>(f, n);
    else if (dim == 2 && order == 8)
        rc = run<2, 8>(f, n);
    else if (dim == 3 && order == 6)
        rc = run<3, 6>(f, n);
    else
        std::printf("FAIL unsupported (dim, order) in this test program\n");
    baobzi_free(f);
    return rc;
}
