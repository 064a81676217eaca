// Micro-benchmark behind DESIGN.md 10-3: what do the RANDOM accesses of a sort-by-index evaluation of a 2D chunk cost on
// their own, with no tree access and no arithmetic?  BASELINE.json configs[3] (2D order 6, 1 M leaves, 383 MB tree) is
// evaluated in chunks of 16 Mi points: indices sorted by grid-cell bucket, then per point a 16-byte gather of x[perm[s]]
// and an 8-byte scattered store res[perm[s]].  If these two alone take most of the kernel's 1.06 ms, no reorganisation
// of the sort can bring the configuration near its streaming roofline.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o c4_floor c4_floor.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

// slot(i) = (i * A) mod 2^k: a bijection for odd A that scatters consecutive i over the whole array
__global__ void gather16_scatter8(const double2 *__restrict__ x, double *__restrict__ res, uint32_t n, uint32_t mask, uint32_t mult) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t s = (i * mult) & mask;
        const double2 p = __ldg(x + s);
        res[s] = p.x + p.y;
    }
}
__global__ void gather16_linear8(const double2 *__restrict__ x, double *__restrict__ res, uint32_t n, uint32_t mask, uint32_t mult) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double2 p = __ldg(x + ((i * mult) & mask));
        res[i] = p.x + p.y;
    }
}
__global__ void linear16_scatter8(const double2 *__restrict__ x, double *__restrict__ res, uint32_t n, uint32_t mask, uint32_t mult) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double2 p = __ldg(x + i);
        res[(i * mult) & mask] = p.x + p.y;
    }
}
__global__ void linear16_linear8(const double2 *__restrict__ x, double *__restrict__ res, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double2 p = __ldg(x + i);
        res[i] = p.x + p.y;
    }
}

template <typename F>
static float time_it(F f, int reps) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    f();
    f();
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    for (int r = 0; r < reps; ++r)
        f();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const uint32_t mult = 2654435761u | 1u;
    for (uint32_t k = 22; k <= 25; ++k) {
        const uint32_t n = 1u << k, mask = n - 1;
        double2 *x;
        double *res;
        cudaMalloc((void **)&x, (size_t)n * 16);
        cudaMalloc((void **)&res, (size_t)n * 8);
        cudaMemset(x, 0, (size_t)n * 16);
        const int grid = sms * 8, thr = 256, reps = 10;
        printf("%u points: x %.0f MB, res %.0f MB\n", n, n * 16 / 1e6, n * 8 / 1e6);
        float t;
        t = time_it([&] { linear16_linear8<<<grid, thr>>>(x, res, n); }, reps);
        printf("  coalesced 16 B read + coalesced 8 B write   %8.1f us  %.3e points/s  (%.0f GB/s)\n", t * 1e3, n / (t * 1e-3), n * 24 / (t * 1e-3) / 1e9);
        t = time_it([&] { gather16_linear8<<<grid, thr>>>(x, res, n, mask, mult); }, reps);
        printf("  random 16 B gather   + coalesced 8 B write   %8.1f us  %.3e points/s\n", t * 1e3, n / (t * 1e-3));
        t = time_it([&] { linear16_scatter8<<<grid, thr>>>(x, res, n, mask, mult); }, reps);
        printf("  coalesced 16 B read + random 8 B scatter     %8.1f us  %.3e points/s\n", t * 1e3, n / (t * 1e-3));
        t = time_it([&] { gather16_scatter8<<<grid, thr>>>(x, res, n, mask, mult); }, reps);
        printf("  random 16 B gather   + random 8 B scatter     %8.1f us  %.3e points/s\n", t * 1e3, n / (t * 1e-3));
        cudaFree(x);
        cudaFree(res);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
