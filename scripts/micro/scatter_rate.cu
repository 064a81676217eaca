// Micro-benchmark: how fast does a B200 take SCATTERED stores of 4 / 8 / 16 / 32 bytes, and scattered 24-byte gathers?
// (What bounds the fused sort's perm[slot] = i phase: the per-SM request rate, or partial-sector writes in L2?)
// slot(i) = (i * A) mod 2^k is a bijection for odd A, and scatters consecutive i over the whole array.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scatter_rate scatter_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int BYTES>
__global__ void scatter_kernel(unsigned char *__restrict__ out, uint32_t n, uint32_t mask, uint32_t mult) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t slot = (i * mult) & mask;
        unsigned char *p = out + (size_t)slot * BYTES;
        if (BYTES == 4) {
            *reinterpret_cast<uint32_t *>(p) = i;
        } else if (BYTES == 8) {
            *reinterpret_cast<uint2 *>(p) = make_uint2(i, i);
        } else if (BYTES == 16) {
            *reinterpret_cast<uint4 *>(p) = make_uint4(i, i, i, i);
        } else if (BYTES == 32) {
            asm volatile("st.global.v4.b64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"((uint64_t)i), "l"((uint64_t)i),
                         "l"((uint64_t)i), "l"((uint64_t)i)
                         : "memory");
        } else { // 33: 32 bytes as two 16-byte stores
            *reinterpret_cast<uint4 *>(p - (size_t)slot) = make_uint4(i, i, i, i);
            *reinterpret_cast<uint4 *>(p - (size_t)slot + 16) = make_uint4(i, i, i, i);
        }
    }
}

// coalesced stores of the same volume, for reference
__global__ void linear_kernel(uint32_t *__restrict__ out, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        out[i] = i;
}

// scattered gather of 24-byte points (3 x 8-byte loads), result summed into a coalesced 8-byte store
__global__ void gather24_kernel(const double *__restrict__ x, double *__restrict__ res, uint32_t n, uint32_t mask,
                                uint32_t mult) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t slot = (i * mult) & mask;
        const double *p = x + 3ull * slot;
        res[i] = __ldg(p) + __ldg(p + 1) + __ldg(p + 2);
    }
}
// same with 32-byte aligned records, one 256-bit load
__global__ void gather32_kernel(const double *__restrict__ x, double *__restrict__ res, uint32_t n, uint32_t mask,
                                uint32_t mult) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t slot = (i * mult) & mask;
        const double *p = x + 4ull * slot;
        double a, b, c, d;
        asm volatile("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
        res[i] = a + b + c + d;
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
    const uint32_t k = 23, n = 1u << k, mask = n - 1, mult = 2654435761u | 1u;
    unsigned char *buf;
    double *res;
    cudaMalloc((void **)&buf, (size_t)n * 32 + 64);
    cudaMalloc((void **)&res, (size_t)n * 8);
    cudaMemset(buf, 0, (size_t)n * 32 + 64);
    const int reps = 20;
    for (int ctas = 1; ctas <= 4; ctas *= 2) {
        const int grid = sms * ctas, thr = 512;
        printf("grid %d x %d, %u elements\n", grid, thr, n);
        float t;
        t = time_it([&] { linear_kernel<<<grid, thr>>>((uint32_t *)buf, n); }, reps);
        printf("  linear 4 B stores        %8.1f us  %.3e /s\n", t * 1e3, n / (t * 1e-3));
        t = time_it([&] { scatter_kernel<4><<<grid, thr>>>(buf, n, mask, mult); }, reps);
        printf("  scattered 4 B stores     %8.1f us  %.3e /s\n", t * 1e3, n / (t * 1e-3));
        t = time_it([&] { scatter_kernel<8><<<grid, thr>>>(buf, n, mask, mult); }, reps);
        printf("  scattered 8 B stores     %8.1f us  %.3e /s\n", t * 1e3, n / (t * 1e-3));
        t = time_it([&] { scatter_kernel<16><<<grid, thr>>>(buf, n, mask, mult); }, reps);
        printf("  scattered 16 B stores    %8.1f us  %.3e /s\n", t * 1e3, n / (t * 1e-3));
        t = time_it([&] { scatter_kernel<32><<<grid, thr>>>(buf, n, mask, mult); }, reps);
        printf("  scattered 32 B stores    %8.1f us  %.3e /s  (one 256-bit store)\n", t * 1e3, n / (t * 1e-3));
        t = time_it([&] { scatter_kernel<33><<<grid, thr>>>(buf, n, mask, mult); }, reps);
        printf("  scattered 2 x 16 B       %8.1f us  %.3e /s\n", t * 1e3, n / (t * 1e-3));
        t = time_it([&] { gather24_kernel<<<grid, thr>>>((const double *)buf, res, n, mask, mult); }, reps);
        printf("  gather 24 B (3 loads)    %8.1f us  %.3e /s\n", t * 1e3, n / (t * 1e-3));
        t = time_it([&] { gather32_kernel<<<grid, thr>>>((const double *)buf, res, n, mask, mult); }, reps);
        printf("  gather 32 B (1 load)     %8.1f us  %.3e /s\n", t * 1e3, n / (t * 1e-3));
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
