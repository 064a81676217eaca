// Probe: is mma.sync.m8n8k4.f64 on sm_100a bit-identical to the sequential chain
//   d = c; for k = 0..3: d = fma(a[m][k], b[k][n], d)   (round to nearest after every step)?
// Random matrices over a wide exponent range, two chained MMAs (K = 8), C = -0.0 start, first column of A = 1.0.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_order dmma_order.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

__global__ void dmma_k8(const double *A, const double *B, double *D, int ncase) {
    // per case: A[8][8] row-major (m, k), B[8][8] (k, n), D[8][8] (m, n)
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    for (int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < ncase; c += gridDim.x * (blockDim.x >> 5)) {
        const double *a = A + 64 * c, *b = B + 64 * c;
        double d0 = -0.0, d1 = -0.0;
        for (int kb = 0; kb < 2; ++kb) {
            const double fa = a[g * 8 + 4 * kb + t], fb = b[(4 * kb + t) * 8 + g];
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(d0), "+d"(d1) : "d"(fa), "d"(fb));
        }
        D[64 * c + g * 8 + 2 * t] = d0;
        D[64 * c + g * 8 + 2 * t + 1] = d1;
    }
}

static double rnd_wide() {
    const double m = 1.0 + drand48();
    const int e = (int)(drand48() * 40) - 20;
    return (drand48() < 0.5 ? -1 : 1) * ldexp(m, e);
}

int main() {
    const int n = 200000;
    double *hA = (double *)malloc(sizeof(double) * 64 * n), *hB = (double *)malloc(sizeof(double) * 64 * n),
           *hD = (double *)malloc(sizeof(double) * 64 * n);
    srand48(7);
    for (int c = 0; c < n; ++c)
        for (int i = 0; i < 64; ++i) {
            const int mode = c % 4;
            // mode 0: wide range; 1: Chebyshev-like (|a| <= 1, a[m][0] = 1); 2: cancellation-heavy; 3: with zeros / signed zeros
            double a = rnd_wide(), b = rnd_wide();
            if (mode == 1) { a = 2 * drand48() - 1; if (i % 8 == 0) a = 1.0; b = rnd_wide() * 1e-3; }
            if (mode == 2) { a = 1.0 + ldexp(drand48() - 0.5, -20); b = (i % 2 ? -1 : 1) * (1.0 + ldexp(drand48() - 0.5, -22)); }
            if (mode == 3) { if (drand48() < 0.3) a = drand48() < 0.5 ? 0.0 : -0.0; if (drand48() < 0.3) b = drand48() < 0.5 ? 0.0 : -0.0; }
            hA[64 * c + i] = a; hB[64 * c + i] = b;
        }
    double *dA, *dB, *dD;
    cudaMalloc(&dA, sizeof(double) * 64 * n); cudaMalloc(&dB, sizeof(double) * 64 * n); cudaMalloc(&dD, sizeof(double) * 64 * n);
    cudaMemcpy(dA, hA, sizeof(double) * 64 * n, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, sizeof(double) * 64 * n, cudaMemcpyHostToDevice);
    dmma_k8<<<296, 256>>>(dA, dB, dD, n);
    cudaMemcpy(hD, dD, sizeof(double) * 64 * n, cudaMemcpyDeviceToHost);
    long bad = 0, badrev = 0, total = 0;
    for (int c = 0; c < n; ++c)
        for (int m = 0; m < 8; ++m)
            for (int nn = 0; nn < 8; ++nn) {
                double d = -0.0, r = -0.0;
                for (int k = 0; k < 8; ++k) d = fma(hA[64 * c + m * 8 + k], hB[64 * c + k * 8 + nn], d);
                for (int k = 7; k >= 0; --k) r = fma(hA[64 * c + m * 8 + k], hB[64 * c + k * 8 + nn], r);
                const double got = hD[64 * c + m * 8 + nn];
                ++total;
                if (memcmp(&got, &d, 8)) { if (bad < 5) printf("case %d (%d,%d): dmma %a chain %a\n", c, m, nn, got, d); ++bad; }
                if (memcmp(&got, &r, 8)) ++badrev;
            }
    printf("DMMA vs sequential FMA chain (k ascending): %ld mismatches of %ld outputs; vs descending chain: %ld mismatches\n", bad, total, badrev);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return bad != 0;
}
