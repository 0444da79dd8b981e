// Latency microbenchmarks for the dependent chains of the BIG-tier refactorisation (one warp, clock64).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o variants/microbench tools/microbench.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __noinline__ double inv_sqrt_ieee(const double x) { return 1.0 / sqrt(x); }
__device__ __forceinline__ double inv_sqrt_inl(const double x) { return 1.0 / sqrt(x); }
__global__ void k(double* out, long long* cyc, int iters, double seed) {
    double x = seed + threadIdx.x * 1e-3;
    long long t0, t1;
    // 0: out-of-line 1/sqrt chain
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = inv_sqrt_ieee(x) + 1.5;
    t1 = clock64(); if (threadIdx.x == 0) cyc[0] = t1 - t0;
    // 1: inlined
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = inv_sqrt_inl(x) + 1.5;
    t1 = clock64(); if (threadIdx.x == 0) cyc[1] = t1 - t0;
    // 2: rsqrt() (not IEEE)
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = rsqrt(x) + 1.5;
    t1 = clock64(); if (threadIdx.x == 0) cyc[2] = t1 - t0;
    // 3: dependent fma chain
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = fma(x, 0.999, 1.5);
    t1 = clock64(); if (threadIdx.x == 0) cyc[3] = t1 - t0;
    // 4: double shuffle chain
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 1) & 31) + 1.0;
    t1 = clock64(); if (threadIdx.x == 0) cyc[4] = t1 - t0;
    // 5: sqrt alone
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = sqrt(x) + 1.5;
    t1 = clock64(); if (threadIdx.x == 0) cyc[5] = t1 - t0;
    // 6: reciprocal alone
    t0 = clock64();
    for (int i = 0; i < iters; ++i) x = 1.0 / x + 1.5;
    t1 = clock64(); if (threadIdx.x == 0) cyc[6] = t1 - t0;
    // 7: shared-memory store -> load -> fma chain
    __shared__ double sh[64];
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { sh[threadIdx.x] = x; __syncwarp(); x = fma(sh[(threadIdx.x + 1) & 31], 0.5, 1.0); __syncwarp(); }
    t1 = clock64(); if (threadIdx.x == 0) cyc[7] = t1 - t0;
    // 8: fp32 rsqrt seed + double Newton (what a hand-written correctly rounded path would start from)
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { float y0 = rsqrtf((float)x); double y = (double)y0; y = y * fma(-0.5 * x, y * y, 1.5); y = y * fma(-0.5 * x, y * y, 1.5); x = y + 1.5; }
    t1 = clock64(); if (threadIdx.x == 0) cyc[8] = t1 - t0;
    out[threadIdx.x] = x;
}
int main() {
    double* out; long long* cyc; cudaMalloc(&out, 32 * 8); cudaMalloc(&cyc, 16 * 8);
    const int iters = 4096;
    for (int rep = 0; rep < 2; ++rep) k<<<1, 32>>>(out, cyc, iters, 2.0);
    long long h[16]; cudaMemcpy(h, cyc, 16 * 8, cudaMemcpyDeviceToHost);
    const char* names[] = {"1/sqrt out of line (+add)", "1/sqrt inlined (+add)", "rsqrt() (+add)", "dependent fma", "double shfl (+add)", "sqrt (+add)", "1/x (+add)", "sts -> syncwarp -> lds -> fma -> syncwarp", "rsqrtf seed + 2 Newton (+add)"};
    for (int e = 0; e < 9; ++e) printf("%-45s %.1f cycles per link\n", names[e], (double)h[e] / iters);
    return 0;
}
