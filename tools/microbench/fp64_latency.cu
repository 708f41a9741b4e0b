// fp64 latency / throughput probe for sm_100a (one warp for latency, full SM for throughput).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a --fmad=false -o fp64_latency fp64_latency.cu
#include <cstdio>
#include <cuda_runtime.h>
#define N 4096
template <int OP> __global__ void lat(double* out, long long* cyc, double a, double b) {
    double x = a + threadIdx.x * 1e-9, y = b;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; i++) {
        if (OP == 0) x = __fma_rn(x, y, y);
        if (OP == 1) x = x + y;
        if (OP == 2) x = x * y;
        if (OP == 3) x = y / x;
        if (OP == 4) x = sqrt(x) + y;
        if (OP == 5) x = (double)(__double2int_rd(x)) + y;
        if (OP == 6) x = __shfl_xor_sync(0xffffffffu, x, 1);
        if (OP == 7) x = fmin(x, y) + 1e-3;
        if (OP == 8) { float f = (float)x; f = f * 1.0001f + 0.5f; x = (double)f; }
        if (OP == 9) x = (x < y) ? x + 1.0 : x - 1.0;
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char* name, int blocks, int threads) {
    double* out; long long* cyc; cudaMalloc(&out, sizeof(double) * blocks * threads); cudaMalloc(&cyc, sizeof(long long) * blocks);
    lat<OP><<<blocks, threads>>>(out, cyc, 1.000001, 0.9999999);
    lat<OP><<<blocks, threads>>>(out, cyc, 1.000001, 0.9999999);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-28s blocks %4d x %4d thr: %8.2f cycles/op/warp  -> %7.2f warp-ops/cycle/SM\n", name, blocks, threads, (double)h / N, (double)(threads / 32) * N / (double)h);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    const char* names[] = {"dfma (dependent)", "dadd", "dmul", "ddiv", "dsqrt+dadd", "d2i+i2d+dadd", "shfl f64", "fmin+dadd", "d2f,ffma,f2d", "dsetp+sel dadd"};
#define RUN(k) run<k>(names[k], 1, 32); run<k>(names[k], 1, 128); run<k>(names[k], 1, 512); run<k>(names[k], 1, 1024);
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7) RUN(8) RUN(9)
    return 0;
}
