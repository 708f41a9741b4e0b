// Does the fp64 pipe's cost of a warp instruction depend on which lanes are active?  One block per SM, W warps per scheduler,
// 8 independent DFMA chains per thread (pipe bound), lanes selected by a mask; cycles per warp-level DFMA per scheduler.
// The cooperative depenetration loop of the step kernel runs its push on all 32 lanes redundantly: if a half-empty warp
// instruction costs half the pipe time, a 16-lane service doubles what a scheduler sustains.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a --fmad=false -o fp64_lanes fp64_lanes.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2048
#define CHAINS 8
__global__ void dfma_masked(double* out, long long* cyc, unsigned mask, double a, double b) {
    double x[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; k++) x[k] = a + (threadIdx.x + k) * 1e-9;
    __syncthreads();
    const long long t0 = clock64();
    if ((mask >> (threadIdx.x & 31)) & 1u) {
#pragma unroll 4
        for (int i = 0; i < ITERS; i++) {
#pragma unroll
            for (int k = 0; k < CHAINS; k++) x[k] = __fma_rn(x[k], b, a);
        }
    }
    __syncwarp();
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0) cyc[blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32] = t1 - t0;
}
int main() {
    double* out; long long* cyc; cudaMalloc(&out, sizeof(double) * 148 * 512); cudaMalloc(&cyc, sizeof(long long) * 148 * 16);
    const unsigned masks[] = {0xffffffffu, 0x0000ffffu, 0xffff0000u, 0x55555555u, 0x000000ffu, 0x0000000fu, 0x00000001u, 0x00010001u, 0x01010101u};
    const char* names[] = {"all 32", "lanes 0-15", "lanes 16-31", "even lanes", "lanes 0-7", "lanes 0-3", "lane 0", "lanes 0,16", "lanes 0,8,16,24"};
    for (int wps = 1; wps <= 2; wps++) {
        for (int m = 0; m < 9; m++) {
            const int threads = 128 * wps;
            dfma_masked<<<148, threads>>>(out, cyc, masks[m], 1.0000001, 0.9999999);
            dfma_masked<<<148, threads>>>(out, cyc, masks[m], 1.0000001, 0.9999999);
            cudaDeviceSynchronize();
            long long h[148 * 16]; cudaMemcpy(h, cyc, sizeof(long long) * 148 * (threads / 32), cudaMemcpyDeviceToHost);
            double mean = 0; for (int k = 0; k < 148 * (threads / 32); k++) mean += (double)h[k]; mean /= 148 * (threads / 32);
            printf("%d warp(s) per scheduler, %-16s: %.2f cycles per warp DFMA per warp, %.2f per scheduler\n", wps, names[m], mean / (ITERS * CHAINS), mean / (ITERS * CHAINS) / wps);
        }
    }
    return 0;
}
