// Instruction-fetch throughput probe: a straight-line body of NI independent-ish IMADs (16 B each) executed
// REP times by W warps on one SM.  Small bodies hit the per-SMSP L0 i-cache; large ones stream from L1.5/L2.
#include <cstdio>
#include <cuda_runtime.h>
template <int NI> __global__ void body(int* out, long long* cyc, int rep, int a, int b) {
    int x0 = threadIdx.x + a, x1 = a * 3, x2 = b, x3 = b + 7;
    long long t0 = clock64();
    #pragma unroll 1
    for (int r = 0; r < rep; r++) {
#pragma unroll
        for (int i = 0; i < NI / 4; i++) { x0 = x0 * a + b; x1 = x1 * b + a; x2 = x2 * a + x3; x3 = x3 * b + x0; }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int NI> void run(int warps) {
    int* out; long long* cyc; cudaMalloc(&out, 4 * 1024); cudaMalloc(&cyc, 8);
    int rep = (1 << 22) / NI;
    body<NI><<<1, warps * 32>>>(out, cyc, rep, 3, 5); body<NI><<<1, warps * 32>>>(out, cyc, rep, 3, 5);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    double instr = (double)NI * rep * warps;
    printf("body %6d instr (%5d KB) warps %2d: %.3f warp-instr/cycle/SM\n", NI, NI * 16 / 1024, warps, instr / (double)h);
    cudaFree(out); cudaFree(cyc);
}
int main() {
#define R(N) run<N>(1); run<N>(4); run<N>(8); run<N>(16);
    R(256) R(1024) R(2048) R(4096) R(8192) R(16384)
    return 0;
}
