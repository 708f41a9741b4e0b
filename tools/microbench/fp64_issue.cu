// Compute-side ceilings for the step kernel's roofline ("measure with a DFMA micro-benchmark since
// MEASURED_PEAKS.json has no fp64 entry"): whole-GPU throughput of independent DFMA chains (the fp64 pipe) and of independent
// integer IMADs (the issue slots), timed with CUDA events.  Prints one JSON line; bench.py reads the committed copy
// profiles/compute_peaks.json.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a --fmad=false -o fp64_issue fp64_issue.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
#define CHAINS 8
__global__ void __launch_bounds__(256) dfma_kernel(double* out, double a, double b) {
    double x[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; k++) x[k] = a + (threadIdx.x + k) * 1e-9;
#pragma unroll 4
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int k = 0; k < CHAINS; k++) x[k] = __fma_rn(x[k], b, a);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) imad_kernel(int* out, int a, int b) {
    int x[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; k++) x[k] = a + threadIdx.x + k;
#pragma unroll 4
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int k = 0; k < CHAINS; k++) x[k] = x[k] * b + a;
    }
    int s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <class F> float time_ms(F&& launch) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); launch();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    return best;
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const int blocks = p.multiProcessorCount * 8, threads = 256;
    double* outd; int* outi; cudaMalloc(&outd, sizeof(double) * blocks * threads); cudaMalloc(&outi, sizeof(int) * blocks * threads);
    const float ms_d = time_ms([&] { dfma_kernel<<<blocks, threads>>>(outd, 1.0000001, 0.9999999); });
    const float ms_i = time_ms([&] { imad_kernel<<<blocks, threads>>>(outi, 3, 5); });
    const double warp_inst = (double)blocks * (threads / 32) * ITERS * CHAINS;
    const double dfma_wips = warp_inst / (ms_d * 1e-3), imad_wips = warp_inst / (ms_i * 1e-3);
    const double hz = clk_khz * 1e3;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_max_mhz\": %.0f, \"dfma_warp_inst_per_s\": %.4e, \"dfma_tflops\": %.2f, "
           "\"dfma_warp_inst_per_cycle_per_sm\": %.3f, \"imad_warp_inst_per_s\": %.4e, \"imad_warp_inst_per_cycle_per_sm\": %.3f, "
           "\"issue_slots_per_s_nominal\": %.4e}\n",
           p.name, p.multiProcessorCount, hz / 1e6, dfma_wips, dfma_wips * 64 / 1e12, dfma_wips / hz / p.multiProcessorCount,
           imad_wips, imad_wips / hz / p.multiProcessorCount, 4.0 * p.multiProcessorCount * hz);
    return 0;
}
