// Latency of the segment primitives as the step kernel uses them (single warp, dependent chain).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a --fmad=false -I. -o prims tools/microbench/prims.cu
#include "../../nclone_b200/csrc/pool.cuh"
#include <cstdio>
#define K 256
__device__ uint32_t mk_lin(int x1, int y1, int x2, int y2) { return x1 | (y1 << 7) | (x2 << 13) | (y2 << 20); }
__device__ uint32_t mk_arc(int cx, int cy, int hn, int vn, int cv) { return NPP_SEG_ARC | cx | (cy << 7) | (hn << 13) | (vn << 14) | (cv << 15); }
template <int OP> __global__ void k(double* out, long long* cyc, int arcmask, double x0, double y0) {
    int lane = threadIdx.x;
    uint32_t s = ((arcmask >> lane) & 1) ? mk_arc(34, 24, 1, 0, 1) : mk_lin(33, 24, 35, 24);      // floor segment y=288 / arc
    double x = x0 + lane * 0.01, y = y0, acc = 0;
    long long t0 = clock64();
    #pragma unroll 1
    for (int i = 0; i < K; i++) {
        if (OP == 0) { double d2, a, b; int r; bool ix; bool ok = closest_item(s, x, y, d2, a, b, r, ix); acc += ok ? d2 : 0.0; x += acc * 1e-300; }
        if (OP == 1) { bool back; double a, b; seg_closest_point(s, x, y, back, a, b); acc += a + b; x += acc * 1e-300; }
        if (OP == 2) { double t = seg_toi_guarded(s, x, y, 0.5, 1.0, 1.25, 5.0, 25.0); acc += t; x += acc * 1e-300; }
        if (OP == 3) { double sx0, sy0, sx1, sy1; seg_bounds(s, sx0, sy0, sx1, sy1); acc += sx0 + sy1; s ^= (__double2loint(acc) & 1) << 30; }
        if (OP == 4) { SegLin l = dec_lin(s); acc += l.x1 + l.y2; s ^= (__double2loint(acc) & 1) << 30; }
        if (OP == 5) { acc += SQ(x); x += acc * 1e-300; }
        if (OP == 6) { bool ix = false; acc += sq_flag(x, ix); x += (ix ? 1e-300 : 2e-300) * acc; }
        if (OP == 7) { acc += (double)floordiv24(x); x += acc * 1e-300; }
        if (OP == 8) { double w = warp_min_f64(x); acc += w; x += acc * 1e-300; }
        if (OP == 9) { double u = x / 576.0; if (0 > u) u = 0; if (1 < u) u = 1; acc += u; x += acc * 1e-300; }
        if (OP == 10) { double d = sqrt(x * x + y * y); double inv = 1.0 / d; acc += inv; x += acc * 1e-300; }
    }
    long long t1 = clock64();
    out[lane] = acc + x;
    if (lane == 0) *cyc = t1 - t0;
}
// the cooperative closest-point iteration on the wedge of bench level 32 (11 cached segments, one arc in quadrant)
__global__ void coop(double* out, long long* cyc, double x0, double y0) {
    __shared__ WarpPool wp;
    int lane = threadIdx.x;
    uint32_t segs[11] = {mk_lin(67, 24, 66, 24), mk_lin(66, 24, 66, 25), mk_lin(67, 25, 67, 24), mk_lin(67, 26, 67, 25), mk_arc(70, 22, 1, 0, 1),
                         mk_lin(69, 24, 68, 26), mk_lin(68, 26, 69, 26), mk_lin(70, 24, 69, 24), mk_lin(69, 26, 70, 26), mk_lin(70, 25, 70, 24), mk_lin(70, 26, 70, 25)};
    const int ns = 11;
    SegC g = seg_decode(lane < ns ? segs[lane] : 0u);
    Ctx c; c.cell_hdr = nullptr; c.segs = nullptr;
    double x = x0, y = y0, acc = 0;
    long long t0 = clock64();
    #pragma unroll 1
    for (int i = 0; i < K; i++) {
        double a = 0, b = 0;
        int r = warp_closest_point(wp, c, 0, ns, g, 0, 0, 0, 0, x, y, a, b);
        acc += a + b + r; x += acc * 1e-300;
    }
    long long t1 = clock64();
    out[lane] = acc;
    if (lane == 0) *cyc = t1 - t0;
}
void run_coop(const char* name, double x0, double y0) {
    double* out; long long* cyc; cudaMalloc(&out, 256); cudaMalloc(&cyc, 8);
    coop<<<1, 32>>>(out, cyc, x0, y0); coop<<<1, 32>>>(out, cyc, x0, y0);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s %8.1f cycles\n", name, (double)h / K);
}
template <int OP> void run(const char* name, int arcmask, double x0, double y0) {
    double* out; long long* cyc; cudaMalloc(&out, 256); cudaMalloc(&cyc, 8);
    k<OP><<<1, 32>>>(out, cyc, arcmask, x0, y0); k<OP><<<1, 32>>>(out, cyc, arcmask, x0, y0);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s %8.1f cycles\n", name, (double)h / K);
}
int main() {
    run_coop("coop closest point, wedge (810, 280)", 810.0, 280.0);
    run_coop("coop closest point, wedge (810.3, 279.9)", 810.3, 279.9);
    run_coop("coop closest point, far (700, 100)", 700.0, 100.0);
    run<0>("closest_item linear (touching floor)", 0, 400.0, 280.0);
    run<0>("closest_item linear (AABB reject)", 0, 400.0, 200.0);
    run<0>("closest_item arc in quadrant", -1, 400.0, 295.0);
    run<0>("closest_item mixed arc/linear lanes", 0x55555555, 400.0, 283.0);
    run<1>("seg_closest_point linear", 0, 400.0, 280.0);
    run<1>("seg_closest_point arc in quadrant", -1, 400.0, 295.0);
    run<2>("seg_toi_guarded linear (far)", 0, 400.0, 200.0);
    run<2>("seg_toi_guarded linear (near)", 0, 400.0, 281.0);
    run<2>("seg_toi_guarded arc", -1, 400.0, 295.0);
    run<3>("seg_bounds", 0, 0, 0);
    run<4>("dec_lin", 0, 0, 0);
    run<5>("SQ exact (random-ish x)", 0, 7.123456789, 0);
    run<6>("sq_flag", 0, 7.123456789, 0);
    run<7>("floordiv24 + i2d", 0, 800.0, 0);
    run<8>("warp_min_f64", 0, 800.0, 0);
    run<9>("x/576 + clamp", 0, 300.0, 0);
    run<10>("sqrt(x*x+y*y), 1/d", 0, 3.0, 4.0);
    return 0;
}
