// tile_tables.h -- the reference's tile definition tables (tile_definitions.py:84-223), shared by the host map
// compiler (map_compiler.cpp) and the device map compiler (map_compile_device.cu, where they sit in __constant__).
#pragma once
#include <stdint.h>
#include <math.h>
#ifndef NPP_TABLE_QUAL
#define NPP_TABLE_QUAL static const
#endif
#ifndef NPP_HD
#ifdef __CUDACC__
#define NPP_HD __host__ __device__ inline
#else
#define NPP_HD inline
#endif
#endif

// tile_definitions.py:84-132 (grid edges) and 137-185 (oriented orthogonal half-segments),
// one row per tile id 0..33 (34..37 are skipped by the reference), bit-packed would be
// unreadable so they stay as rows: 6 horizontal halves (left->right, top->bottom) then
// 6 vertical halves (top->bottom, left->right).
NPP_TABLE_QUAL int8_t kEdge[34][12] = {
    {0,0,0,0,0,0, 0,0,0,0,0,0}, {1,1,0,0,1,1, 1,1,0,0,1,1}, {1,1,1,1,0,0, 1,0,0,0,1,0}, {0,1,0,0,0,1, 0,0,1,1,1,1},
    {0,0,1,1,1,1, 0,1,0,0,0,1}, {1,0,0,0,1,0, 1,1,1,1,0,0}, {1,1,0,1,1,0, 1,1,0,1,1,0}, {1,1,1,0,0,1, 1,0,0,1,1,1},
    {0,1,1,0,1,1, 0,1,1,0,1,1}, {1,0,0,1,1,1, 1,1,1,0,0,1}, {1,1,0,0,1,1, 1,1,0,0,1,1}, {1,1,0,0,1,1, 1,1,0,0,1,1},
    {1,1,0,0,1,1, 1,1,0,0,1,1}, {1,1,0,0,1,1, 1,1,0,0,1,1}, {1,1,0,1,1,0, 1,1,0,1,1,0}, {1,1,1,0,0,1, 1,0,0,1,1,1},
    {0,1,1,0,1,1, 0,1,1,0,1,1}, {1,0,0,1,1,1, 1,1,1,0,0,1}, {1,1,1,1,0,0, 1,0,0,0,1,0}, {1,1,1,1,0,0, 1,0,0,0,1,0},
    {0,0,1,1,1,1, 0,1,0,0,0,1}, {0,0,1,1,1,1, 0,1,0,0,0,1}, {1,1,0,0,1,1, 1,1,0,0,1,1}, {1,1,0,0,1,1, 1,1,0,0,1,1},
    {1,1,0,0,1,1, 1,1,0,0,1,1}, {1,1,0,0,1,1, 1,1,0,0,1,1}, {1,0,0,0,1,0, 1,1,1,1,0,0}, {0,1,0,0,0,1, 0,0,1,1,1,1},
    {0,1,0,0,0,1, 0,0,1,1,1,1}, {1,0,0,0,1,0, 1,1,1,1,0,0}, {1,1,0,0,1,1, 1,1,0,0,1,1}, {1,1,0,0,1,1, 1,1,0,0,1,1},
    {1,1,0,0,1,1, 1,1,0,0,1,1}, {1,1,0,0,1,1, 1,1,0,0,1,1},
};
NPP_TABLE_QUAL int8_t kOrtho[34][12] = {
    {0,0,0,0,0,0, 0,0,0,0,0,0}, {-1,-1,0,0,1,1, -1,-1,0,0,1,1}, {-1,-1,1,1,0,0, -1,0,0,0,1,0}, {0,-1,0,0,0,1, 0,0,-1,-1,1,1},
    {0,0,-1,-1,1,1, 0,-1,0,0,0,1}, {-1,0,0,0,1,0, -1,-1,1,1,0,0}, {-1,-1,0,0,0,0, -1,-1,0,0,0,0}, {-1,-1,0,0,0,0, 0,0,0,0,1,1},
    {0,0,0,0,1,1, 0,0,0,0,1,1}, {0,0,0,0,1,1, -1,-1,0,0,0,0}, {-1,-1,0,0,0,0, -1,-1,0,0,0,0}, {-1,-1,0,0,0,0, 0,0,0,0,1,1},
    {0,0,0,0,1,1, 0,0,0,0,1,1}, {0,0,0,0,1,1, -1,-1,0,0,0,0}, {-1,-1,0,0,0,0, -1,-1,0,0,0,0}, {-1,-1,0,0,0,0, 0,0,0,0,1,1},
    {0,0,0,0,1,1, 0,0,0,0,1,1}, {0,0,0,0,1,1, -1,-1,0,0,0,0}, {-1,-1,0,0,0,0, -1,0,0,0,0,0}, {-1,-1,0,0,0,0, 0,0,0,0,1,0},
    {0,0,0,0,1,1, 0,0,0,0,0,1}, {0,0,0,0,1,1, 0,-1,0,0,0,0}, {-1,-1,0,0,0,0, -1,-1,0,0,1,0}, {-1,-1,0,0,0,0, -1,0,0,0,1,1},
    {0,0,0,0,1,1, 0,-1,0,0,1,1}, {0,0,0,0,1,1, -1,-1,0,0,0,1}, {-1,0,0,0,0,0, -1,-1,0,0,0,0}, {0,-1,0,0,0,0, 0,0,0,0,1,1},
    {0,0,0,0,0,1, 0,0,0,0,1,1}, {0,0,0,0,1,0, -1,-1,0,0,0,0}, {-1,-1,0,0,1,0, -1,-1,0,0,0,0}, {-1,-1,0,0,0,1, 0,0,0,0,1,1},
    {0,-1,0,0,1,1, 0,0,0,0,1,1}, {-1,0,0,0,1,1, -1,-1,0,0,0,0},
};
// tile_definitions.py:189-223 in half-tile units (x12): diagonal end points, or arc centre + quadrant + convex
struct Slanted { int8_t kind; int8_t a, b, c, d; int8_t convex; };   // kind 0 none, 1 diagonal (a,b)->(c,d), 2 arc centre (a,b) quadrant (c,d)
NPP_TABLE_QUAL Slanted kSlanted[34] = {
    {0,0,0,0,0,0},{0,0,0,0,0,0},{0,0,0,0,0,0},{0,0,0,0,0,0},{0,0,0,0,0,0},{0,0,0,0,0,0},
    {1,0,2,2,0,0},{1,0,0,2,2,0},{1,2,0,0,2,0},{1,2,2,0,0,0},
    {2,0,0,1,1,1},{2,2,0,-1,1,1},{2,2,2,-1,-1,1},{2,0,2,1,-1,1},
    {2,2,2,-1,-1,0},{2,0,2,1,-1,0},{2,0,0,1,1,0},{2,2,0,-1,1,0},
    {1,0,1,2,0,0},{1,0,0,2,1,0},{1,2,1,0,2,0},{1,2,2,0,1,0},
    {1,0,2,2,1,0},{1,0,1,2,2,0},{1,2,0,0,1,0},{1,2,1,0,0,0},
    {1,0,2,1,0,0},{1,1,0,2,2,0},{1,2,0,1,2,0},{1,1,2,0,0,0},
    {1,1,2,2,0,0},{1,0,0,1,2,0},{1,1,0,0,2,0},{1,2,2,1,0,0},
};


NPP_HD uint32_t pack_linear(int x1, int y1, int x2, int y2) {      // units of 12 px
    return (uint32_t)x1 | ((uint32_t)y1 << 7) | ((uint32_t)x2 << 13) | ((uint32_t)y2 << 20);
}
NPP_HD uint32_t pack_arc(int cx, int cy, int hor, int ver, int convex) {
    return NPP_SEG_ARC | (uint32_t)cx | ((uint32_t)cy << 7) | ((hor < 0 ? 1u : 0u) << 13) | ((ver < 0 ? 1u : 0u) << 14) | ((convex ? 1u : 0u) << 15);
}
// clamp_cell(math.floor(v / 24)) for one axis (entities.py:233, physics.py:11-13), clamped in the double domain so
// that garbage coordinates (|v| beyond int range) land where Python's unbounded ints put them; NaN -> 0.
NPP_HD int cell_axis(double v, int hi) {
    double f = floor(v / 24);
    if (!(f >= 0)) return 0;
    return f > hi ? hi : (int)f;
}
NPP_HD int cell_of(double x, double y) { return cell_axis(x, 43) * NPP_NCY + cell_axis(y, 24); }
NPP_HD bool is_int(double v) { return v == floor(v); }
