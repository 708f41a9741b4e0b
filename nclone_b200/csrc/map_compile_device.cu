// map_compile_device.cu -- SURVEY.md 8(f) f3: Simulator.load's static part for a whole batch of levels, on the device.
//
// One thread block compiles one level: map_data row (map_loader format, float64) -> one blob (map_blob.h),
// byte-for-byte the blob the host compiler (map_compiler.cpp) emits for the same row.  Curriculum training loads
// a fresh generated level per episode (env_map_loader.py:163-204): at 65,536 envs that is 65,536 host compiles
// (~0.1 ms each) per turnover in the reference's design, one ~1 ms launch here.
//
// Reference work replaced (paths under /root/reference/nclone/):
//   MapLoader.load_map_tiles                          map_loader.py:18-41
//   TileSegmentFactory.create_segments_for_simulator  utils/tile_segment_factory.py:169-344
//   SpatialSegmentIndex._build_index                  utils/spatial_segment_index.py:69-110
//   Simulator.reset_map_tile_data                     nsim.py:194-219
//   MapLoader.load_map_entities + entity constructors map_loader.py:69-145, utils/entity_factory.py:145-233
//
// Organisation (integer / byte work, bound by the 41 KB blob write per level):
//   * the 44x25 tile ids sit in shared memory; everything derived from tiles is a *gather*: a half-cell edge
//     (x, y) of the 89x51 grids receives contributions from at most two tiles, so edge bitmaps and the ortho
//     segment states are recomputed where they are needed instead of being scattered with atomics;
//   * one thread per cell emits the cell's ordered segment row (own diagonal / arc, horizontal halves in x-major
//     half-cell order, vertical halves) as two 16-byte stores plus the header word;
//   * entity records have a data-dependent stride (doors: 9 or 10 values, map_loader.py:126-141), so lane 0 of the
//     last warp walks them serially into shared-memory drafts while the other warps do the tile work; every later
//     ordering (entity_dic order, dynamic slots, door records, per-cell static lists in creation and in entity_dic
//     order) is a rank computed per draft in parallel.
#include <cuda_runtime.h>
#include <atomic>
#include <mutex>
#include <cstdio>

#include "map_blob.h"
#define NPP_TABLE_QUAL __constant__ const
#include "tile_tables.h"
#include "../../include/nppsim.h"

namespace {

constexpr int MC_THREADS = 256;
constexpr int MC_MAX_ENT = NPPSIM_DEVICE_COMPILE_MAX_ENT;
constexpr int MC_STAGE = 512;              // doubles of the entity section staged in shared memory (102 plain records)
constexpr uint32_t MC_FRONT = 112 + NPP_NCELL * 4 + NPP_NCELL * NPP_SEGS_PER_CELL * 4 + 2 * 576;   // header .. vedge, all 16-byte aligned

struct Draft {                 // one entity in creation order (40 bytes)
    double x, y;
    uint8_t kind, type, p0, p1;
    uint8_t key, roaming, has_door, vertical;
    uint8_t dx0, dy0, dx1, dy1;
    int16_t parent;            // exit switch: creation position of its door
    uint16_t index, cell, rank, id;
    uint16_t pad;
};
static_assert(sizeof(Draft) == 40, "Draft layout");

struct Shared {
    unsigned long long cw[NPP_NCELL];       // per cell: its tile's packed table row (pack_tables_kernel)
    uint8_t tile[NPP_NCELL];
    uint16_t cell_cnt[NPP_NCELL + 4];       // static entities per cell, then the exclusive scan (ecell_off)
    Draft d[MC_MAX_ENT];
    double stage[MC_STAGE];                 // map_data[1230 : 1230 + MC_STAGE], the entity records the serial walk reads
    int n, status, n_static, n_dyn, n_doors, n_seg;
    uint32_t kind_mask;
    int key_cnt[32], key_base[32], type_cnt[40];
    int warp_tot[MC_THREADS / 32];
};

__device__ __forceinline__ uint32_t align16(uint32_t v) { return (v + 15u) & ~15u; }

// The tile tables are indexed by a per-thread tile id; from __constant__ that serialises a warp up to 32-fold (the
// first version of this kernel spent >90 % of its time there).  They are packed once into one 64-bit word per tile id
//   bits  0-11 kEdge row | 12-23 kOrtho row > 0 | 24-35 kOrtho row < 0 | 36-37 slanted kind | 38-45 a,b,c+1,d+1 (2 bits
//   each; for diagonals c,d are stored as is) | 46 convex
// and every block spreads them over its cells in shared memory.
__device__ unsigned long long g_tile_word[34];
__global__ void pack_tables_kernel() {
    const int t = threadIdx.x;
    if (t >= 34) return;
    unsigned long long w = 0;
    for (int j = 0; j < 12; j++) {
        if (kEdge[t][j]) w |= 1ull << j;
        if (kOrtho[t][j] > 0) w |= 1ull << (12 + j);
        if (kOrtho[t][j] < 0) w |= 1ull << (24 + j);
    }
    const Slanted sl = kSlanted[t];
    w |= (unsigned long long)sl.kind << 36;
    w |= (unsigned long long)sl.a << 38 | (unsigned long long)sl.b << 40;
    if (sl.kind == 2) w |= (unsigned long long)(sl.c + 1) << 42 | (unsigned long long)(sl.d + 1) << 44;
    else w |= (unsigned long long)sl.c << 42 | (unsigned long long)sl.d << 44;
    w |= (unsigned long long)sl.convex << 46;
    g_tile_word[t] = w;
}
__device__ __forceinline__ void half_bits(unsigned long long w, int j, int& sum, int& edge) {
    edge ^= (int)(w >> j) & 1;
    sum += ((int)(w >> (12 + j)) & 1) - ((int)(w >> (24 + j)) & 1);
}
// ortho state and grid-edge parity of one horizontal half-cell edge (X in 0..88, Y in 0..50): at most two tiles contribute
__device__ __forceinline__ void hor_half(const unsigned long long* cw, int X, int Y, int& sum, int& edge) {
    sum = 0; edge = (Y == 0 || Y == NPP_NHY - 1) ? 1 : 0;                       // nsim.py:208-215
    const int x = X >> 1, c = X & 1;
    if (x >= NPP_NCX) return;
    const int y = Y >> 1, r = Y & 1;
    if (y < NPP_NCY) half_bits(cw[x * NPP_NCY + y], 2 * r + c, sum, edge);
    if (r == 0 && y >= 1) half_bits(cw[x * NPP_NCY + y - 1], 4 + c, sum, edge);
}
__device__ __forceinline__ void ver_half(const unsigned long long* cw, int X, int Y, int& sum, int& edge) {
    sum = 0; edge = (X == 0 || X == NPP_NHX - 1) ? 1 : 0;
    const int y = Y >> 1, r = Y & 1;
    if (y >= NPP_NCY) return;
    const int x = X >> 1, c = X & 1;
    if (x < NPP_NCX) half_bits(cw[x * NPP_NCY + y], 6 + 2 * c + r, sum, edge);
    if (c == 0 && x >= 1) half_bits(cw[(x - 1) * NPP_NCY + y], 10 + r, sum, edge);
}
__device__ __forceinline__ void seg_box(uint32_t s, int& x0, int& y0, int& x1, int& y1) {   // entities.py:36-41, 121-125
    int ax, ay, bx, by;
    if (s & NPP_SEG_ARC) { ax = s & 127; ay = (s >> 7) & 63; bx = ax + (((s >> 13) & 1) ? -2 : 2); by = ay + (((s >> 14) & 1) ? -2 : 2); }
    else { ax = s & 127; ay = (s >> 7) & 63; bx = (s >> 13) & 127; by = (s >> 20) & 63; }
    x0 = min(ax, bx); x1 = max(ax, bx); y0 = min(ay, by); y1 = max(ay, by);
}
// physics.py:317-332: the orientation dict has keys 0..7 only
__device__ __forceinline__ bool orientation_vec(double o, double& vx, double& vy) {
    if (!is_int(o) || o < 0 || o > 7) return false;
    const double d = 0.70710678118654757;          // sqrt(2)/2 as the host computes it
    const int k = (int)o;
    vx = (k == 0) ? 1.0 : (k == 4) ? -1.0 : (k == 2 || k == 6) ? 0.0 : (k == 1 || k == 7) ? d : -d;
    vy = (k == 2) ? 1.0 : (k == 6) ? -1.0 : (k == 0 || k == 4) ? 0.0 : (k == 1 || k == 3) ? d : -d;
    return true;
}

// map_loader.py:88-141 + the entity constructors, serial (one thread)
__device__ int parse_entities(const double* __restrict__ mg, int len, Shared& S) {
    int n = 0;
    const double dexit = mg[1156];
    // a dependent global load per record would cost ~1 us each; the records were staged by the whole block
    auto M = [&](int i) -> double { const int r = i - 1230; return (r >= 0 && r < MC_STAGE) ? S.stage[r] : mg[i]; };
    if (!is_int(dexit)) return NPPSIM_E_MAP;
    auto make = [&](int kind, int type, int key, double xc, double yc) -> int {
        if (n >= MC_MAX_ENT) return -1;
        Draft& d = S.d[n];
        d.kind = (uint8_t)kind; d.type = (uint8_t)type; d.key = (uint8_t)key; d.p0 = 0; d.p1 = 0;
        d.x = xc * 6; d.y = yc * 6;
        d.index = (uint16_t)S.type_cnt[type]++;
        d.cell = (uint16_t)cell_of(d.x, d.y);
        d.roaming = 0; d.has_door = 0; d.vertical = 0; d.dx0 = d.dy0 = d.dx1 = d.dy1 = 0; d.parent = -1; d.pad = 0;
        d.rank = (uint16_t)S.key_cnt[key]++;
        return n++;
    };
    auto make_door = [&](Draft& d, double orientation, double swx, double swy) -> int {   // entity_door_base.py:52-97
        double vx, vy;
        if (!orientation_vec(orientation, vx, vy)) return NPPSIM_E_MAP;
        const bool vertical = (orientation == 0 || orientation == 4);
        const int hx = 2 * (cell_axis(d.x - 12 * vx, 43) + 1), hy = 2 * (cell_axis(d.y - 12 * vy, 24) + 1);
        int x0, y0, x1, y1;
        if (vertical) { x0 = hx; y0 = hy - 2; x1 = hx; y1 = hy - 1; }
        else { x0 = hx - 2; y0 = hy; x1 = hx - 1; y1 = hy; }
        if (x0 >= NPP_NHX || x1 >= NPP_NHX || y0 >= NPP_NHY || y1 >= NPP_NHY) return NPPSIM_E_MAP;
        d.has_door = 1; d.vertical = vertical ? 1 : 0; d.p0 = d.vertical;
        d.dx0 = (uint8_t)x0; d.dy0 = (uint8_t)y0; d.dx1 = (uint8_t)x1; d.dy1 = (uint8_t)y1;
        d.x = 6 * swx; d.y = 6 * swy;                                           // the entity moves to its switch
        d.cell = (uint16_t)cell_of(d.x, d.y);
        return 0;
    };
    int index = 1230;
    while (index < len) {
        if (index + 4 >= len) break;
        const double type = M(index), xc = M(index + 1), yc = M(index + 2), orientation = M(index + 3), mode = M(index + 4);
        const int t = (is_int(type) && type >= 0 && type < 40) ? (int)type : -1;
        int k = 0, rc = 0;
        switch (t) {
        case 1: case 21: k = make(NPP_K_MINE, t, t, xc, yc); if (k >= 0) S.d[k].p0 = (t == 1) ? 0 : 1; break;
        case 2: k = make(NPP_K_GOLD, 2, 2, xc, yc); break;
        case 3: {                                                               // entity_factory.py:185-197
            k = make(NPP_K_EXIT, 3, 3, xc, yc);
            if (k < 0) break;
            S.d[k].roaming = 1;
            const double cd = (double)index + 5.0 * dexit + 1;
            if (!(cd >= 0 && cd + 1 < len)) return NPPSIM_E_MAP;
            const int ci = (int)cd;
            const int s = make(NPP_K_EXIT_SWITCH, 4, 3, M(ci), M(ci + 1));
            if (s < 0) { k = -1; break; }
            S.d[s].roaming = 1; S.d[s].parent = (int16_t)k;
            break; }
        case 5: k = make(NPP_K_DOOR_REGULAR, 5, 5, xc, yc); if (k >= 0) rc = make_door(S.d[k], orientation, xc, yc); break;
        case 6: case 8:
            if (index + 7 >= len) return NPPSIM_E_MAP;
            k = make(t == 6 ? NPP_K_DOOR_LOCKED : NPP_K_DOOR_TRAP, t, t, xc, yc);
            if (k >= 0) rc = make_door(S.d[k], orientation, M(index + 6), M(index + 7));
            break;
        case 10: case 11: {
            double vx, vy;
            if (!orientation_vec(orientation, vx, vy)) return NPPSIM_E_MAP;
            k = make(t == 10 ? NPP_K_LAUNCH_PAD : NPP_K_ONE_WAY, t, t, xc, yc); if (k >= 0) S.d[k].p0 = (uint8_t)(int)orientation; break; }
        case 14: case 26: {                                                     // entity_drone_base.py:66-78
            const double dir = floor(orientation / 2);
            if (!(dir >= 0 && dir <= 3) || !(mode == 0 || mode == 1 || mode == 2 || mode == 3)) return NPPSIM_E_MAP;
            k = make(t == 14 ? NPP_K_DRONE_ZAP : NPP_K_MINI_DRONE, t, t, xc, yc);
            if (k >= 0) { S.d[k].p0 = (uint8_t)(int)dir; S.d[k].p1 = (uint8_t)(int)mode; S.d[k].roaming = 1; }
            break; }
        case 17: k = make(NPP_K_BOUNCE, 17, 17, xc, yc); if (k >= 0) S.d[k].roaming = 1; break;
        case 20: k = make(NPP_K_THWUMP, 20, 20, xc, yc);                        // entity_thwump.py:92-101
            if (k >= 0) { S.d[k].roaming = 1; S.d[k].p0 = (uint8_t)(((orientation == 0 || orientation == 4) ? 1 : 0) | ((orientation == 0 || orientation == 2) ? 0 : 2)); }
            break;
        case 24: k = make(NPP_K_BOOST, 24, 24, xc, yc); break;
        case 25: k = make(NPP_K_DEATH_BALL, 25, 25, xc, yc); if (k >= 0) S.d[k].roaming = 1; break;
        case 28: k = make(NPP_K_SHWUMP, 28, 28, xc, yc); if (k >= 0) S.d[k].roaming = 1; break;
        default: break;                                                         // unknown types are skipped
        }
        if (rc) return rc;
        if (k < 0) return NPPSIM_E_CAPACITY;
        if (t == 6 || t == 8) {                                                 // map_loader.py:126-141
            if (index + 9 < len) index += (M(index + 7) != 0 && M(index + 8) == 0 && M(index + 9) == 0) ? 10 : 9;
            else index += 9;
        } else index += 5;
    }
    S.n = n;
    return 0;
}

__global__ void __launch_bounds__(MC_THREADS) compile_maps_kernel(const double* __restrict__ maps, long long row_stride, const int32_t* __restrict__ map_len,
                                                                   int default_len, uint8_t* __restrict__ blobs, long long blob_stride,
                                                                   long long* __restrict__ map_off, int32_t* __restrict__ status, int32_t* __restrict__ summary) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    Shared& S = *reinterpret_cast<Shared*>(smem_raw);
    const int map = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double* __restrict__ m = maps + (size_t)map * row_stride;
    const int len = map_len ? map_len[map] : default_len;
    uint8_t* __restrict__ out = blobs + (size_t)map * blob_stride;
    if (map_off && tid == 0) map_off[map] = (long long)map * blob_stride;
    if (len < 1235 || len > row_stride) {
        if (tid == 0) { status[map] = NPPSIM_E_MAP; *reinterpret_cast<uint32_t*>(out) = 0; atomicAdd(summary + 3, 1); }
        return;
    }
    // ---- tiles into shared memory (map_loader.py:18-41; border cells solid), read in input order (coalesced) --
    for (int c = tid; c < NPP_NCELL; c += MC_THREADS) {
        const int x = c / NPP_NCY, y = c - x * NPP_NCY;
        if (x == 0 || x == NPP_NCX - 1 || y == 0 || y == NPP_NCY - 1) S.tile[c] = 1;
        S.cell_cnt[c] = 0;
    }
    for (int i = tid; i < 42 * 23; i += MC_THREADS) {
        const int y = i / 42, x = i - y * 42;
        const double v = m[184 + i];
        // ids outside the integers 1..33 behave as empty (tile_segment_factory.py:184-191, 278-283)
        S.tile[(x + 1) * NPP_NCY + y + 1] = (is_int(v) && v >= 1 && v <= 33) ? (uint8_t)(int)v : (uint8_t)0;
    }
    for (int i = tid; i < MC_STAGE && 1230 + i < len; i += MC_THREADS) S.stage[i] = m[1230 + i];
    if (tid < 32) { S.key_cnt[tid] = 0; S.key_base[tid] = 0; }
    if (tid < 40) S.type_cnt[tid] = 0;
    if (tid == 0) { S.n = 0; S.status = 0; S.n_static = 0; S.n_dyn = 0; S.n_doors = 0; S.n_seg = 0; S.kind_mask = 1u << NPP_K_NONE; S.cell_cnt[NPP_NCELL] = 0; }
    __syncthreads();
    for (int c = tid; c < NPP_NCELL; c += MC_THREADS) S.cw[c] = g_tile_word[S.tile[c]];
    __syncthreads();

    // ---- the last warp's lane 0 walks the entity records while everybody else does the tile-derived tables ---
    if (warp == MC_THREADS / 32 - 1) {
        if (lane == 0) { const int rc = parse_entities(m, len, S); if (rc) S.status = rc; }
    } else {
        const int nt = MC_THREADS - 32;
        uint32_t* __restrict__ cell_hdr = reinterpret_cast<uint32_t*>(out + 112);
        uint4* __restrict__ segs = reinterpret_cast<uint4*>(out + 112 + NPP_NCELL * 4);
        int my_seg = 0;
        for (int c = tid; c < NPP_NCELL; c += nt) {
            const int cx = c / NPP_NCY, cy = c - cx * NPP_NCY;
            uint32_t row[NPP_SEGS_PER_CELL];
#pragma unroll
            for (int k = 0; k < NPP_SEGS_PER_CELL; k++) row[k] = 0;
            int cnt = 0;
            auto push = [&](uint32_t s) {
#pragma unroll
                for (int k = 0; k < NPP_SEGS_PER_CELL; k++) if (k == cnt) row[k] = s;
                cnt++;
            };
            const unsigned long long w = S.cw[c];
            const int sk = (int)(w >> 36) & 3, sa = (int)(w >> 38) & 3, sb = (int)(w >> 40) & 3, sc = (int)(w >> 42) & 3, sd = (int)(w >> 44) & 3;
            if (sk == 1) push(pack_linear(2 * cx + sa, 2 * cy + sb, 2 * cx + sc, 2 * cy + sd));
            else if (sk == 2) push(pack_arc(2 * cx + sa, 2 * cy + sb, sc - 1, sd - 1, (int)(w >> 46) & 1));
            // tile_segment_factory.py:231-262: state s != 0 -> one segment in cell floor((coord - 0.1 s) / 2),
            // end points swapped only for s == -1; x-major half-cell order inside the cell
#pragma unroll
            for (int dx = 0; dx < 2; dx++)
#pragma unroll
                for (int dy = 0; dy < 3; dy++) {
                    const int X = 2 * cx + dx, Y = 2 * cy + dy;
                    int s, e; hor_half(S.cw, X, Y, s, e);
                    if (s != 0 && (int)floor((Y - 0.1 * s) / 2) == cy) push(s == -1 ? pack_linear(X + 1, Y, X, Y) : pack_linear(X, Y, X + 1, Y));
                }
#pragma unroll
            for (int dx = 0; dx < 3; dx++)
#pragma unroll
                for (int dy = 0; dy < 2; dy++) {
                    const int X = 2 * cx + dx, Y = 2 * cy + dy;
                    int s, e; ver_half(S.cw, X, Y, s, e);
                    if (s != 0 && (int)floor((X - 0.1 * s) / 2) == cx) push(s == -1 ? pack_linear(X, Y, X, Y + 1) : pack_linear(X, Y + 1, X, Y));
                }
            if (cnt > NPP_SEGS_PER_CELL) { S.status = NPPSIM_E_CAPACITY; cnt = NPP_SEGS_PER_CELL; }
            uint32_t hdr = 0;
            if (cnt) {
                int bx0 = 1 << 20, by0 = 1 << 20, bx1 = -1, by1 = -1;
#pragma unroll
                for (int k = 0; k < NPP_SEGS_PER_CELL; k++) if (k < cnt) {
                    int x0, y0, x1, y1; seg_box(row[k], x0, y0, x1, y1);
                    bx0 = min(bx0, x0); by0 = min(by0, y0); bx1 = max(bx1, x1); by1 = max(by1, y1);
                }
                hdr = (uint32_t)cnt | ((uint32_t)bx0 << 4) | ((uint32_t)by0 << 11) | ((uint32_t)bx1 << 17) | ((uint32_t)by1 << 24);
            }
            cell_hdr[c] = hdr;
            segs[2 * c] = make_uint4(row[0], row[1], row[2], row[3]);
            segs[2 * c + 1] = make_uint4(row[4], row[5], row[6], row[7]);
            my_seg += cnt;
        }
        if (my_seg) atomicAdd(&S.n_seg, my_seg);
        // edge bitmaps: bit x*51+y of 142 words each (+2 padding words up to the 16-byte boundary)
        uint32_t* __restrict__ hb = reinterpret_cast<uint32_t*>(out + 112 + NPP_NCELL * 4 + NPP_NCELL * NPP_SEGS_PER_CELL * 4);
        for (int w = tid; w < 2 * 144; w += nt) {
            const int vertical = w >= 144, ww = vertical ? w - 144 : w;
            uint32_t bits = 0;
            if (ww < NPP_EDGE_WORDS) {
                for (int b = 0; b < 32; b++) {
                    const int bit = ww * 32 + b;
                    if (bit >= NPP_NHX * NPP_NHY) break;
                    const int X = bit / NPP_NHY, Y = bit - X * NPP_NHY;
                    int s, e;
                    if (vertical) ver_half(S.cw, X, Y, s, e); else hor_half(S.cw, X, Y, s, e);
                    bits |= (uint32_t)(e & 1) << b;
                }
            }
            hb[w] = bits;
        }
    }
    __syncthreads();
    if (S.status) {
        if (tid == 0) { status[map] = S.status; *reinterpret_cast<uint32_t*>(out) = 0; atomicAdd(summary + 3, 1); }
        return;
    }
    // ---- ranks: entity_dic order (keys ascending, creation order inside a key; nsim.py:190-192) ---------------
    const int n = S.n;
    if (tid == 0) { int b = 0; for (int k = 1; k <= 28; k++) { S.key_base[k] = b; b += S.key_cnt[k]; } }
    __syncthreads();
    for (int k = tid; k < n; k += MC_THREADS) {
        Draft& d = S.d[k];
        d.id = (uint16_t)(S.key_base[d.key] + d.rank);
        if (!d.roaming) atomicAdd(&S.n_static, 1);
        if (d.roaming || d.kind == NPP_K_DOOR_REGULAR) atomicAdd(&S.n_dyn, 1);
        if (d.has_door) atomicAdd(&S.n_doors, 1);
        atomicOr(&S.kind_mask, 1u << d.kind);
    }
    __syncthreads();
    // static per-cell counts -> exclusive scan (ecell_off).  u16 counters: two share a 32-bit word, so count in
    // registers per thread-owned cell instead of atomics.
    for (int c = tid; c < NPP_NCELL; c += MC_THREADS) {
        int cnt = 0;
        for (int k = 0; k < n; k++) cnt += (!S.d[k].roaming && S.d[k].cell == c) ? 1 : 0;
        S.cell_cnt[c] = (uint16_t)cnt;
    }
    __syncthreads();
    {   // block-wide exclusive scan of 1100 counts: 5 consecutive cells per thread
        const int base = tid * 5;
        int v[5], sum = 0;
#pragma unroll
        for (int k = 0; k < 5; k++) { v[k] = (base + k < NPP_NCELL) ? S.cell_cnt[base + k] : 0; sum += v[k]; }
        int incl = sum;
#pragma unroll
        for (int dlt = 1; dlt < 32; dlt <<= 1) { const int o = __shfl_up_sync(0xffffffffu, incl, dlt); if (lane >= dlt) incl += o; }
        if (lane == 31) S.warp_tot[warp] = incl;
        __syncthreads();
        int prefix = incl - sum;
        for (int w = 0; w < warp; w++) prefix += S.warp_tot[w];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 5; k++) { if (base + k <= NPP_NCELL) S.cell_cnt[base + k] = (uint16_t)prefix; prefix += v[k]; }
    }
    __syncthreads();
    // ---- layout of the entity tail -------------------------------------------------------------------------
    const uint32_t n_static = (uint32_t)S.n_static, n_dyn = (uint32_t)S.n_dyn, n_doors = (uint32_t)S.n_doors;
    const uint32_t off_ents = MC_FRONT;
    const uint32_t off_ecell_off = align16(off_ents + (uint32_t)n * 32u);
    const uint32_t off_ids = align16(off_ecell_off + (NPP_NCELL + 1) * 2u);
    const uint32_t off_ids_dic = align16(off_ids + n_static * 2u);
    const uint32_t off_dyn = align16(off_ids_dic + n_static * 2u);
    const uint32_t off_doors = align16(off_dyn + n_dyn * 4u);
    const uint32_t off_tiles = align16(off_doors + n_doors * 8u);
    const uint32_t total = align16(off_tiles + NPP_NCELL);
    if ((long long)total > blob_stride) {
        if (tid == 0) { status[map] = NPPSIM_E_CAPACITY; *reinterpret_cast<uint32_t*>(out) = 0; atomicAdd(summary + 3, 1); }
        return;
    }
    // zero the tail (padding and alignment gaps must match the host's memset), then fill it
    for (uint32_t o = off_ents + tid * 16u; o < total; o += MC_THREADS * 16u) *reinterpret_cast<uint4*>(out + o) = make_uint4(0, 0, 0, 0);
    __syncthreads();
    uint16_t* __restrict__ ecell_off = reinterpret_cast<uint16_t*>(out + off_ecell_off);
    for (int c = tid; c <= NPP_NCELL; c += MC_THREADS) ecell_off[c] = S.cell_cnt[c];
    for (int c = tid; c < NPP_NCELL; c += MC_THREADS) {
        out[off_tiles + c] = S.tile[c];
    }
    NppStaticEnt* __restrict__ ents = reinterpret_cast<NppStaticEnt*>(out + off_ents);
    uint16_t* __restrict__ ids_load = reinterpret_cast<uint16_t*>(out + off_ids);
    uint16_t* __restrict__ ids_dic = reinterpret_cast<uint16_t*>(out + off_ids_dic);
    uint32_t* __restrict__ dyn_ent = reinterpret_cast<uint32_t*>(out + off_dyn);
    NppDoor* __restrict__ doors = reinterpret_cast<NppDoor*>(out + off_doors);
    for (int k = tid; k < n; k += MC_THREADS) {
        const Draft d = S.d[k];
        const bool needs_slot = d.roaming || d.kind == NPP_K_DOOR_REGULAR;
        int slot = 0, door_id = 0, pos_load = 0, pos_dic = 0;
        for (int j = 0; j < n; j++) {
            const Draft& o = S.d[j];
            const bool before = o.id < d.id;
            slot += (before && (o.roaming || o.kind == NPP_K_DOOR_REGULAR)) ? 1 : 0;
            door_id += (before && o.has_door) ? 1 : 0;
            const bool same_cell = !o.roaming && o.cell == d.cell;
            pos_load += (same_cell && j < k) ? 1 : 0;
            pos_dic += (same_cell && before) ? 1 : 0;
        }
        NppStaticEnt se;
        se.x = d.x; se.y = d.y; se.kind = d.kind; se.type = d.type; se.p0 = d.p0; se.p1 = d.p1;
        se.dyn = needs_slot ? (int16_t)slot : (int16_t)-1;
        se.aux = (d.kind == NPP_K_EXIT_SWITCH) ? (int16_t)S.d[d.parent].id : (d.has_door ? (int16_t)door_id : (int16_t)-1);
        se.cell = d.cell; se.index = d.index; se.load_order = (uint16_t)k; se.pad = 0;
        const uint4* src = reinterpret_cast<const uint4*>(&se);
        uint4* dst = reinterpret_cast<uint4*>(&ents[d.id]);
        dst[0] = src[0]; dst[1] = src[1];
        if (needs_slot) {
            const uint32_t p = (d.kind == NPP_K_DRONE_ZAP || d.kind == NPP_K_MINI_DRONE) ? d.p1 : d.p0;
            dyn_ent[slot] = (uint32_t)d.id | ((uint32_t)d.kind << 16) | (p << 24);
        }
        if (d.has_door) {
            NppDoor dr; dr.ent = d.id; dr.vertical = d.vertical; dr.x0 = d.dx0; dr.y0 = d.dy0; dr.x1 = d.dx1; dr.y1 = d.dy1; dr.pad = 0;
            doors[door_id] = dr;
        }
        if (!d.roaming) {
            const int b = S.cell_cnt[d.cell];
            ids_load[b + pos_load] = d.id;
            ids_dic[b + pos_dic] = d.id;
        }
    }
    if (tid == 0) {
        NppMapHeader h;
        h.magic = NPP_BLOB_MAGIC; h.total_bytes = total; h.n_ent = (uint32_t)n; h.n_dyn = n_dyn; h.n_doors = n_doors;
        h.n_static_cell = n_static; h.n_segments = (uint32_t)S.n_seg;
        const uint32_t edge_readers = (1u << NPP_K_DRONE_ZAP) | (1u << NPP_K_MINI_DRONE) | (1u << NPP_K_THWUMP) | (1u << NPP_K_SHWUMP);
        h.flags = (S.kind_mask & edge_readers) ? 1u : 0u;
        h.spawn_x = m[1231] * 6; h.spawn_y = m[1232] * 6; h.db_count = m[1200];
        h.off_cell_hdr = 112; h.off_segs = 112 + NPP_NCELL * 4; h.off_hedge = h.off_segs + NPP_NCELL * NPP_SEGS_PER_CELL * 4;
        h.off_vedge = h.off_hedge + 576; h.off_ents = off_ents; h.off_ecell_off = off_ecell_off; h.off_ecell_ids = off_ids;
        h.off_ecell_ids_dic = off_ids_dic; h.off_dyn_ent = off_dyn; h.off_doors = off_doors; h.off_tiles = off_tiles;
        h.kind_mask = S.kind_mask;
        uint64_t buf[14];
#pragma unroll
        for (int k = 0; k < 14; k++) buf[k] = 0;
        memcpy(buf, &h, sizeof(h));
        uint4* dst = reinterpret_cast<uint4*>(out);
#pragma unroll
        for (int k = 0; k < 7; k++) dst[k] = make_uint4((uint32_t)buf[2 * k], (uint32_t)(buf[2 * k] >> 32), (uint32_t)buf[2 * k + 1], (uint32_t)(buf[2 * k + 1] >> 32));
        status[map] = 0;
        atomicMax(summary + 0, n);
        atomicMax(summary + 1, (int)n_dyn);
        atomicOr(reinterpret_cast<unsigned int*>(summary + 2), S.kind_mask);
    }
}

}  // namespace

extern thread_local char g_nppsim_err[256];       // nppsim_kernels.cu
namespace { char (&g_err)[256] = g_nppsim_err; }

extern std::atomic<long long> g_nppsim_launches;


extern "C" int nppsim_compile_maps_device(const double* map_data, int64_t row_stride, const int32_t* map_len, int32_t default_len,
                                          int32_t n_maps, uint8_t* blobs, int64_t blob_stride, int64_t* map_off, int32_t* status,
                                          int32_t* summary, void* stream) {
    g_err[0] = 0;
    if (!map_data || !blobs || !status || !summary || n_maps <= 0) { snprintf(g_err, sizeof(g_err), "bad arguments"); return NPPSIM_E_ARG; }
    if (blob_stride < (int64_t)MC_FRONT + 3312 || (blob_stride & 15) || ((uintptr_t)blobs & 15)) {
        snprintf(g_err, sizeof(g_err), "blob_stride must be a multiple of 16 and at least nppsim_map_blob_bound(1235); blobs 16-byte aligned");
        return NPPSIM_E_ARG;
    }
    if (row_stride < 1235) { snprintf(g_err, sizeof(g_err), "row_stride shorter than 1235 values"); return NPPSIM_E_ARG; }
    // one-time setup PER DEVICE (function attributes and the __device__ tile table belong to the current device's context)
    {
        static std::mutex init_mu;
        static bool init_done[64] = {false};
        int dev = 0;
        cudaError_t ie = cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lock(init_mu);
        if (ie == cudaSuccess && (dev < 0 || dev >= 64 || !init_done[dev])) {
            ie = cudaFuncSetAttribute(compile_maps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Shared));
            if (ie == cudaSuccess) {
                pack_tables_kernel<<<1, 64>>>();      // legacy stream: ordered before any later launch on a blocking stream; synchronise for the rest
                ie = cudaGetLastError();
            }
            if (ie == cudaSuccess) ie = cudaDeviceSynchronize();
            if (ie == cudaSuccess && dev >= 0 && dev < 64) init_done[dev] = true;
        }
        if (ie != cudaSuccess) { snprintf(g_err, sizeof(g_err), "device compiler setup: %s", cudaGetErrorString(ie)); return (int)ie > 0 ? (int)ie : 1; }
    }
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(summary, 0, 4 * sizeof(int32_t), s);
    if (e == cudaSuccess) {
        compile_maps_kernel<<<n_maps, MC_THREADS, sizeof(Shared), s>>>(map_data, (long long)row_stride, map_len, default_len, blobs,
                                                                      (long long)blob_stride, (long long*)map_off, status, summary);
        g_nppsim_launches++;
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) { snprintf(g_err, sizeof(g_err), "compile_maps_kernel: %s", cudaGetErrorString(e)); return (int)e > 0 ? (int)e : 1; }
    return 0;
}
