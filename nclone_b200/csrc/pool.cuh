// pool.cuh -- warp-pooled tile-segment queries.
//
// Why: segment queries are sparse (mean 1.65 segments per query, median 0).  With one
// thread per env and a per-thread loop over "my" segments, a warp executed the per-segment maths with
// 1-3 of 32 lanes active (profiles/r01_a: 4.9 lanes on average).  Here the 32 envs of a warp put
// their (owner lane, segment) pairs into a shared-memory queue and all 32 lanes sweep that queue, so
// the expensive fp64 work runs on dense lanes; each owner then folds its own items in reference order.
//
// Exactness: SQ() (glibc pow emulation) is expensive, so the comparison-only sites use x*x plus a
// guard band: a decision is taken from the cheap value only when it is provably the same for any
// squares within 1 ulp of it; otherwise the literal reference expression with exact SQ() is evaluated.
#pragma once
#include "geom.cuh"

#ifdef NPP_PROFILE
// NPP_PROFILE: per-warp cycle counters per phase (debug builds only; tools/phase_profile.py)
__device__ unsigned long long g_phase_cycles[16];      // [0..10] cycles per phase, [11..14] event counters, [15] kernel cycles
__device__ unsigned long long g_warp_cycles[1 << 16];      // per warp: kernel duration of the last launch
__device__ unsigned long long g_warp_span[1 << 16][2];    // per warp: %globaltimer (ns) at start / end of the last launch
DEV unsigned long long global_ns() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ unsigned long long g_warp_phase[1 << 16][4];   // per warp: cycles in entities / sweep / closest / post of the last launch
__device__ unsigned long long g_solo_log[8192][12];        // per env served by a solo warp: env | t_first << 24 | converted << 32 | initial << 33 | resume substep << 34, ns at pickup, at end, at publication (b), cooperative iterations, %smid, cycles in entities / sweep / closest-point loop / post_collision
__device__ unsigned int g_solo_n;
__device__ unsigned long long g_hand_pub[1 << 12];        // per entry of (b): ns at publication
__device__ unsigned long long g_kind_cycles[2][32][2];   // roaming-entity phases: [move | think][kind][cycles, calls] (lane 0 of every warp)
__device__ unsigned long long g_coop_cycles[8];        // cooperative depenetration loop: [0] evaluate, [1] select, [2] push, [3] iterations, [4] services, [5] park/take/back
#define COOP_T(acc, k, t) { long long _n = clock64(); (acc)[k] += _n - (t); (t) = _n; }
#define PH_COUNT(k) atomicAdd(&g_phase_cycles[k], 1ULL)
#define PH_T0 long long _q0 = clock64();
#define PH_ACC(k) { long long _q1 = clock64(); if ((threadIdx.x & 31) == 0) atomicAdd(&g_phase_cycles[k], (unsigned long long)(_q1 - _q0)); _q0 = _q1; }
#define PH_BEGIN long long _ph_t = clock64();
#define PH_MARK(k) { long long _t2 = clock64(); if ((threadIdx.x & 31) == (__ffs(__activemask()) - 1)) { atomicAdd(&g_phase_cycles[k], (unsigned long long)(_t2 - _ph_t)); int _w = wp.group; int _b = (k) <= 3 ? 0 : ((k) == 5 ? 1 : ((k) == 6 ? 2 : (((k) == 9 || (k) == 10) ? 3 : -1))); if (_b >= 0 && _w >= 0 && _w < (1 << 16)) g_warp_phase[_w][_b] += (unsigned long long)(_t2 - _ph_t); if (_b >= 0 && _w < 0) wp.sph[_b] += (unsigned long long)(_t2 - _ph_t); } _ph_t = _t2; }
#else
#define PH_BEGIN
#define PH_MARK(k)
#define PH_COUNT(k)
#define PH_T0
#define COOP_T(acc, k, t)
#define PH_ACC(k)
#endif

// NPP_PHASE_SYNC: the warps of a block walk the frame's phases together (block-wide votes and barriers at
// the phase boundaries), so that they fetch the same code at the same time -- the kernel is bound by
// instruction fetch (sm__icc hit rate 62 %, profiles/), not by issue slots or memory.
#ifndef NPP_PHASE_SYNC
#define NPP_PHASE_SYNC 1
#endif
// (a solo warp -- Ctx::solo, block-uniform -- runs handed-off envs by itself and takes no part in block barriers)
#if NPP_PHASE_SYNC
// (NPP_SOLO is the including variant's compile-time switch, sim_variant.inc: only a kernel with solo warps tests the flag)
#define NPP_BLOCK_ANY(p) ((NPP_SOLO && c.solo) ? (__any_sync(0xffffffffu, (p)) != 0) : (__syncthreads_or((int)(p)) != 0))
#define NPP_BLOCK_SYNC() do { if (!(NPP_SOLO && c.solo)) __syncthreads(); } while (0)
#else
#define NPP_BLOCK_ANY(p) __any_sync(FULL_MASK, (p))
#define NPP_BLOCK_SYNC()
#endif

#define POOL_CAP 256                 // 32 lanes x at most 8 segments per cell per round
#define FULL_MASK 0xffffffffu
#define POOL_LCAP 24                 // larger gathers fall back to the out-of-line exact walk

struct WarpPool {
    double par[5][32];               // per-owner query parameters
    double res[POOL_CAP];            // per-item result (time of impact | adjusted distance^2 | wall term)
    double ra[POOL_CAP], rb[POOL_CAP];   // per-item closest point
    uint32_t seg[POOL_CAP];
    uint8_t own[POOL_CAP], aux[POOL_CAP];
    uint32_t lseg[32][POOL_LCAP];   // per-lane copy of the segments gathered for the current substep (ninja.py:282-285)
    int coop;                       // cooperative depenetration iterations this launch: the warp's cost for longest-first scheduling
    int group;                      // the env group this warp works on (profile build: per-warp phase attribution)
#ifdef NPP_PROFILE
    unsigned long long sph[6];      // solo warps: cycles per phase group (entities / sweep / closest-point loop / post_collision)
#endif
};

// ---- gather: the segments of a query box, per lane, in SpatialSegmentIndex.query_region order -----------
// (utils/spatial_segment_index.py:112-158: cells x-major / y-inner, a cell is skipped when absent or when
// its AABB misses the box, segments in cell order).  Each lane fills its own list wp.lseg[lane][0..n); no
// warp-level step is needed here, and the loads of two cells are issued before anything waits on them (a box
// is at most 2 x 2 cells unless the ninja moves more than a cell per frame).  Returns n, or -1 when the list
// does not fit (callers then run the literal per-thread version).
DEV bool cell_pass(uint32_t hdr, double min_x, double min_y, double max_x, double max_y) {
    if (!(hdr & 15)) return false;
    const double bx0 = u12((hdr >> 4) & 127), by0 = u12((hdr >> 11) & 63), bx1 = u12((hdr >> 17) & 127), by1 = u12((hdr >> 24) & 63);
    return !(max_x < bx0 || min_x > bx1 || max_y < by0 || min_y > by1);                    // _bounds_intersect :175-188
}
DEV int list_append(uint32_t* dst, int n, const uint4& a, const uint4& b, int cnt) {
    if (n < 0 || n + cnt > POOL_LCAP) return -1;
    dst[n] = a.x;
    if (cnt > 1) dst[n + 1] = a.y;
    if (cnt > 2) dst[n + 2] = a.z;
    if (cnt > 3) dst[n + 3] = a.w;
    if (cnt > 4) dst[n + 4] = b.x;
    if (cnt > 5) dst[n + 5] = b.y;
    if (cnt > 6) dst[n + 6] = b.z;
    if (cnt > 7) dst[n + 7] = b.w;
    return n + cnt;
}
DEV int gather_list(WarpPool& wp, const Ctx& c, bool active, double x1, double y1, double x2, double y2) {
    int n = 0;
    PH_T0
    if (active) {
        const double min_x = (x2 < x1) ? x2 : x1, min_y = (y2 < y1) ? y2 : y1;
        const double max_x = (x2 > x1) ? x2 : x1, max_y = (y2 > y1) ? y2 : y1;
        const int cx0 = clampi(floordiv24(min_x), 0, 43), cx1 = clampi(floordiv24(max_x), 0, 43);
        const int cy0 = clampi(floordiv24(min_y), 0, 24), cy1 = clampi(floordiv24(max_y), 0, 24);
        uint32_t* dst = wp.lseg[threadIdx.x & 31];
        const uint4 z = make_uint4(0, 0, 0, 0);
        #pragma unroll 1
        for (int xc = cx0; xc <= cx1; xc++) {
            #pragma unroll 1
            for (int yc = cy0; yc <= cy1; yc += 2) {
                const int ci = xc * NPP_NCY + yc;
                const uint32_t h0 = NPP_LDG(c.cell_hdr + ci), h1 = (yc < cy1) ? NPP_LDG(c.cell_hdr + ci + 1) : 0u;
                const bool p0 = cell_pass(h0, min_x, min_y, max_x, max_y), p1 = cell_pass(h1, min_x, min_y, max_x, max_y);
                const uint4* sp = reinterpret_cast<const uint4*>(c.segs + ci * NPP_SEGS_PER_CELL);
                uint4 a0 = z, b0 = z, a1 = z, b1 = z;
                if (p0) { a0 = NPP_LDG(sp); if ((h0 & 15) > 4) b0 = NPP_LDG(sp + 1); }
                if (p1) { a1 = NPP_LDG(sp + 2); if ((h1 & 15) > 4) b1 = NPP_LDG(sp + 3); }
                if (p0) n = list_append(dst, n, a0, b0, h0 & 15);
                if (p1) n = list_append(dst, n, a1, b1, h1 & 15);
            }
        }
    }
    __syncwarp();
    PH_ACC(11)
    return n;
}

// ---- evaluate + fold: the lists of all 32 lanes as one flat queue ---------------------------------------------
// All 32 lanes call this together; lane l contributes its first n entries of wp.lseg[l] (n <= 0: none).
// item(j, seg, owner) fills wp.res/ra/rb/aux[j] for queue slot j; fold(off, cnt) lets an owner consume its
// slots [off, off+cnt) in list order and returns false to stop early.  The queue holds POOL_CAP slots, so
// more than that (never observed: it needs 8+ segments on average around every ninja of the warp) takes
// several passes; an owner's fold state lives in its caller's variables and carries over.
template <class ItemF, class FoldF>
DEV void pooled_eval(WarpPool& wp, int n, ItemF&& item, FoldF&& fold) {
    const int lane = threadIdx.x & 31;
    const int cnt = n > 0 ? n : 0;
    PH_T0
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { int v = __shfl_up_sync(FULL_MASK, incl, d); if (lane >= d) incl += v; }
    const int off = incl - cnt;
    const int total = __shfl_sync(FULL_MASK, incl, 31);
    bool go = true;
    #pragma unroll 1
    for (int base = 0; base < total; base += POOL_CAP) {
        const int lo = max(off, base), hi = min(off + cnt, base + POOL_CAP);
        #pragma unroll 1
        for (int j = lo; j < hi; j++) { wp.seg[j - base] = wp.lseg[lane][j - off]; wp.own[j - base] = (uint8_t)lane; }
        __syncwarp();
        PH_ACC(12)
        const int m = min(total - base, POOL_CAP);
        #pragma unroll 1
        for (int j = lane; j < m; j += 32) item(j, wp.seg[j], (int)wp.own[j]);
        __syncwarp();
        PH_ACC(13)
        if (go && hi > lo) go = fold(lo - base, hi - lo);
        __syncwarp();
        PH_ACC(14)
    }
}

// ---- time of impact, guarded ------------------------------------------------------------------------
// literal physics.py:247-261 (out of line: reached only when the guarded version cannot decide)
__device__ __noinline__ double toi_circle_exact(double dx, double dy, double dot_prod, double vx, double vy, double rr) {
    double vel_sq = SQ(vx) + SQ(vy);
    double dist_sq = SQ(dx) + SQ(dy);
    if (dist_sq - rr > 0) {
        double radicand = SQ(dot_prod) - vel_sq * (dist_sq - rr);
        if (vel_sq > 0.0001 && dot_prod < 0 && radicand >= 0) return DIV(-dot_prod - SQRT(radicand), vel_sq);
        return 1;
    }
    return 0;
}
// Same value as the reference's get_time_of_intersection_circle_vs_circle whenever that value is < 1
// or == 0; otherwise some value >= 1 (callers only ever test `== 0` and `< shortest` with
// shortest <= 1: physics.py:119-126, entities.py:96).  vel_a = vx*vx + vy*vy (cheap squares).
DEV double toi_circle_guarded(double xpos, double ypos, double vx, double vy, double vel_a, double a, double b, double rr) {
    const double dx = xpos - a, dy = ypos - b;
    const double dot_prod = dx * vx + dy * vy;
    const double m = dx * dx + dy * dy;                 // within 3 ulp of dx**2 + dy**2
    const double diff = m - rr;
    const double g = 1e-13 * (m + rr);
    if (diff > g) {                                       // certainly outside the circle
        if (!(dot_prod < 0)) return 1;
        if (vel_a < 0.0001 * (1 - 1e-12)) return 1;       // vel_sq > 0.0001 certainly false
        if (vel_a > 0.0001 * (1 + 1e-12)) {
            const double d2 = dot_prod * dot_prod;
            const double R = d2 - vel_a * diff;
            const double E = 1e-13 * (d2 + vel_a * (m + rr));
            if (R < -E) return 1;                         // radicand certainly negative
            if (R > E) {
                const double t_lo = DIV(-dot_prod - SQRT(R + E), vel_a);   // below the true root
                if (t_lo > 1.000000001) return 1;         // certainly later than this frame
            }
        }
    } else if (diff < -g) {
        return 0;                                         // certainly inside
    }
    return toi_circle_exact(dx, dy, dot_prod, vx, vy, rr);
}
// physics.py:264-285 for a tile segment.  Its direction (wx, wy) is one of (+-12, 0), (0, +-12), (+-24, +-24),
// (+-24, +-12), (+-12, +-24), so seg_len = sqrt(wx**2 + wy**2) and the unit vector wx / seg_len are table
// values: seg_len in {12, sqrt(720), sqrt(1152)} and RN(wx / seg_len) = (wx / 12) * RN(12 / seg_len), the
// factor wx / 12 in {0, +-1, +-2} being an exact scaling.  No sqrt, no division.
DEV double toi_lineseg_int(uint32_t s, double xpos, double ypos, double dx, double dy, double radius) {
    const int x1 = s & 127, y1 = (s >> 7) & 63, wxi = (int)((s >> 13) & 127) - x1, wyi = (int)((s >> 20) & 63) - y1;
    const int l2 = wxi * wxi + wyi * wyi;                                        // 1, 5 or 8 (units of 144)
    const double seg_len = l2 == 1 ? 12.0 : (l2 == 5 ? 26.832815729997478 : 33.941125496954278);   // sqrt(144), sqrt(720), sqrt(1152)
    const double k = l2 == 1 ? 1.0 : (l2 == 5 ? 12.0 / 26.832815729997478 : 12.0 / 33.941125496954278);
    const double nx = s2d(wxi) * k, ny = s2d(wyi) * k;
    const double a1 = u12(x1), b1 = u12(y1);
    double normal_proj = (xpos - a1) * ny - (ypos - b1) * nx;
    double hor_proj = (xpos - a1) * nx + (ypos - b1) * ny;
    if (fabs(normal_proj) >= radius) {
        double dir = dx * ny - dy * nx;
        if (dir * normal_proj < 0) {
            double q = DIV(fabs(normal_proj) - radius, fabs(dir));
            double t = (1 < q) ? 1 : q;
            double hor_proj2 = hor_proj + t * (dx * nx + dy * ny);
            if (0 <= hor_proj2 && hor_proj2 <= seg_len) return t;
        }
    } else {
        if (0 <= hor_proj && hor_proj <= seg_len) return 0;
    }
    return 1;
}
// physics.py:288-314
__device__ __noinline__ double toi_arc_exact(double xpos, double ypos, double vx, double vy, double a, double b, double hor, double ver, double radius_circle) {
    double dx = xpos - a, dy = ypos - b;
    double dist_sq = SQ(dx) + SQ(dy);
    double vel_sq = SQ(vx) + SQ(vy);
    double dot_prod = dx * vx + dy * vy;
    double radius1 = 24.0 + radius_circle, radius2 = 24.0 - radius_circle;
    double rr1 = radius1 * radius1, rr2 = radius2 * radius2;             // small integers or x.5 halves: exact
    double t = 1;
    if (dist_sq > rr1) {
        double radicand = SQ(dot_prod) - vel_sq * (dist_sq - rr1);
        if (vel_sq > 0.0001 && dot_prod < 0 && radicand >= 0) t = DIV(-dot_prod - SQRT(radicand), vel_sq);
    } else if (dist_sq < rr2) {
        double radicand = SQ(dot_prod) - vel_sq * (dist_sq - rr2);
        if (vel_sq > 0.0001) { double q = DIV(-dot_prod + SQRT(radicand), vel_sq); t = (1 < q) ? 1 : q; }
    } else {
        t = 0;
    }
    if ((dx + t * vx) * hor > 0 && (dy + t * vy) * ver > 0) return t;
    return 1;
}
// entities.py:82-96,180-203 for the ninja's sweep radius (an exactly representable square)
DEV double seg_toi_guarded(uint32_t s, double xpos, double ypos, double dx, double dy, double vel_a, double radius, double rr) {
    if (!(s & NPP_SEG_ARC)) {
        SegLin l = dec_lin(s);
        double t1 = toi_circle_guarded(xpos, ypos, dx, dy, vel_a, l.x1, l.y1, rr);
        double t2 = toi_circle_guarded(xpos, ypos, dx, dy, vel_a, l.x2, l.y2, rr);
        double t3 = toi_lineseg_int(s, xpos, ypos, dx, dy, radius);
        return min3(t1, t2, t3);
    } else {
        SegArc a = dec_arc(s);
        double t1 = toi_circle_guarded(xpos, ypos, dx, dy, vel_a, a.cx + 24.0 * a.hor, a.cy, rr);
        double t2 = toi_circle_guarded(xpos, ypos, dx, dy, vel_a, a.cx, a.cy + 24.0 * a.ver, rr);
        double t3 = toi_arc_exact(xpos, ypos, dx, dy, a.cx, a.cy, a.hor, a.ver, radius);
        return min3(t1, t2, t3);
    }
}

// ---- pooled sweep_circle_vs_tiles (physics.py:104-128), ninja radius 5 -----------------------------------
// The reference returns 0 as soon as a segment reports exactly 0, else the minimum; the owner's fold
// walks its items in list order with the same early exit.
// literal per-thread sweep (out of line: only for a box whose segment list does not fit the per-lane cache)
__device__ __noinline__ double sweep_exact(const uint32_t* cell_hdr, const uint32_t* segs, double xpos_old, double ypos_old, double dx, double dy, double radius) {
    Ctx c; c.cell_hdr = cell_hdr; c.segs = segs;
    return sweep_circle_vs_tiles(c, xpos_old, ypos_old, dx, dy, radius);
}
DEV double pooled_sweep(WarpPool& wp, const Ctx& c, bool active, double xpos_old, double ypos_old, double dx, double dy) {
    const int lane = threadIdx.x & 31;
    const double radius = NINJA_RADIUS * 0.5;
    double xpos_new = xpos_old + dx, ypos_new = ypos_old + dy;
    double width = radius + 1;
    double x1 = (xpos_old < xpos_new ? xpos_old : xpos_new) - width;
    double y1 = (ypos_old < ypos_new ? ypos_old : ypos_new) - width;
    double x2 = (xpos_old > xpos_new ? xpos_old : xpos_new) + width;
    double y2 = (ypos_old > ypos_new ? ypos_old : ypos_new) + width;
    wp.par[0][lane] = xpos_old; wp.par[1][lane] = ypos_old; wp.par[2][lane] = dx; wp.par[3][lane] = dy; wp.par[4][lane] = dx * dx + dy * dy;
    const int n = gather_list(wp, c, active, x1, y1, x2, y2);
    __syncwarp();
    double shortest = 1;
    pooled_eval(wp, n,
        [&](int j, uint32_t s, int o) {
            wp.res[j] = seg_toi_guarded(s, wp.par[0][o], wp.par[1][o], wp.par[2][o], wp.par[3][o], wp.par[4][o], radius, radius * radius);
        },
        [&](int off, int cnt) {
            #pragma unroll 1
            for (int k = 0; k < cnt; k++) {
                double t = wp.res[off + k];
                if (t == 0) { shortest = 0; return false; }
                if (t < shortest) shortest = t;
            }
            return true;
        });
    if (n < 0) shortest = sweep_exact(c.cell_hdr, c.segs, xpos_old, ypos_old, dx, dy, radius);
    return shortest;
}

// literal per-thread get_single_closest_point with exact squares (out of line, rare)
__device__ __noinline__ int closest_point_exact(const uint32_t* cell_hdr, const uint32_t* segs, double gx1, double gy1, double gx2, double gy2,
                                                double xpos, double ypos, double& pa, double& pb) {
    Ctx c; c.cell_hdr = cell_hdr; c.segs = segs;
    return get_single_closest_point(c, gx1, gy1, gx2, gy2, xpos, ypos, NINJA_RADIUS, pa, pb);
}

// ---- pooled get_single_closest_point (physics.py:131-180), radius 10 -----------------------------------------
// Items carry the cheap adjusted squared distance; the owner's strict-< scan accepts a comparison only
// when it cannot flip under 1-ulp changes of the squares (or both items are the same point and side);
// otherwise it re-runs the literal per-thread version with exact SQ().
// One candidate of get_single_closest_point's strict-< scan (physics.py:155-179).  d2 was computed
// with cheap squares; `ix` says it may differ from the reference's value by an ulp.  Between two exact
// values the reference's own comparison is used; otherwise the comparison is accepted only outside a guard band.
DEV void closest_fold(double d2, double a, double b, int r, bool ix, double& best, double& ba, double& bb, int& result, bool& bix, bool& uncertain) {
    if (!ix && !bix) { if (d2 < best) { best = d2; ba = a; bb = b; result = r; } return; }
    const double G = 1e-10;                                  // >> 3 ulp of any d2 here (<= 1000)
    if (d2 < best - G) { best = d2; ba = a; bb = b; result = r; bix = ix; }
    else if (d2 > best + G) { }
    else if (a == ba && b == bb && r == result) { }          // identical inputs: exact values tie, first wins
    else uncertain = true;
}
// A tile segment decoded for repeated closest-point queries (the depenetration loop re-tests the same
// segments up to 32 times): AABB as get_bounds() (entities.py:36-41,121-125) + the closest-point constants.
struct SegC {
    double bx0, by0, bx1, by1;
    double x1, y1, px, py, lensq;        // linear: first end, direction, squared length | arc: centre, (hor, ver), convex ? 1 : 0
    bool arc;
};
DEV SegC seg_decode(uint32_t s) {
    SegC g;
    g.arc = (s & NPP_SEG_ARC) != 0;
    const int ax = s & 127, ay = (s >> 7) & 63;
    int bx, by;
    if (g.arc) {
        const int hn = (s >> 13) & 1, vn = (s >> 14) & 1;
        bx = ax + (hn ? -2 : 2); by = ay + (vn ? -2 : 2);
        g.px = hn ? -1.0 : 1.0; g.py = vn ? -1.0 : 1.0; g.lensq = ((s >> 15) & 1) ? 1.0 : 0.0;
    } else {
        bx = (s >> 13) & 127; by = (s >> 20) & 63;
        const int wx = bx - ax, wy = by - ay;
        g.px = s2d(12 * wx); g.py = s2d(12 * wy); g.lensq = s2d(144 * (wx * wx + wy * wy));
    }
    g.x1 = u12(ax); g.y1 = u12(ay);
    g.bx0 = u12(min(ax, bx)); g.bx1 = u12(max(ax, bx)); g.by0 = u12(min(ay, by)); g.by1 = u12(max(ay, by));
    return g;
}
// closest point + adjusted squared distance (physics.py:155-179); false when the AABB test rejects the segment
DEV bool closest_eval(const SegC& g, double x, double y, double& d2, double& a, double& b, int& r, bool& ix) {
    if (g.bx1 < x - NINJA_RADIUS || g.bx0 > x + NINJA_RADIUS || g.by1 < y - NINJA_RADIUS || g.by0 > y + NINJA_RADIUS) return false;
    const double dx = x - g.x1, dy = y - g.y1;
    bool back;
    if (!g.arc) {                                                            // entities.py:43-59
        double u = div_seglen(dx * g.px + dy * g.py, g.lensq);
        if (0 > u) u = 0;
        if (1 < u) u = 1;
        a = g.x1 + u * g.px; b = g.y1 + u * g.py;
        back = (dy * g.px - dx * g.py < 0);
    } else {                                                                 // entities.py:127-157
        back = false;
        if (dx * g.px > 0 && dy * g.py > 0) {
            double dist = SQRT(SQ(dx) + SQ(dy));
            double qa, qb;
            div2(24.0 * dx, 24.0 * dy, dist, qa, qb);
            a = g.x1 + qa; b = g.y1 + qb;
            back = (g.lensq != 0) ? (dist < 24.0) : (dist > 24.0);
        } else if (dx * g.px > dy * g.py) { a = g.x1 + 24.0 * g.px; b = g.y1; }
        else { a = g.x1; b = g.y1 + 24.0 * g.py; }
    }
    ix = false;
    d2 = sq_flag(x - a, ix) + sq_flag(y - b, ix);
    if (!back) d2 -= 0.1;
    r = back ? -1 : 1;
    return true;
}
DEV bool closest_item(uint32_t s, double x, double y, double& d2, double& a, double& b, int& r, bool& ix) {
    return closest_eval(seg_decode(s), x, y, d2, a, b, r, ix);
}

// One closest-point query per participating lane, pooled over the warp, on the lists in wp.lseg (n entries each,
// n = -1: list overflow, the lane runs the literal version).  (gx..) is the box the list was gathered from.
DEV int pooled_list_closest(WarpPool& wp, const Ctx& c, bool active, int n, double gx1, double gy1, double gx2, double gy2,
                            double xpos, double ypos, double& pa, double& pb) {
    const int lane = threadIdx.x & 31;
    wp.par[0][lane] = xpos; wp.par[1][lane] = ypos;
    __syncwarp();
    double best = __longlong_as_double(0x7ff0000000000000LL), ba = 0, bb = 0;
    int result = 0;
    bool uncertain = false, bix = false;
    pooled_eval(wp, active ? n : 0,
        [&](int j, uint32_t s, int o) {
            double d2, a, b; int r; bool ix;
            if (!closest_item(s, wp.par[0][o], wp.par[1][o], d2, a, b, r, ix)) { wp.aux[j] = 2; return; }
            wp.res[j] = d2; wp.ra[j] = a; wp.rb[j] = b; wp.aux[j] = (uint8_t)((r < 0 ? 1 : 0) | (ix ? 4 : 0));
        },
        [&](int off, int cnt) {
            #pragma unroll 1
            for (int k = 0; k < cnt; k++) {
                int ax = wp.aux[off + k];
                if (ax == 2) continue;                                   // AABB-rejected
                closest_fold(wp.res[off + k], wp.ra[off + k], wp.rb[off + k], (ax & 1) ? -1 : 1, (ax & 4) != 0, best, ba, bb, result, bix, uncertain);
            }
            return true;
        });
    if (active && (uncertain || n < 0)) return closest_point_exact(c.cell_hdr, c.segs, gx1, gy1, gx2, gy2, xpos, ypos, pa, pb);
    pa = ba; pb = bb;
    return result;
}
// First iteration of the depenetration loop: gathers each lane's segment list around the post-sweep
// position (ninja.py:282-285; it stays in wp.lseg[lane][0..nseg) for the later iterations) and queries it.
DEV int pooled_closest_point(WarpPool& wp, const Ctx& c, bool active, double gx1, double gy1, double gx2, double gy2,
                             double xpos, double ypos, double& pa, double& pb, int& nseg) {
    nseg = gather_list(wp, c, active, gx1, gy1, gx2, gy2);
    return pooled_list_closest(wp, c, active, nseg, gx1, gy1, gx2, gy2, xpos, ypos, pa, pb);
}
// Iterations 2..31 of one owner lane `o`, served by the whole warp: lane k evaluates the owner's k-th
// cached segment (decoded once per owner by the caller: `g`).  With exact squares everywhere the reference's strict-< scan picks the first minimum,
// which is a warp min + ballot; when some square is flagged inexact the owner folds the items in list
// order with the guard band (closest_fold) as in the per-lane version.  All 32 lanes call this; only
// lane o's (result, pa, pb) are meaningful.
// min over the warp of non-NaN doubles: order-preserving 64-bit keys, reduced 32 bits at a time with redux.sync
DEV double warp_min_f64(double v) {
    long long b = __double_as_longlong(v);
    unsigned long long key = (unsigned long long)(b ^ ((b >> 63) | (long long)0x8000000000000000ULL));   // monotone in v
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mhi = __reduce_min_sync(FULL_MASK, hi);
    const unsigned mlo = __reduce_min_sync(FULL_MASK, hi == mhi ? lo : 0xffffffffu);
    key = ((unsigned long long)mhi << 32) | mlo;
    b = (long long)key;
    b = (b < 0) ? (b ^ (long long)0x8000000000000000ULL) : ~b;
    return __longlong_as_double(b);
}
DEV int warp_closest_point(WarpPool& wp, const Ctx& c, int o, int ns, const SegC& g, double gx1, double gy1, double gx2, double gy2,
                           double xpos, double ypos, double& pa, double& pb) {
    const int lane = threadIdx.x & 31;
    const double x = __shfl_sync(FULL_MASK, xpos, o), y = __shfl_sync(FULL_MASK, ypos, o);
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    double d2 = inf, a = 0, b = 0; int r = 0; bool ix = false, ok = false;
    if (lane < ns) ok = closest_eval(g, x, y, d2, a, b, r, ix);
    if (!ok) d2 = inf;
    int result = 0;
    if (__ballot_sync(FULL_MASK, ok && ix) == 0) {
        const double mn = warp_min_f64(d2);
        const unsigned eq = __ballot_sync(FULL_MASK, ok && d2 == mn);
        if (eq) {
            const int w = __ffs(eq) - 1;
            pa = __shfl_sync(FULL_MASK, a, w); pb = __shfl_sync(FULL_MASK, b, w); result = __shfl_sync(FULL_MASK, r, w);
        }
        return result;
    }
    wp.res[lane] = d2; wp.ra[lane] = a; wp.rb[lane] = b; wp.aux[lane] = (uint8_t)(!ok ? 2 : ((r < 0 ? 1 : 0) | (ix ? 4 : 0)));
    __syncwarp();
    if (lane == o) {
        double best = inf, ba = 0, bb = 0;
        bool uncertain = false, bix = false;
        #pragma unroll 1
        for (int k = 0; k < ns; k++) {
            int ax = wp.aux[k];
            if (ax == 2) continue;
            closest_fold(wp.res[k], wp.ra[k], wp.rb[k], (ax & 1) ? -1 : 1, (ax & 4) != 0, best, ba, bb, result, bix, uncertain);
        }
        if (uncertain) { result = closest_point_exact(c.cell_hdr, c.segs, gx1, gy1, gx2, gy2, xpos, ypos, pa, pb); }
        else { pa = ba; pb = bb; }
    }
    __syncwarp();
    return result;
}

// closest_eval without the early AABB return: the bounds test only gates `ok` and stays off the dependency chain
// (in the cooperative loop every lane owns one segment, so skipping the arithmetic would save nothing).
DEV bool closest_eval_nb(const SegC& g, double x, double y, double& d2, double& a, double& b, int& r) {
    const bool inb = !(g.bx1 < x - NINJA_RADIUS || g.bx0 > x + NINJA_RADIUS || g.by1 < y - NINJA_RADIUS || g.by0 > y + NINJA_RADIUS);
    const double dx = x - g.x1, dy = y - g.y1;
    bool back;
    if (!g.arc) {                                                            // entities.py:43-59
        double u = div_seglen(dx * g.px + dy * g.py, g.lensq);
        if (0 > u) u = 0;
        if (1 < u) u = 1;
        a = g.x1 + u * g.px; b = g.y1 + u * g.py;
        back = (dy * g.px - dx * g.py < 0);
    } else {                                                                 // entities.py:127-157
        back = false;
        if (dx * g.px > 0 && dy * g.py > 0) {
            double dist = SQRT(SQ(dx) + SQ(dy));
            double qa, qb;
            div2(24.0 * dx, 24.0 * dy, dist, qa, qb);
            a = g.x1 + qa; b = g.y1 + qb;
            back = (g.lensq != 0) ? (dist < 24.0) : (dist > 24.0);
        } else if (dx * g.px > dy * g.py) { a = g.x1 + 24.0 * g.px; b = g.y1; }
        else { a = g.x1; b = g.y1 + 24.0 * g.py; }
    }
    // exact squares (pow2_exact: one FMA test, the table path for the 3 % near a rounding midpoint): no "inexact" flag to
    // resolve afterwards -- with one segment per lane a flagged lane costs less here than the owner's ordered re-fold
    d2 = SQ(x - a) + SQ(y - b);
    if (!back) d2 -= 0.1;
    r = back ? -1 : 1;
    return inb;
}
// First minimum of d2 over the lanes with ok set: ballot of the winner(s) -- 0 when no lane is ok, else the lowest
// set bit is the reference's strict-< scan result (physics.py:170-176).  Usually one redux on the high words decides.
DEV unsigned warp_first_min(double d2, bool ok) {
    const long long bits = __double_as_longlong(d2);
    const unsigned long long key = (unsigned long long)(bits ^ ((bits >> 63) | (long long)0x8000000000000000ULL));   // monotone in d2
    const unsigned hi = ok ? (unsigned)(key >> 32) : 0xffffffffu, lo = (unsigned)key;
    const unsigned mhi = __reduce_min_sync(FULL_MASK, hi);
    unsigned cand = __ballot_sync(FULL_MASK, ok && hi == mhi);
    if (cand & (cand - 1)) {
        const unsigned mlo = __reduce_min_sync(FULL_MASK, (ok && hi == mhi) ? lo : 0xffffffffu);
        cand = __ballot_sync(FULL_MASK, ok && hi == mhi && lo == mlo);
    }
    return cand;
}

// Ordered fold of the 32 lanes' evaluated segments by the owner lane (a square was flagged inexact: the guard-banded
// strict-< scan of closest_fold, the literal walk when that is uncertain too), result broadcast to every lane.
DEV int coop_ordered_fold(WarpPool& wp, const Ctx& c, int o, int ns, bool ok, double d2, double a, double b, int r, bool ix,
                          double gx1, double gy1, double gx2, double gy2, double x, double y, double& pa, double& pb) {
    const int lane = threadIdx.x & 31;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    int result = 0;
    wp.res[lane] = d2; wp.ra[lane] = a; wp.rb[lane] = b; wp.aux[lane] = (uint8_t)(!ok ? 2 : ((r < 0 ? 1 : 0) | (ix ? 4 : 0)));
    __syncwarp();
    double fa = 0, fb = 0;
    if (lane == o) {
        double best = inf, ba = 0, bb = 0;
        bool uncertain = false, bix = false;
        #pragma unroll 1
        for (int k = 0; k < ns; k++) {
            int ax = wp.aux[k];
            if (ax == 2) continue;
            closest_fold(wp.res[k], wp.ra[k], wp.rb[k], (ax & 1) ? -1 : 1, (ax & 4) != 0, best, ba, bb, result, bix, uncertain);
        }
        if (uncertain) { double ea, eb; result = closest_point_exact(c.cell_hdr, c.segs, gx1, gy1, gx2, gy2, x, y, ea, eb); fa = ea; fb = eb; }
        else { fa = ba; fb = bb; }
    }
    __syncwarp();
    pa = __shfl_sync(FULL_MASK, fa, o); pb = __shfl_sync(FULL_MASK, fb, o);
    return __shfl_sync(FULL_MASK, result, o);
}
// Selection half of the uniform query: from every lane's evaluated segment (ok, d2, a, b, r, ix) to the reference's strict-<
// scan result on every lane; `win` = the winning lane (-1: none, or decided by the ordered fold).
DEV int warp_select_uniform(WarpPool& wp, const Ctx& c, int o, int ns, bool ok, double d2, double a, double b, int r, bool ix,
                            double gx1, double gy1, double gx2, double gy2, double x, double y, double& pa, double& pb, int& win) {
    win = -1;
    int result = 0;
    if (__ballot_sync(FULL_MASK, ok && ix) == 0) {
        const long long bits = __double_as_longlong(d2);
        const unsigned long long key = (unsigned long long)(bits ^ ((bits >> 63) | (long long)0x8000000000000000ULL));   // monotone in d2
        const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
        const unsigned mhi = __reduce_min_sync(FULL_MASK, hi);
        unsigned cand = __ballot_sync(FULL_MASK, ok && hi == mhi);
        if (cand & (cand - 1)) {                                             // several lanes share the high word: decide on the low one
            const unsigned mlo = __reduce_min_sync(FULL_MASK, (ok && hi == mhi) ? lo : 0xffffffffu);
            cand = __ballot_sync(FULL_MASK, ok && hi == mhi && lo == mlo);
        }
        if (cand) {                                                          // first minimum in list order (strict < scan, physics.py:170-176)
            const int w = __ffs(cand) - 1;
            win = w;
            pa = __shfl_sync(FULL_MASK, a, w); pb = __shfl_sync(FULL_MASK, b, w); result = __shfl_sync(FULL_MASK, r, w);
        }
        return result;
    }
    return coop_ordered_fold(wp, c, o, ns, ok, d2, a, b, r, ix, gx1, gy1, gx2, gy2, x, y, pa, pb);
}
// Uniform variant for the cooperative depenetration loop: (x, y) is the served owner's position, already held by
// every lane, and the result (a, b, result) is returned to every lane, so that all 32 lanes can run the push
// redundantly -- no divergent owner-only section, no broadcast of the position and of the continue flag per
// iteration.  The minimum is usually decided by the high words alone (one redux + one ballot).
DEV int warp_closest_uniform(WarpPool& wp, const Ctx& c, int o, int ns, const SegC& g, double gx1, double gy1, double gx2, double gy2,
                             double x, double y, double& pa, double& pb, int& win, long long* cacc, long long& ct) {
    const int lane = threadIdx.x & 31;
    win = -1;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    double d2 = inf, a = 0, b = 0; int r = 0; bool ix = false, ok = false;
    if (lane < ns) ok = closest_eval(g, x, y, d2, a, b, r, ix);
    if (!ok) d2 = inf;
    COOP_T(cacc, 0, ct)
    return warp_select_uniform(wp, c, o, ns, ok, d2, a, b, r, ix, gx1, gy1, gx2, gy2, x, y, pa, pb, win);
}
// stash slots of the cooperative loop: 13 values per lane in the idle tails of the pool's result arrays
DEV double& coop_slot(WarpPool& wp, int k, int l) { return k < 7 ? wp.res[32 * (k + 1) + l] : wp.ra[32 * (k - 6) + l]; }

// ---- pooled wall test of post_collision (ninja.py:422-441) ----------------------------------------------------
// Items evaluate the literal reference expression on cheap squares; the owner adds the terms to
// wall_normal in list order.  `uncertain` when a term's squares were flagged inexact (table path needed).
DEV void pooled_wall_test(WarpPool& wp, const Ctx& c, bool active, double xpos, double ypos, double& wall_normal, bool& uncertain) {
    const int lane = threadIdx.x & 31;
    const double rad = NINJA_RADIUS + 0.1;
    wp.par[0][lane] = xpos; wp.par[1][lane] = ypos;
    __syncwarp();
    uncertain = false;
    const int n = gather_list(wp, c, active, xpos - rad, ypos - rad, xpos + rad, ypos + rad);
    if (n < 0) uncertain = true;                                         // list overflow: the caller's literal version
    pooled_eval(wp, n,
        [&](int j, uint32_t s, int o) {
            const double x = wp.par[0][o], y = wp.par[1][o];
            bool back; double a, b;
            seg_closest_point(s, x, y, back, a, b);
            double dx = x - a, dy = y - b;
            int code = 0;                                                // 0 no term, 1 term in res[j], 3 needs the exact path
            if (fabs(dy) < 0.00001) {
                bool ix = false;
                double dist = SQRT(sq_flag(dx, ix) + sq_flag(dy, ix));
                if (ix) code = 3;
                else if (0 < dist && dist <= rad) { wp.res[j] = DIV(dx, dist); code = 1; }
            }
            wp.aux[j] = (uint8_t)code;
        },
        [&](int off, int cnt) {
            #pragma unroll 1
            for (int k = 0; k < cnt; k++) {
                int code = wp.aux[off + k];
                if (code == 1) wall_normal += wp.res[off + k]; else if (code == 3) uncertain = true;
            }
            return true;
        });
}
