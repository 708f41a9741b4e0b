// ninja.cuh -- the ninja's per-frame update.  Reference: /root/reference/nclone/ninja.py.
// Included once per kernel variant, after entities.cuh (sim_variant.inc).


DEV void ninja_integrate(Nin& n) {                                       // ninja.py:198-206
    n.f[F_XSPEED] *= n.f[F_DRAG];
    n.f[F_YSPEED] *= n.f[F_DRAG];
    n.f[F_YSPEED] += n.f[F_GRAVITY];
    n.f[F_XPOS_OLD] = n.f[F_XPOS]; n.f[F_YPOS_OLD] = n.f[F_YPOS];
    n.f[F_XPOS] += n.f[F_XSPEED]; n.f[F_YPOS] += n.f[F_YSPEED];
}

// ninja.py:208-222.  The reference caches the gathered entity list for the four substeps; the
// list cannot change in between (no grid_move / activation happens during ninja physics), so it is
// enough to remember where it was gathered and whether it holds anything physical at all.
DEV bool ninja_pre_collision(Nin& n, Ctx& c, double& gx, double& gy) {
    n.f[F_XSPEED_OLD] = n.f[F_XSPEED]; n.f[F_YSPEED_OLD] = n.f[F_YSPEED];
    n.i[I_FLOOR_COUNT] = 0; n.i[I_CEILING_COUNT] = 0;                   // wall_count stays 0 forever
    n.f[F_FLOOR_NX] = 0; n.f[F_FLOOR_NY] = 0; n.f[F_CEIL_NX] = 0; n.f[F_CEIL_NY] = 0;
    n.i[I_CRUSHABLE] = 0;
    n.f[F_XCRUSH] = 0; n.f[F_YCRUSH] = 0; n.f[F_CRUSH_LEN] = 0;
    gx = n.f[F_XPOS]; gy = n.f[F_YPOS];
    if (!HAS_ANY(NPP_PHYS_MASK)) return false;
    int nphys = 0;
    gather_entities(c, gx, gy, [&](int id) { if (is_physical_kind(ent_kind(c, id))) nphys++; });
    return nphys > 0;
}

DEV void ninja_collide_vs_objects(Nin& n, Ctx& c, double gx, double gy) {   // ninja.py:224-267
    gather_entities(c, gx, gy, [&](int id) {
        if (!is_physical_kind(ent_kind(c, id))) return;
        EntS e = ld_ent(c, id);
        double depen_x, depen_y, depen_len;
        if (!ent_physical_collision(n, c, id, e, depen_x, depen_y, depen_len)) return;
        double pop_x = depen_x * depen_len, pop_y = depen_y * depen_len;
        n.f[F_XPOS] += pop_x; n.f[F_YPOS] += pop_y;
        if (e.kind != NPP_K_BOUNCE) { n.f[F_XCRUSH] += pop_x; n.f[F_YCRUSH] += pop_y; n.f[F_CRUSH_LEN] += depen_len; }
        if (e.kind == NPP_K_THWUMP) n.i[I_CRUSHABLE] = 1;
        if (e.kind == NPP_K_BOUNCE || e.kind == NPP_K_THWUMP || e.kind == NPP_K_SHWUMP) { n.f[F_XSPEED] += pop_x; n.f[F_YSPEED] += pop_y; }
        if (e.kind == NPP_K_ONE_WAY) {
            double xspeed_new = (n.f[F_XSPEED] * depen_y - n.f[F_YSPEED] * depen_x) * depen_y;
            // axis-aligned normals are Python ints (physics.py:322-331): -(int 0) is 0, not -0.0
            double yspeed_new = (n.f[F_XSPEED] * depen_y - n.f[F_YSPEED] * depen_x) * (depen_x == 0 ? 0.0 : -depen_x);
            n.f[F_XSPEED] = xspeed_new; n.f[F_YSPEED] = yspeed_new;
        }
        if (depen_y >= -0.0001) { n.i[I_CEILING_COUNT] += 1; n.f[F_CEIL_NX] += depen_x; n.f[F_CEIL_NY] += depen_y; }
        else { n.i[I_FLOOR_COUNT] += 1; n.f[F_FLOOR_NX] += depen_x; n.f[F_FLOOR_NY] += depen_y; }
    });
}

// ninja.py:269-379, warp-synchronous: every lane of the warp calls this; `alive` lanes collide.
// Returns true when the substep changed nothing (position bit-identical, no depenetration push).
DEV bool ninja_collide_vs_tiles(Nin& n, Ctx& c, WarpPool& wp, bool alive) {
    const double xin = n.f[F_XPOS], yin = n.f[F_YPOS];
    bool pushed = false;
    double dx = n.f[F_XPOS] - n.f[F_XPOS_OLD], dy = n.f[F_YPOS] - n.f[F_YPOS_OLD];
    PH_BEGIN
    double time = pooled_sweep(wp, c, alive, n.f[F_XPOS_OLD], n.f[F_YPOS_OLD], dx, dy);
    PH_MARK(5)
    NPP_BLOCK_SYNC();
    double xpos = n.f[F_XPOS_OLD] + time * dx;       // recomputed even when time == 1 (not a no-op in fp)
    double ypos = n.f[F_YPOS_OLD] + time * dy;
    // segments are gathered once, around the post-sweep position (ninja.py:282-285)
    const double gx1 = xpos - NINJA_RADIUS, gy1 = ypos - NINJA_RADIUS, gx2 = xpos + NINJA_RADIUS, gy2 = ypos + NINJA_RADIUS;
    double xspeed = n.f[F_XSPEED], yspeed = n.f[F_YSPEED];
    // One depenetration push (ninja.py:303-379) from the closest point (a, b), in two parts: push_eval is a pure function of
    // (result, a, b) and the position -- the dependent chain of a wedged ninja -- and push_commit applies it.  The split lets the
    // cooperative loop evaluate the push from a PREDICTED winner while the warp minimum is still being found (below).
    struct Push { bool go; double dx, dy, inv_dist, norm_dx, norm_dy, depen_len; };
    auto push_eval = [&](int result, double a, double b) -> Push {
        Push t; t.go = false; t.dx = 0; t.dy = 0; t.inv_dist = 0; t.norm_dx = 0; t.norm_dy = 0; t.depen_len = 0;
        if (result == 0) return t;
        double ex = xpos - a; const double ey = ypos - b;
        if (fabs(ex) <= 0.0000001) {                  // ninja.py:312-318 band-aid
            ex = 0;
            if (xpos == 50.51197510492316 || xpos == 49.23232124849253) ex = -0x1p-47;
            if (xpos == 49.153536108584795) ex = 0x1p-47;
        }
        t.dx = ex; t.dy = ey;
        const double dist_sq = ex * ex + ey * ey;
        if (dist_sq < 1e-16) return t;
        const double dist = SQRT(dist_sq);
        t.depen_len = NINJA_RADIUS - (result < 0 ? -dist : dist);     // dist * result, result = +-1: exact either way
        if (t.depen_len < 0.0000001) return t;
        t.inv_dist = DIV(1.0, dist);
        t.norm_dx = ex * t.inv_dist; t.norm_dy = ey * t.inv_dist;
        t.go = true;
        return t;
    };
    auto push_commit = [&](const Push& t) {
        dx = t.dx; dy = t.dy;
        const double depen_x = t.norm_dx * t.depen_len, depen_y = t.norm_dy * t.depen_len;
        pushed = true;
        xpos += depen_x; ypos += depen_y;
        n.f[F_XCRUSH] += depen_x; n.f[F_YCRUSH] += depen_y; n.f[F_CRUSH_LEN] += t.depen_len;
        const double dot_product = xspeed * dx + yspeed * dy;
        if (dot_product < 0) {
            const double cross_product = xspeed * dy - yspeed * dx;
            const double inv_dist_sq = t.inv_dist * t.inv_dist;
            xspeed = cross_product * inv_dist_sq * dy;
            // dx == 0 only via the band-aid's `dx = 0` (a Python int): -(int 0) is 0, not -0.0
            yspeed = cross_product * inv_dist_sq * (dx == 0 ? 0.0 : -dx);
        }
        if (dy >= -0.0001) { n.i[I_CEILING_COUNT] += 1; n.f[F_CEIL_NX] += t.norm_dx; n.f[F_CEIL_NY] += t.norm_dy; }
        else { n.i[I_FLOOR_COUNT] += 1; n.f[F_FLOOR_NX] += t.norm_dx; n.f[F_FLOOR_NY] += t.norm_dy; }
    };
    auto depen = [&](int result, double a, double b) -> bool {
        const Push t = push_eval(result, a, b);
        if (t.go) push_commit(t);
        return t.go;
    };
    // Iteration 0 is pooled over the warp (and caches each lane's gathered list); iteration 1 runs per
    // lane on the cached list (5 % of substeps need it, i.e. most warps); iterations 2..31 are rare per
    // lane (0.2 %) but long (a wedged ninja takes all 32), so the warp serves one such lane at a time
    // with its cached segments spread over the lanes.
    double a0 = 0, b0 = 0; int nseg = 0;
    bool need;
    int it_first = 2;                                       // first iteration of the cooperative service below (warp-uniform)
#if NPP_SOLO
    // A solo warp runs ONE env (sim_variant.inc): the pooled machinery for iterations 0 and 1 -- queue scan, shared-memory
    // staging, ordered fold, built for 32 sparse queries -- cost it ~7 k cycles per substep for two evaluations.  It gathers
    // the env's list and enters the cooperative service at iteration 0 instead: the same query and the same push per
    // iteration, on the same inputs.
    if (c.solo) {
        nseg = gather_list(wp, c, alive, gx1, gy1, gx2, gy2);
        __syncwarp();
        need = alive;
        it_first = 0;
    } else
#endif
    {
    int result0 = pooled_closest_point(wp, c, alive, gx1, gy1, gx2, gy2, xpos, ypos, a0, b0, nseg);
    need = alive && depen(result0, a0, b0);
    NPP_BLOCK_SYNC();
    if (__any_sync(FULL_MASK, need)) {                      // iteration 1: 5 % of lanes, i.e. most warps -> pooled again
        double a = 0, b = 0;
        int result = pooled_list_closest(wp, c, need, nseg, gx1, gy1, gx2, gy2, xpos, ypos, a, b);
        if (need) need = depen(result, a, b);
    }
    }
    unsigned todo;
    // Cooperative service, uniform: every lane parks its own loop state in shared memory and takes over the served
    // owner's, so the whole iteration (query + push) runs redundantly on all 32 lanes without a divergent section;
    // a lone wedged ninja -- the kernel's critical path -- pays per iteration for one broadcast of (a, b, result) only.
#define COOP_STATE(X) X(0, xpos) X(1, ypos) X(2, xspeed) X(3, yspeed) X(4, n.f[F_XCRUSH]) X(5, n.f[F_YCRUSH]) X(6, n.f[F_CRUSH_LEN]) \
                      X(7, n.f[F_CEIL_NX]) X(8, n.f[F_CEIL_NY]) X(9, n.f[F_FLOOR_NX]) X(10, n.f[F_FLOOR_NY])
    int ncoop = 0;
    long long cacc[6] = {0, 0, 0, 0, 0, 0}, ct = 0;
    (void)cacc; (void)ct;
    #pragma unroll 1
    while ((todo = __ballot_sync(FULL_MASK, need)) != 0) {
#ifdef NPP_PROFILE
        ct = clock64(); cacc[4]++;
#endif
        const int lane = threadIdx.x & 31;
        const bool pushed_own = pushed;
#define COOP_PARK(k, v) coop_slot(wp, k, lane) = v;
#define COOP_TAKE(k, v) v = coop_slot(wp, k, o);
#define COOP_BACK(k, v) v = coop_slot(wp, k, lane);
#if NPP_GROUPS
        // Several lanes wait for service (the envs of a level share a warp, and a level that wedges one ninja wedges others):
        // serve up to four of them AT ONCE, each by a group of W = 8 (or two by W = 16) lanes, instead of one after the other --
        // the loop is a dependent chain per owner, so the chains of the groups overlap in the warp's instruction stream.
        // Group g takes over the state of its owner exactly as the whole warp does below; its lane l evaluates the owner's
        // l-th cached segment with exact squares (closest_eval_nb: no flagged squares, so the reference's strict-< scan is the
        // first minimum in list order: a segmented butterfly minimum of the order-preserving keys and a ballot); the push is
        // the same push_eval / push_commit on the same inputs.  Owners whose list does not fit a group (or overflowed) are
        // left to the whole-warp service.
        if (todo & (todo - 1)) {
            int W = (__popc(todo) == 2) ? 16 : 8;
            unsigned elig = __ballot_sync(FULL_MASK, need && nseg >= 0 && nseg <= W);
            if (W == 8 && !(elig & (elig - 1))) { W = 16; elig = __ballot_sync(FULL_MASK, need && nseg >= 0 && nseg <= W); }
            if (elig & (elig - 1)) {
                const int G = 32 / W, grp = lane / W, l = lane - grp * W;
                const unsigned gmask = (W == 16 ? 0xffffu : 0xffu) << (grp * W);
                const unsigned fo = __fns(elig, 0, grp + 1);                 // the grp-th eligible owner (0xffffffff: none left)
                const bool act = fo < 32u;
                const int o = act ? (int)fo : lane;
                const int ns_o = __shfl_sync(FULL_MASK, nseg, o);
                const int ns = act ? ns_o : 0;
                const SegC g = seg_decode((act && l < ns) ? wp.lseg[o][l] : 0u);
                const bool served = need && ((elig >> lane) & 1u) && __popc(elig & ((1u << lane) - 1u)) < G;
                COOP_STATE(COOP_PARK)
                wp.seg[lane] = (uint32_t)n.i[I_CEILING_COUNT]; wp.seg[32 + lane] = (uint32_t)n.i[I_FLOOR_COUNT];
                __syncwarp();
                COOP_STATE(COOP_TAKE)
                n.i[I_CEILING_COUNT] = (int)wp.seg[o]; n.i[I_FLOOR_COUNT] = (int)wp.seg[32 + o];
                __syncwarp();                                                // every lane has taken before a leader writes back
                bool live = act;
                int it;
                #pragma unroll 1
                for (it = 2; it < 32; it++) {
                    ncoop++;
                    const double inf = __longlong_as_double(0x7ff0000000000000LL);
                    double d2 = inf, a = 0, b = 0; int r = 0; bool ok = false;
                    if (live && l < ns) ok = closest_eval_nb(g, xpos, ypos, d2, a, b, r);
                    const long long bits = __double_as_longlong(d2);
                    const unsigned long long key = ok ? (unsigned long long)(bits ^ ((bits >> 63) | (long long)0x8000000000000000ULL)) : ~0ULL;   // monotone in d2
                    unsigned long long kmin = key;
                    #pragma unroll
                    for (int d = 1; d < 16; d <<= 1) {
                        if (d < W) { const unsigned long long other = __shfl_xor_sync(FULL_MASK, kmin, d); kmin = other < kmin ? other : kmin; }
                    }
                    const unsigned cand = __ballot_sync(FULL_MASK, ok && key == kmin) & gmask;
                    const int w = cand ? __ffs(cand) - 1 : lane;             // first minimum in list order (strict < scan, physics.py:170-176)
                    a = __shfl_sync(FULL_MASK, a, w); b = __shfl_sync(FULL_MASK, b, w); r = __shfl_sync(FULL_MASK, r, w);
                    const Push t = push_eval(cand ? r : 0, a, b);
                    if (live && t.go) push_commit(t); else live = false;
                    if (!__any_sync(FULL_MASK, live)) break;
                }
                // the group's first lane hands the owner's state back through the owner's slots; a chain that ran all 30
                // iterations marks a wedged ninja (sim_variant.inc)
                if (act && l == 0) {
#define COOP_PUT(k, v) coop_slot(wp, k, o) = v;
                    COOP_STATE(COOP_PUT)
#undef COOP_PUT
                    wp.seg[o] = (uint32_t)n.i[I_CEILING_COUNT]; wp.seg[32 + o] = (uint32_t)n.i[I_FLOOR_COUNT];
                    wp.aux[o] = (uint8_t)((it == 32 && live) ? 1 : 0);
                }
                __syncwarp();
                COOP_STATE(COOP_BACK)
                n.i[I_CEILING_COUNT] = (int)wp.seg[lane]; n.i[I_FLOOR_COUNT] = (int)wp.seg[32 + lane];
                pushed = pushed_own;
                if (served) { need = false; if (wp.aux[lane]) c.wedged = 1; }
                __syncwarp();
                continue;
            }
        }
#endif
        const int o = __ffs(todo) - 1;
        const int ns = __shfl_sync(FULL_MASK, nseg, o);
        const SegC g = seg_decode(lane < ns ? wp.lseg[o][lane] : 0u);
        COOP_STATE(COOP_PARK)
        wp.seg[lane] = (uint32_t)n.i[I_CEILING_COUNT]; wp.seg[32 + lane] = (uint32_t)n.i[I_FLOOR_COUNT];
        __syncwarp();
        COOP_STATE(COOP_TAKE)
        n.i[I_CEILING_COUNT] = (int)wp.seg[o]; n.i[I_FLOOR_COUNT] = (int)wp.seg[32 + o];
        COOP_T(cacc, 5, ct)
        int w1 = -1, w2 = -1;                                // winning lanes of the last two calls (-1: unknown)
        int it;
        #pragma unroll 1
        for (it = it_first; it < 32; it++) {
            ncoop++;
            double a = 0, b = 0; int result = 0, w = -1;
            if (ns < 0) {                                    // list overflow: the owner's literal walk, then broadcast
                double ea = 0, eb = 0;
                if (lane == o) result = closest_point_exact(c.cell_hdr, c.segs, gx1, gy1, gx2, gy2, xpos, ypos, ea, eb);
                a = __shfl_sync(FULL_MASK, ea, o); b = __shfl_sync(FULL_MASK, eb, o); result = __shfl_sync(FULL_MASK, result, o);
            }
#ifndef NPP_COOP_NOPRED
            // Winner prediction.  A wedged ninja is pushed out of two segments in turn, so the closest segment of this call is
            // the one of two calls back in 96 % of the calls (tools/iter_stats.py).  Every lane takes the predicted lane's
            // closest point (two shuffles) and evaluates the push from it; ONE vote, issued behind the push's square root,
            // checks that no other lane can win the reference's strict-< scan.  The push is committed only after the vote; on a
            // miss the full selection (redux + ballots, ordered fold for flagged squares) runs and the push is evaluated again
            // from the real winner.  The selection's dependent collectives leave the chain evaluate -> select -> push; results
            // are those of the unpredicted loop by construction (profiles/r01_d_step_kernel.md).
            else if (w2 >= 0) {
                const double inf = __longlong_as_double(0x7ff0000000000000LL);
                double d2 = inf, ea = 0, eb = 0; int er = 0; bool ix = false, ok = false;
                if (lane < ns) ok = closest_eval(g, xpos, ypos, d2, ea, eb, er, ix);
                if (!ok) { d2 = inf; er = 0; ix = false; }
                const double sa = __shfl_sync(FULL_MASK, ea, w2), sb = __shfl_sync(FULL_MASK, eb, w2);
                const int scode = __shfl_sync(FULL_MASK, (er & 3) | (ix ? 4 : 0), w2);
                const int sr = (scode & 3) == 3 ? -1 : (scode & 3);
                COOP_T(cacc, 0, ct)
                // The speculative push, flat: push_eval's expressions without its early returns.
                Push t;
                t.dx = xpos - sa; t.dy = ypos - sb;
                if (fabs(t.dx) <= 0.0000001) {                // ninja.py:312-318 band-aid
                    t.dx = 0;
                    if (xpos == 50.51197510492316 || xpos == 49.23232124849253) t.dx = -0x1p-47;
                    if (xpos == 49.153536108584795) t.dx = 0x1p-47;
                }
                const double dist_sq = t.dx * t.dx + t.dy * t.dy;
                const double dist = SQRT(dist_sq);
                // The prediction holds iff no other lane can beat the predicted one in the reference's strict-< scan: it has the
                // same closest point and side (then it does not matter which of the two wins), or a larger d2 (or the same d2 later
                // in the list) -- beyond closest_fold's guard band when one of the two squares was flagged inexact.  ONE vote,
                // issued behind the square root and read after the push.
                const double sd = __shfl_sync(FULL_MASK, d2, w2);
                const bool same = er == sr && __double_as_longlong(ea) == __double_as_longlong(sa) && __double_as_longlong(eb) == __double_as_longlong(sb);
                const bool loses = (ix || (scode & 4)) ? (d2 > sd + 1e-10) : (d2 > sd || (d2 == sd && lane > w2));
                const unsigned bad = __ballot_sync(FULL_MASK, ok && lane != w2 && !same && !loses);
                t.depen_len = NINJA_RADIUS - (sr < 0 ? -dist : dist);
                t.inv_dist = DIV(1.0, dist);
                t.norm_dx = t.dx * t.inv_dist; t.norm_dy = t.dy * t.inv_dist;
                t.go = sr != 0 && !(dist_sq < 1e-16) && !(t.depen_len < 0.0000001);
                COOP_T(cacc, 2, ct)
                w = w2;
                if (bad) {                                   // mispredicted (or a flagged square): the full selection, the push again
                    result = warp_select_uniform(wp, c, o, ns, ok, d2, ea, eb, er, ix, gx1, gy1, gx2, gy2, xpos, ypos, a, b, w);
                    t = push_eval(result, a, b);
                    // the ordered fold names no lane: keep predicting when it confirmed the predicted lane's point
                    if (w < 0 && result == sr && __double_as_longlong(a) == __double_as_longlong(sa) && __double_as_longlong(b) == __double_as_longlong(sb)) w = w2;
                }
                COOP_T(cacc, 1, ct)
                w2 = w1; w1 = w;
                if (!t.go) break;
                push_commit(t);
                continue;
            }
#endif
            else {
                result = warp_closest_uniform(wp, c, o, ns, g, gx1, gy1, gx2, gy2, xpos, ypos, a, b, w, cacc, ct);
            }
            COOP_T(cacc, 1, ct)
            w2 = w1; w1 = w;
            const bool go = depen(result, a, b);             // uniform: same inputs on every lane
            COOP_T(cacc, 2, ct)
            if (!go) break;
        }
        __syncwarp();
        if (it == 32 && lane == o) c.wedged = 1;             // all 32 iterations used: a wedged ninja (handoff, sim_variant.inc)
        if (lane != o) {
            COOP_STATE(COOP_BACK)
            n.i[I_CEILING_COUNT] = (int)wp.seg[lane]; n.i[I_FLOOR_COUNT] = (int)wp.seg[32 + lane];
            pushed = pushed_own;
        } else need = false;
        __syncwarp();
        COOP_T(cacc, 5, ct)
    }
    if ((threadIdx.x & 31) == 0) wp.coop += ncoop;
#ifdef NPP_PROFILE
    if ((threadIdx.x & 31) == 0 && ncoop) { for (int k = 0; k < 6; k++) if (k != 3) atomicAdd(&g_coop_cycles[k], (unsigned long long)cacc[k]); atomicAdd(&g_coop_cycles[3], (unsigned long long)ncoop); }
    if ((threadIdx.x & 31) == 0 && ncoop && wp.group < 0) { wp.sph[4] += (unsigned long long)(cacc[0] + cacc[1] + cacc[2]); wp.sph[5] += (unsigned long long)cacc[5]; }
#endif
#undef COOP_STATE
#undef COOP_PARK
#undef COOP_TAKE
#undef COOP_BACK
    PH_MARK(6)
    if (alive) { n.f[F_XPOS] = xpos; n.f[F_YPOS] = ypos; n.f[F_XSPEED] = xspeed; n.f[F_YSPEED] = yspeed; }
    return !pushed && __double_as_longlong(xpos) == __double_as_longlong(xin) && __double_as_longlong(ypos) == __double_as_longlong(yin);
}

// literal per-thread wall test (ninja.py:422-441) with exact squares: fallback of the pooled version
__device__ __noinline__ double wall_terms_exact(const uint32_t* cell_hdr, const uint32_t* segs, double xpos, double ypos, double wall_normal) {
    Ctx c; c.cell_hdr = cell_hdr; c.segs = segs;
    const double rad = NINJA_RADIUS + 0.1;
    query_region(c, xpos - rad, ypos - rad, xpos + rad, ypos + rad, [&](uint32_t s) {
        bool back; double a, b;
        seg_closest_point(s, xpos, ypos, back, a, b);
        double dx = xpos - a, dy = ypos - b;
        double dist = sqrt(SQ(dx) + SQ(dy));
        if (fabs(dy) < 0.00001 && 0 < dist && dist <= rad) wall_normal += dx / dist;
        return true;
    });
    return wall_normal;
}

// ninja.py:381-537, warp-synchronous
DEV void ninja_post_collision(Nin& n, Ctx& c, WarpPool& wp, bool alive) {
    PH_BEGIN
    double wall_normal = 0;
    if (alive) {
        gather_entities(c, n.f[F_XPOS], n.f[F_YPOS], [&](int id) {
            if (ent_kind(c, id) != NPP_K_BOOST) ent_logical_collision(n, c, id, wall_normal);   // boost pads are not logical-collidable
        });
        commit_pending_doors(c);
    }
    PH_MARK(9)
    NPP_BLOCK_SYNC();
    const double wall_entities = wall_normal;
    bool uncertain;
    pooled_wall_test(wp, c, alive, n.f[F_XPOS], n.f[F_YPOS], wall_normal, uncertain);
    PH_MARK(10)
    if (!alive) return;
    if (uncertain) wall_normal = wall_terms_exact(c.cell_hdr, c.segs, n.f[F_XPOS], n.f[F_YPOS], wall_entities);
    n.i[I_AIRBORN_OLD] = n.i[I_AIRBORN];
    n.i[I_AIRBORN] = 1; n.i[I_WALLED] = 0;
    if (wall_normal != 0) { n.i[I_WALLED] = 1; n.f[F_WALL_NORMAL] = DIV(wall_normal, fabs(wall_normal)); }
    if (n.i[I_FLOOR_COUNT] > 0) {
        n.i[I_AIRBORN] = 0;
        double floor_scalar = SQRT(SQ(n.f[F_FLOOR_NX]) + SQ(n.f[F_FLOOR_NY]));
        if (floor_scalar == 0) { n.f[F_FNX] = 0; n.f[F_FNY] = -1; }
        else { n.f[F_FNX] = DIV(n.f[F_FLOOR_NX], floor_scalar); n.f[F_FNY] = DIV(n.f[F_FLOOR_NY], floor_scalar); }
        if (n.i[I_STATE] != 8 && n.i[I_AIRBORN_OLD]) {
            double impact_vel = -(n.f[F_FNX] * n.f[F_XSPEED_OLD] + n.f[F_FNY] * n.f[F_YSPEED_OLD]);
            if (impact_vel > MAX_SURVIVABLE_IMPACT - 4.0 / 3 * fabs(n.f[F_FNY])) {
                n.f[F_XSPEED] = n.f[F_XSPEED_OLD]; n.f[F_YSPEED] = n.f[F_YSPEED_OLD];
                ninja_kill(n, c, n.f[F_XPOS], n.f[F_YPOS], n.f[F_XSPEED] * 0.5, n.f[F_YSPEED] * 0.5, 2);
                GI(c, I_TERMINAL_IMPACT) = 1;
            }
        }
    }
    if (n.i[I_AIRBORN]) n.i[I_FRAMES_AIRBORNE] += 1; else n.i[I_FRAMES_AIRBORNE] = 0;
    if (!n.i[I_AIRBORN]) n.i[I_CONSEC_FLOOR] += 1; else n.i[I_CONSEC_FLOOR] = 0;
    if (n.i[I_WALLED]) n.i[I_CONSEC_WALL] += 1; else n.i[I_CONSEC_WALL] = 0;
    if (n.i[I_CEILING_COUNT] > 0) {
        double ceiling_scalar = SQRT(SQ(n.f[F_CEIL_NX]) + SQ(n.f[F_CEIL_NY]));
        double cnx, cny;
        if (ceiling_scalar == 0) { cnx = 0; cny = 1; }
        else { cnx = DIV(n.f[F_CEIL_NX], ceiling_scalar); cny = DIV(n.f[F_CEIL_NY], ceiling_scalar); }
        GF(c, F_CNX) = cnx; GF(c, F_CNY) = cny;
        if (n.i[I_STATE] != 8) {
            double impact_vel = -(cnx * n.f[F_XSPEED_OLD] + cny * n.f[F_YSPEED_OLD]);
            if (impact_vel > MAX_SURVIVABLE_IMPACT - 4.0 / 3 * fabs(cny)) {
                n.f[F_XSPEED] = n.f[F_XSPEED_OLD]; n.f[F_YSPEED] = n.f[F_YSPEED_OLD];
                ninja_kill(n, c, n.f[F_XPOS], n.f[F_YPOS], n.f[F_XSPEED] * 0.5, n.f[F_YSPEED] * 0.5, 2);
                GI(c, I_TERMINAL_IMPACT) = 1;
            }
        }
    }
    if (n.i[I_CRUSHABLE] && n.f[F_CRUSH_LEN] > 0) {
        if (DIV(SQRT(SQ(n.f[F_XCRUSH]) + SQ(n.f[F_YCRUSH])), n.f[F_CRUSH_LEN]) < MIN_SURVIVABLE_CRUSHING)
            ninja_kill(n, c, n.f[F_XPOS], n.f[F_YPOS], 0, 0, 0);
    }
}

DEV void ninja_floor_jump(Nin& n, bool was_buffered, int hor_input, int& ljb) {                   // ninja.py:539-579
    n.i[I_JUMP_BUFFER] = -1; n.i[I_FLOOR_BUFFER] = -1; n.i[I_LP_BUFFER] = -1;
    if (was_buffered) ljb = 1;
    n.i[I_STATE] = 3;
    n.f[F_GRAVITY] = GRAVITY_JUMP;
    double jx, jy;
    if (n.f[F_FNX] == 0) { jx = 0; jy = -2; }
    else {
        double dx = n.f[F_FNX], dy = n.f[F_FNY];
        if (n.f[F_XSPEED] * dx >= 0) {
            if (n.f[F_XSPEED] * hor_input >= 0) { jx = 2.0 / 3 * dx; jy = 2 * dy; }
            else { jx = 0; jy = -1.4; }
        } else {
            if (n.f[F_XSPEED] * hor_input > 0) { jx = 0; jy = -1.4; }
            else { n.f[F_XSPEED] = 0; jx = 2.0 / 3 * dx; jy = 2 * dy; }
        }
    }
    if (n.f[F_YSPEED] > 0) n.f[F_YSPEED] = 0;
    n.f[F_XSPEED] += jx; n.f[F_YSPEED] += jy;
    n.f[F_XPOS] += jx; n.f[F_YPOS] += jy;
    n.i[I_JUMP_DURATION] = 0;
}
DEV void ninja_wall_jump(Nin& n, bool was_buffered, int hor_input, int& ljb) {                    // ninja.py:581-608
    double jx, jy;
    if (hor_input * n.f[F_WALL_NORMAL] < 0 && n.i[I_STATE] == 5) { jx = 2.0 / 3; jy = -1; }
    else { jx = 1; jy = -1.4; }
    if (was_buffered) ljb = 1;
    n.i[I_STATE] = 3;
    n.f[F_GRAVITY] = GRAVITY_JUMP;
    if (n.f[F_XSPEED] * n.f[F_WALL_NORMAL] < 0) n.f[F_XSPEED] = 0;
    if (n.f[F_YSPEED] > 0) n.f[F_YSPEED] = 0;
    n.f[F_XSPEED] += jx * n.f[F_WALL_NORMAL]; n.f[F_YSPEED] += jy;
    n.f[F_XPOS] += jx * n.f[F_WALL_NORMAL]; n.f[F_YPOS] += jy;
    n.i[I_JUMP_BUFFER] = -1; n.i[I_WALL_BUFFER] = -1; n.i[I_LP_BUFFER] = -1;
    n.i[I_JUMP_DURATION] = 0;
}
DEV void ninja_lp_jump(Nin& n, const Ctx& c) {                                         // ninja.py:610-626
    n.i[I_FLOOR_BUFFER] = -1; n.i[I_WALL_BUFFER] = -1; n.i[I_JUMP_BUFFER] = -1; n.i[I_LP_BUFFER] = -1;
    const double xlp = GF(c, F_XLP), ylp = GF(c, F_YLP);
    double boost_scalar = 2 * fabs(xlp) + 2;
    if (boost_scalar == 2) boost_scalar = 1.7;
    n.f[F_XSPEED] += xlp * boost_scalar * (2.0 / 3);
    n.f[F_YSPEED] += ylp * boost_scalar * (2.0 / 3);
}

DEV void ninja_think_body(Nin& n, const Ctx& c, const int hor_input, const int jump_input, int& ljb) {   // ninja.py:849-1059
    if (n.i[I_STATE] != n.i[I_PREVIOUS_STATE]) { n.i[I_STATE_CHANGE_FRAME] = 0; n.i[I_PREVIOUS_STATE] = n.i[I_STATE]; }
    else n.i[I_STATE_CHANGE_FRAME] += 1;
    bool new_jump_check = jump_input ? (n.i[I_JUMP_INPUT_OLD] == 0) : false;
    n.i[I_JUMP_INPUT_OLD] = jump_input;
    int b;
    b = n.i[I_LP_BUFFER];    b = (-1 < b && b < 3) ? b + 1 : -1; n.i[I_LP_BUFFER] = b;    bool in_lp_buffer = -1 < b && b < 4;
    b = n.i[I_JUMP_BUFFER];  b = (-1 < b && b < 5) ? b + 1 : -1; n.i[I_JUMP_BUFFER] = b;  bool in_jump_buffer = -1 < b && b < 5;
    b = n.i[I_WALL_BUFFER];  b = (-1 < b && b < 5) ? b + 1 : -1; n.i[I_WALL_BUFFER] = b;  bool in_wall_buffer = -1 < b && b < 5;
    b = n.i[I_FLOOR_BUFFER]; b = (-1 < b && b < 5) ? b + 1 : -1; n.i[I_FLOOR_BUFFER] = b; bool in_floor_buffer = -1 < b && b < 5;
    bool just = false;
    if (new_jump_check && n.i[I_AIRBORN]) { n.i[I_JUMP_BUFFER] = 0; just = true; }
    if (n.i[I_WALLED]) n.i[I_WALL_BUFFER] = 0;
    if (!n.i[I_AIRBORN]) n.i[I_FLOOR_BUFFER] = 0;
    if (jump_input) {
        bool has_active;
        if (just) has_active = true;
        else { int jb = n.i[I_JUMP_BUFFER]; has_active = (-1 < jb && jb < 5) || in_floor_buffer || in_wall_buffer || in_lp_buffer; }
        GI(c, I_JUMP_USELESS) = (n.i[I_AIRBORN] && n.i[I_STATE] != 3 && !has_active) ? 1 : 0;
    } else GI(c, I_JUMP_USELESS) = 0;

    int state = n.i[I_STATE];
    if (state == 6 || state == 9) return;
    if (state == 7) { n.i[I_STATE] = 6; return; }                 // think_awaiting_death :1061-1063
    if (state == 8) { n.f[F_DRAG] = n.i[I_AIRBORN] ? DRAG_REGULAR : DRAG_SLOW; return; }

    if (!n.i[I_AIRBORN]) {
        double xspeed_new = n.f[F_XSPEED] + GROUND_ACCEL * hor_input;
        if (fabs(xspeed_new) < MAX_HOR_SPEED) n.f[F_XSPEED] = xspeed_new;
        if (state > 2) {
            if (state == 3) n.f[F_GRAVITY] = GRAVITY_FALL;
            state = (n.f[F_XSPEED] * hor_input <= 0) ? 2 : 1;
            n.i[I_STATE] = state;
        }
        if (!in_jump_buffer && !new_jump_check) {
            if (state == 2) {
                double projection = fabs(n.f[F_YSPEED] * n.f[F_FNX] - n.f[F_XSPEED] * n.f[F_FNY]);
                if (hor_input * projection * n.f[F_XSPEED] > 0) { n.i[I_STATE] = 1; return; }
                if (projection < 0.1 && n.f[F_FNX] == 0) { n.i[I_STATE] = 0; return; }
                if (n.f[F_YSPEED] < 0 && n.f[F_FNX] != 0) {
                    double speed_scalar = SQRT(SQ(n.f[F_XSPEED]) + SQ(n.f[F_YSPEED]));
                    double fric_force = fabs(n.f[F_XSPEED] * (1 - FRICTION_GROUND) * n.f[F_FNY]);
                    double fric_force2 = speed_scalar - fric_force * SQ(n.f[F_FNY]);
                    n.f[F_XSPEED] = DIV(n.f[F_XSPEED], speed_scalar) * fric_force2;
                    n.f[F_YSPEED] = DIV(n.f[F_YSPEED], speed_scalar) * fric_force2;
                    return;
                }
                n.f[F_XSPEED] *= FRICTION_GROUND;
                return;
            }
            if (state == 1) {
                double projection = fabs(n.f[F_YSPEED] * n.f[F_FNX] - n.f[F_XSPEED] * n.f[F_FNY]);
                if (hor_input * projection * n.f[F_XSPEED] > 0) {
                    if (hor_input * n.f[F_FNX] >= 0) return;
                    if (fabs(xspeed_new) < MAX_HOR_SPEED) {
                        double boost = GROUND_ACCEL / 2 * hor_input;
                        double xboost = boost * n.f[F_FNY] * n.f[F_FNY];
                        double yboost = boost * n.f[F_FNY] * -n.f[F_FNX];
                        n.f[F_XSPEED] += xboost; n.f[F_YSPEED] += yboost;
                    }
                    return;
                }
                n.i[I_STATE] = 2;
            } else {
                if (hor_input) { n.i[I_STATE] = 1; return; }
                double projection = fabs(n.f[F_YSPEED] * n.f[F_FNX] - n.f[F_XSPEED] * n.f[F_FNY]);
                if (projection < 0.1) { n.f[F_XSPEED] *= FRICTION_GROUND_SLOW; return; }
                n.i[I_STATE] = 2;
            }
            return;
        }
        ninja_floor_jump(n, false, hor_input, ljb);
        return;
    } else {
        double xspeed_new = n.f[F_XSPEED] + AIR_ACCEL * hor_input;
        if (fabs(xspeed_new) < MAX_HOR_SPEED) n.f[F_XSPEED] = xspeed_new;
        if (state < 3) { n.i[I_STATE] = 4; return; }
        if (state == 3) {
            n.i[I_JUMP_DURATION] += 1;
            if (!jump_input || n.i[I_JUMP_DURATION] > MAX_JUMP_DURATION) { n.f[F_GRAVITY] = GRAVITY_FALL; n.i[I_STATE] = 4; return; }
        }
        if (in_jump_buffer || new_jump_check) {
            if (n.i[I_WALLED] || in_wall_buffer) { ninja_wall_jump(n, in_wall_buffer, hor_input, ljb); return; }
            if (in_floor_buffer) { ninja_floor_jump(n, true, hor_input, ljb); return; }
            if (in_lp_buffer && new_jump_check) { ninja_lp_jump(n, c); return; }
        }
        if (!n.i[I_WALLED]) { if (state == 5) n.i[I_STATE] = 4; }
        else {
            if (state == 5) {
                if (hor_input * n.f[F_WALL_NORMAL] <= 0) n.f[F_YSPEED] *= FRICTION_WALL;
                else n.i[I_STATE] = 4;
            } else {
                if (n.f[F_YSPEED] > 0 && hor_input * n.f[F_WALL_NORMAL] < 0) {
                    if (state == 3) n.f[F_GRAVITY] = GRAVITY_FALL;
                    n.i[I_STATE] = 5;
                }
            }
        }
    }
}

DEV void ninja_think(Nin& n, const Ctx& c, int hor_input, int jump_input) {
    int ljb = 0;                                   // last_jump_was_buffered = False at the start of each frame
    ninja_think_body(n, c, hor_input, jump_input, ljb);
    GI(c, I_LAST_JUMP_BUFFERED) = ljb;
}

// Simulator.tick: nsim.py:221-292 (headless: enable_anim, log_data false).  Warp-synchronous: all 32
// lanes call it together; `run` lanes advance their env by one frame.
// s_resume > 0 (solo warps, warp-uniform): the tick was started by the env's home lane, which left it after substep
// s_resume - 1 (mid-tick handoff, below); everything before the substep loop has run and is in the env's state.
DEV void sim_tick(Nin& n, Ctx& c, WarpPool& wp, bool run, int hor_input, int jump_input, int s_resume = 0) {
    bool alive = false, phys = false;
    double gx = 0, gy = 0;
    PH_BEGIN
#if NPP_SOLO
    if (run && s_resume > 0) { alive = true; c.wedged = 1; }       // (only a live, wedged ninja is handed over)
    else
#endif
    if (run) {
        c.wedged = 0;
        n.i[I_FRAME] += 1;
        GI(c, I_HOR_INPUT) = hor_input; GI(c, I_JUMP_INPUT) = jump_input;
        GF(c, F_X_AT_ACTION) = n.f[F_XPOS]; GF(c, F_Y_AT_ACTION) = n.f[F_YPOS];   // ninja.py:841-847
        // all move() then all think(): roaming entities in slot (= entity_dic) order, mines and boost
        // pads by neighbourhood (they are independent of every other entity's update)
        dyn_move_phase(c);
        PH_MARK(0)
        static_update_phase(n, c);
        PH_MARK(1)
        dyn_think_phase(n, c);
        PH_MARK(2)
        int state = n.i[I_STATE];
        alive = (state != 9 && state != 6);      // dead: ragdoll is None when enable_anim=False (ninja.py:122-128)
        if (alive) {
            ninja_integrate(n);
            phys = ninja_pre_collision(n, c, gx, gy);
        }
    }
    PH_MARK(3)
    // The four substeps (nsim.py:263-267).  A substep without entity pushes that leaves the position
    // bit-identical and performs no depenetration is a pure function of unchanged inputs, so the remaining
    // substeps would repeat it verbatim: they are skipped (typical: free flight, or resting after substep 2).
    bool sub = alive;
    #pragma unroll 1
    for (int s = (NPP_SOLO ? s_resume : 0); s < 4 && NPP_BLOCK_ANY(sub); s++) {
        if (HAS_ANY(NPP_PHYS_MASK) && sub && phys) ninja_collide_vs_objects(n, c, gx, gy);
        bool idle = ninja_collide_vs_tiles(n, c, wp, sub);
        if (idle && !phys) sub = false;
#if NPP_SOLO
        // Mid-tick handoff: a substep that used all 32 iterations marks a wedged ninja; the three substeps that follow are
        // the same 30-iteration chain again.  The home lane leaves the tick HERE -- between two substeps everything the
        // rest of the tick reads is in the ninja record and the env's tables (nothing physical on these levels: no phys /
        // gx / gy) -- reserves a queue entry and sits out; step_kernel stores the state and publishes (env, tick, s + 1).
        if (!HAS_ANY(NPP_PHYS_MASK) && sub && s < 3 && c.wedged && c.hq && *reinterpret_cast<const volatile int32_t*>(c.hq) < c.hand_cap) {
            const int slot = atomicAdd(c.hq, 1);
            if (slot < NPPSIM_HANDOFF_CAP) { c.mid_slot = slot; c.mid_s = s + 1; sub = false; alive = false; run = false; }
        }
#endif
    }
    PH_MARK(4)
    NPP_BLOCK_SYNC();
    ninja_post_collision(n, c, wp, alive);
    PH_MARK(7)
    NPP_BLOCK_SYNC();
    if (run && n.i[I_STATE] != 9) ninja_think(n, c, hor_input, jump_input);
    PH_MARK(8)
}
