/*
 * nsim_oracle.c -- CPU restatement of nclone's headless Simulator.tick path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the checker, never the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load the library built from it.  Nothing under
 * nclone_b200/ links, imports or executes it.
 *
 * It follows the reference one function at a time (citations are path:line
 * under /root/reference/nclone/), single environment, plain serial C, AoS
 * structs -- deliberately the opposite layout from the CUDA path so that the
 * two are independent statements of the same algorithm.
 *
 * Parity status: PINNED.  oracle/gen_golden.py runs the unmodified reference
 * (imported from /root/reference) and tests/test_oracle_golden.py checks this
 * file against the resulting per-frame records of all 130 bc_replays and of
 * the generated-map random-action trajectories, bit for bit.
 *
 * Arithmetic: IEEE double, built with -ffp-contract=off.  Python's `x**2` on a
 * float is libm pow(x, 2.0) (CPython floatobject.c:float_pow), so SQ() calls
 * pow() and the file must be compiled with -fno-builtin-pow (gcc would fold
 * pow(x,2) to x*x, which differs in ~1e-3 of inputs on this glibc).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SQ(x) pow((x), 2.0)

#define NCX 44
#define NCY 25
#define NHX 89
#define NHY 51
#define MAX_ENTS 4096
#define NINJA_RADIUS 10.0

/* constants/physics_constants.py:54-60, 282-329 */
static const double GRAVITY_FALL = 0.06666666666666665;
static const double GRAVITY_JUMP = 0.01111111111111111;
static const double GROUND_ACCEL = 0.06666666666666665;
static const double AIR_ACCEL = 0.04444444444444444;
static const double DRAG_REGULAR = 0.9933221725495059;
static const double DRAG_SLOW = 0.8617738760127536;
static const double FRICTION_GROUND = 0.9459290248857720;
static const double FRICTION_GROUND_SLOW = 0.8617738760127536;
static const double FRICTION_WALL = 0.9113380468927672;
static const double MAX_HOR_SPEED = 3.333;
#define MAX_JUMP_DURATION 45
static const double MAX_SURVIVABLE_IMPACT = 6.0;
static const double MIN_SURVIVABLE_CRUSHING = 0.05;

/* ------------------------------------------------------------------------- */
/* tile tables: tile_definitions.py:84-223                                    */
/* ------------------------------------------------------------------------- */
static const signed char GRID_EDGE[38][12] = {
    {0,0,0,0,0,0,0,0,0,0,0,0}, {1,1,0,0,1,1,1,1,0,0,1,1},
    {1,1,1,1,0,0,1,0,0,0,1,0}, {0,1,0,0,0,1,0,0,1,1,1,1},
    {0,0,1,1,1,1,0,1,0,0,0,1}, {1,0,0,0,1,0,1,1,1,1,0,0},
    {1,1,0,1,1,0,1,1,0,1,1,0}, {1,1,1,0,0,1,1,0,0,1,1,1},
    {0,1,1,0,1,1,0,1,1,0,1,1}, {1,0,0,1,1,1,1,1,1,0,0,1},
    {1,1,0,0,1,1,1,1,0,0,1,1}, {1,1,0,0,1,1,1,1,0,0,1,1},
    {1,1,0,0,1,1,1,1,0,0,1,1}, {1,1,0,0,1,1,1,1,0,0,1,1},
    {1,1,0,1,1,0,1,1,0,1,1,0}, {1,1,1,0,0,1,1,0,0,1,1,1},
    {0,1,1,0,1,1,0,1,1,0,1,1}, {1,0,0,1,1,1,1,1,1,0,0,1},
    {1,1,1,1,0,0,1,0,0,0,1,0}, {1,1,1,1,0,0,1,0,0,0,1,0},
    {0,0,1,1,1,1,0,1,0,0,0,1}, {0,0,1,1,1,1,0,1,0,0,0,1},
    {1,1,0,0,1,1,1,1,0,0,1,1}, {1,1,0,0,1,1,1,1,0,0,1,1},
    {1,1,0,0,1,1,1,1,0,0,1,1}, {1,1,0,0,1,1,1,1,0,0,1,1},
    {1,0,0,0,1,0,1,1,1,1,0,0}, {0,1,0,0,0,1,0,0,1,1,1,1},
    {0,1,0,0,0,1,0,0,1,1,1,1}, {1,0,0,0,1,0,1,1,1,1,0,0},
    {1,1,0,0,1,1,1,1,0,0,1,1}, {1,1,0,0,1,1,1,1,0,0,1,1},
    {1,1,0,0,1,1,1,1,0,0,1,1}, {1,1,0,0,1,1,1,1,0,0,1,1},
    {1,1,0,0,0,0,0,0,0,0,0,0}, {0,0,0,0,0,0,0,0,0,0,1,1},
    {0,0,0,0,1,1,0,0,0,0,0,0}, {0,0,0,0,0,0,1,1,0,0,0,0},
};
static const signed char SEG_ORTHO[38][12] = {
    {0,0,0,0,0,0,0,0,0,0,0,0}, {-1,-1,0,0,1,1,-1,-1,0,0,1,1},
    {-1,-1,1,1,0,0,-1,0,0,0,1,0}, {0,-1,0,0,0,1,0,0,-1,-1,1,1},
    {0,0,-1,-1,1,1,0,-1,0,0,0,1}, {-1,0,0,0,1,0,-1,-1,1,1,0,0},
    {-1,-1,0,0,0,0,-1,-1,0,0,0,0}, {-1,-1,0,0,0,0,0,0,0,0,1,1},
    {0,0,0,0,1,1,0,0,0,0,1,1}, {0,0,0,0,1,1,-1,-1,0,0,0,0},
    {-1,-1,0,0,0,0,-1,-1,0,0,0,0}, {-1,-1,0,0,0,0,0,0,0,0,1,1},
    {0,0,0,0,1,1,0,0,0,0,1,1}, {0,0,0,0,1,1,-1,-1,0,0,0,0},
    {-1,-1,0,0,0,0,-1,-1,0,0,0,0}, {-1,-1,0,0,0,0,0,0,0,0,1,1},
    {0,0,0,0,1,1,0,0,0,0,1,1}, {0,0,0,0,1,1,-1,-1,0,0,0,0},
    {-1,-1,0,0,0,0,-1,0,0,0,0,0}, {-1,-1,0,0,0,0,0,0,0,0,1,0},
    {0,0,0,0,1,1,0,0,0,0,0,1}, {0,0,0,0,1,1,0,-1,0,0,0,0},
    {-1,-1,0,0,0,0,-1,-1,0,0,1,0}, {-1,-1,0,0,0,0,-1,0,0,0,1,1},
    {0,0,0,0,1,1,0,-1,0,0,1,1}, {0,0,0,0,1,1,-1,-1,0,0,0,1},
    {-1,0,0,0,0,0,-1,-1,0,0,0,0}, {0,-1,0,0,0,0,0,0,0,0,1,1},
    {0,0,0,0,0,1,0,0,0,0,1,1}, {0,0,0,0,1,0,-1,-1,0,0,0,0},
    {-1,-1,0,0,1,0,-1,-1,0,0,0,0}, {-1,-1,0,0,0,1,0,0,0,0,1,1},
    {0,-1,0,0,1,1,0,0,0,0,1,1}, {-1,0,0,0,1,1,-1,-1,0,0,0,0},
    {-1,-1,0,0,0,0,0,0,0,0,0,0}, {0,0,0,0,0,0,0,0,0,0,1,1},
    {0,0,0,0,1,1,0,0,0,0,0,0}, {0,0,0,0,0,0,-1,-1,0,0,0,0},
};
/* tile id -> diagonal (x1,y1,x2,y2), -1 = none (tile_definitions.py:189-210) */
static const signed char SEG_DIAG[34][4] = {
    {-1,0,0,0},{-1,0,0,0},{-1,0,0,0},{-1,0,0,0},{-1,0,0,0},{-1,0,0,0},
    {0,24,24,0},{0,0,24,24},{24,0,0,24},{24,24,0,0},
    {-1,0,0,0},{-1,0,0,0},{-1,0,0,0},{-1,0,0,0},{-1,0,0,0},{-1,0,0,0},{-1,0,0,0},{-1,0,0,0},
    {0,12,24,0},{0,0,24,12},{24,12,0,24},{24,24,0,12},
    {0,24,24,12},{0,12,24,24},{24,0,0,12},{24,12,0,0},
    {0,24,12,0},{12,0,24,24},{24,0,12,24},{12,24,0,0},
    {12,24,24,0},{0,0,12,24},{12,0,0,24},{24,24,12,0},
};
/* tile id 10..17 -> centre x,y, quadrant hor,ver, convex (tile_definitions.py:214-223) */
static const signed char SEG_CIRC[8][5] = {
    {0,0,1,1,1},{24,0,-1,1,1},{24,24,-1,-1,1},{0,24,1,-1,1},
    {24,24,-1,-1,0},{0,24,1,-1,0},{0,0,1,1,0},{24,0,-1,1,0},
};

/* ------------------------------------------------------------------------- */
/* types                                                                       */
/* ------------------------------------------------------------------------- */
typedef struct {
    int kind;                 /* 0 linear (entities.py:9-96), 1 circular (entities.py:99-203) */
    double x1, y1, x2, y2, px, py, seg_lensq;
    int oriented;
    double cx, cy, hor, ver, radius;
    int convex;
    double phx, phy, pvx, pvy;
    double b[4];              /* min_x, min_y, max_x, max_y */
} Seg;

typedef struct { int *v; int n, cap; } IVec;

typedef struct {
    int present;              /* cell key exists in SpatialSegmentIndex.cell_data */
    double b[4];
    IVec segs;
} Cell;

enum { K_NONE = 0, K_MINE, K_GOLD, K_EXIT, K_EXIT_SWITCH, K_DOOR_REGULAR, K_DOOR_LOCKED,
       K_DOOR_TRAP, K_LAUNCH_PAD, K_ONE_WAY, K_DRONE_ZAP, K_BOUNCE, K_THWUMP, K_BOOST,
       K_DEATH_BALL, K_MINI_DRONE, K_SHWUMP };

typedef struct {
    int kind, type, dic_type, index, active;
    double xpos, ypos;
    int cellx, celly;
    int logical, physical, movable, thinkable;
    /* initial values for fast_reset (the *_initial_* attributes) */
    double ix, iy; int icx, icy, istate;
    int state;                /* mine / thwump / shwump */
    double radius;            /* mine RADIUS, drone RADIUS */
    int parent, switch_hit;
    int closed, is_vertical, open_timer, edge[2][2];
    double normal_x, normal_y;
    double xspeed, yspeed, xorigin, yorigin;
    double speed, xtarget, ytarget; int dir, mode; double grid_width;
    int is_horizontal, direction;
    int touching;
    double xdir, ydir; int activated;
} Ent;

typedef struct {
    double xpos, ypos, xspeed, yspeed, xpos_old, ypos_old, xspeed_old, yspeed_old;
    double applied_gravity, applied_drag, applied_friction;
    double floor_normalized_x, floor_normalized_y, ceiling_normalized_x, ceiling_normalized_y;
    double wall_normal, xlp_boost_normalized, ylp_boost_normalized;
    double floor_normal_x, floor_normal_y, ceiling_normal_x, ceiling_normal_y;
    double x_crush, y_crush, crush_len;
    double death_xpos, death_ypos, death_xspeed, death_yspeed;
    double xpos_at_action, ypos_at_action;
    double initial_xpos, initial_ypos;
    int state, airborn, airborn_old, walled, jump_input_old, jump_duration;
    int jump_buffer, floor_buffer, wall_buffer, launch_pad_buffer;
    int is_jump_useless, last_jump_was_buffered;
    int floor_count, wall_count, ceiling_count, is_crushable;
    int gold_collected, doors_opened, terminal_impact, death_cause;
    int frames_airborne, state_change_frame, previous_state;
    int consecutive_floor_frames, consecutive_wall_frames;
    int hor_input, jump_input;
    IVec cached_entities;
} Ninja;

typedef struct OSim {
    int frame;
    double *map_data; int map_len;
    double tile[NCX][NCY];
    Seg *segs; int nsegs, segcap;
    Cell cells[NCX][NCY];
    int hor_edge[NHX][NHY], ver_edge[NHX][NHY];
    int hor_segd[NHX][NHY], ver_segd[NHX][NHY];
    Ent *ents; int nents;
    int dic_order[MAX_ENTS]; int ndic;      /* entity_dic.values() flattened */
    IVec grid[NCX][NCY];                    /* grid_entity */
    int entity_counts[40]; int have_counts;
    Ninja ninja; int have_ninja;
    IVec tmp_segs, tmp_ents, movable, thinkable;
    int events;                              /* bit0 exit switch, bit1 locked door */
} OSim;

/* ------------------------------------------------------------------------- */
static void iv_push(IVec *a, int x) {
    if (a->n == a->cap) { a->cap = a->cap ? 2 * a->cap : 8; a->v = (int *)realloc(a->v, sizeof(int) * a->cap); }
    a->v[a->n++] = x;
}
static void iv_remove(IVec *a, int x) {    /* list.remove: first occurrence */
    for (int i = 0; i < a->n; i++) if (a->v[i] == x) { memmove(a->v + i, a->v + i + 1, sizeof(int) * (a->n - i - 1)); a->n--; return; }
}
static int clampi(int n, int a, int b) { return n < a ? a : (n > b ? b : n); }  /* physics.py:6-8 */
static int ifloor(double x) { return (int)floor(x); }

/* ------------------------------------------------------------------------- */
/* geometry: physics.py / entities.py                                          */
/* ------------------------------------------------------------------------- */
/* physics.py:247-261 */
static double toi_circle_vs_circle(double xpos, double ypos, double vx, double vy, double a, double b, double radius) {
    double dx = xpos - a, dy = ypos - b;
    double dist_sq = SQ(dx) + SQ(dy);
    double vel_sq = SQ(vx) + SQ(vy);
    double dot_prod = dx * vx + dy * vy;
    if (dist_sq - SQ(radius) > 0) {
        double radicand = SQ(dot_prod) - vel_sq * (dist_sq - SQ(radius));
        if (vel_sq > 0.0001 && dot_prod < 0 && radicand >= 0)
            return (-dot_prod - sqrt(radicand)) / vel_sq;
        return 1;
    }
    return 0;
}
/* physics.py:264-285 */
static double toi_circle_vs_lineseg(double xpos, double ypos, double dx, double dy, double a1, double b1, double a2, double b2, double radius) {
    double wx = a2 - a1, wy = b2 - b1;
    double seg_len = sqrt(SQ(wx) + SQ(wy));
    double nx = wx / seg_len, ny = wy / seg_len;
    double normal_proj = (xpos - a1) * ny - (ypos - b1) * nx;
    double hor_proj = (xpos - a1) * nx + (ypos - b1) * ny;
    if (fabs(normal_proj) >= radius) {
        double dir = dx * ny - dy * nx;
        if (dir * normal_proj < 0) {
            double q = (fabs(normal_proj) - radius) / fabs(dir);
            double t = (1 < q) ? 1 : q;                     /* min(q, 1) keeps q on ties */
            double hor_proj2 = hor_proj + t * (dx * nx + dy * ny);
            if (0 <= hor_proj2 && hor_proj2 <= seg_len) return t;
        }
    } else {
        if (0 <= hor_proj && hor_proj <= seg_len) return 0;
    }
    return 1;
}
/* physics.py:288-314 */
static double toi_circle_vs_arc(double xpos, double ypos, double vx, double vy, double a, double b, double hor, double ver, double radius_arc, double radius_circle) {
    double dx = xpos - a, dy = ypos - b;
    double dist_sq = SQ(dx) + SQ(dy);
    double vel_sq = SQ(vx) + SQ(vy);
    double dot_prod = dx * vx + dy * vy;
    double radius1 = radius_arc + radius_circle, radius2 = radius_arc - radius_circle;
    double t = 1;
    if (dist_sq > SQ(radius1)) {
        double radicand = SQ(dot_prod) - vel_sq * (dist_sq - SQ(radius1));
        if (vel_sq > 0.0001 && dot_prod < 0 && radicand >= 0)
            t = (-dot_prod - sqrt(radicand)) / vel_sq;
    } else if (dist_sq < SQ(radius2)) {
        double radicand = SQ(dot_prod) - vel_sq * (dist_sq - SQ(radius2));
        if (vel_sq > 0.0001) {
            double q = (-dot_prod + sqrt(radicand)) / vel_sq;
            t = (1 < q) ? 1 : q;
        }
    } else {
        t = 0;
    }
    if ((dx + t * vx) * hor > 0 && (dy + t * vy) * ver > 0) return t;
    return 1;
}
static double min3(double a, double b, double c) {          /* Python min(a,b,c): first minimum wins */
    double m = a;
    if (b < m) m = b;
    if (c < m) m = c;
    return m;
}
/* entities.py:82-96, 180-203 */
static double seg_intersect_with_ray(const Seg *s, double xpos, double ypos, double dx, double dy, double radius) {
    if (s->kind == 0) {
        double t1 = toi_circle_vs_circle(xpos, ypos, dx, dy, s->x1, s->y1, radius);
        double t2 = toi_circle_vs_circle(xpos, ypos, dx, dy, s->x2, s->y2, radius);
        double t3 = toi_circle_vs_lineseg(xpos, ypos, dx, dy, s->x1, s->y1, s->x2, s->y2, radius);
        return min3(t1, t2, t3);
    } else {
        double t1 = toi_circle_vs_circle(xpos, ypos, dx, dy, s->phx, s->phy, radius);
        double t2 = toi_circle_vs_circle(xpos, ypos, dx, dy, s->pvx, s->pvy, radius);
        double t3 = toi_circle_vs_arc(xpos, ypos, dx, dy, s->cx, s->cy, s->hor, s->ver, s->radius, radius);
        return min3(t1, t2, t3);
    }
}
/* entities.py:43-59, 127-157 */
static void seg_closest_point(const Seg *s, double xpos, double ypos, int *back, double *a, double *b) {
    if (s->kind == 0) {
        double dx = xpos - s->x1, dy = ypos - s->y1;
        double u = (dx * s->px + dy * s->py) / s->seg_lensq;
        if (0 > u) u = 0;          /* max(u, 0) */
        if (1 < u) u = 1;          /* min(u, 1) */
        *a = s->x1 + u * s->px;
        *b = s->y1 + u * s->py;
        *back = (dy * s->px - dx * s->py < 0) && s->oriented;
    } else {
        double dx = xpos - s->cx, dy = ypos - s->cy;
        *back = 0;
        if (dx * s->hor > 0 && dy * s->ver > 0) {
            double dist_sq = SQ(dx) + SQ(dy);
            double dist = sqrt(dist_sq);
            if (dist == 0) {       /* unreachable: the branch needs dx,dy != 0 */
                if (dx * s->hor > dy * s->ver) { *a = s->phx; *b = s->phy; } else { *a = s->pvx; *b = s->pvy; }
                return;
            }
            *a = s->cx + s->radius * dx / dist;
            *b = s->cy + s->radius * dy / dist;
            *back = s->convex ? (dist < s->radius) : (dist > s->radius);
        } else {
            if (dx * s->hor > dy * s->ver) { *a = s->phx; *b = s->phy; } else { *a = s->pvx; *b = s->pvy; }
        }
    }
}

/* utils/spatial_segment_index.py:112-158 (the only path physics.py:39-52 takes) */
static void query_region(OSim *S, double x1, double y1, double x2, double y2, IVec *out) {
    out->n = 0;
    double min_x = (x2 < x1) ? x2 : x1, min_y = (y2 < y1) ? y2 : y1;
    double max_x = (x2 > x1) ? x2 : x1, max_y = (y2 > y1) ? y2 : y1;
    int cx0 = clampi(ifloor(min_x / 24), 0, 43), cx1 = clampi(ifloor(max_x / 24), 0, 43);
    int cy0 = clampi(ifloor(min_y / 24), 0, 24), cy1 = clampi(ifloor(max_y / 24), 0, 24);
    for (int xc = cx0; xc <= cx1; xc++)
        for (int yc = cy0; yc <= cy1; yc++) {
            Cell *c = &S->cells[xc][yc];
            if (!c->present) continue;
            /* _bounds_intersect(query, cell) spatial_segment_index.py:175-188 */
            if (max_x < c->b[0] || min_x > c->b[2] || max_y < c->b[1] || min_y > c->b[3]) continue;
            for (int k = 0; k < c->segs.n; k++) iv_push(out, c->segs.v[k]);
        }
}
/* physics.py:79-101 */
static void gather_entities(OSim *S, double xpos, double ypos, IVec *out) {
    out->n = 0;
    int cx = clampi(ifloor(xpos / 24), 0, 43), cy = clampi(ifloor(ypos / 24), 0, 24);
    int x0 = cx - 1 > 0 ? cx - 1 : 0, x1 = cx + 1 < 43 ? cx + 1 : 43;
    int y0 = cy - 1 > 0 ? cy - 1 : 0, y1 = cy + 1 < 24 ? cy + 1 : 24;
    for (int x = x0; x <= x1; x++)
        for (int y = y0; y <= y1; y++) {
            IVec *l = &S->grid[x][y];
            for (int k = 0; k < l->n; k++) if (S->ents[l->v[k]].active) iv_push(out, l->v[k]);
        }
}
/* physics.py:104-128 */
static double sweep_circle_vs_tiles(OSim *S, double xpos_old, double ypos_old, double dx, double dy, double radius) {
    double xpos_new = xpos_old + dx, ypos_new = ypos_old + dy;
    double width = radius + 1;
    double x1 = (xpos_old < xpos_new ? xpos_old : xpos_new) - width;
    double y1 = (ypos_old < ypos_new ? ypos_old : ypos_new) - width;
    double x2 = (xpos_old > xpos_new ? xpos_old : xpos_new) + width;
    double y2 = (ypos_old > ypos_new ? ypos_old : ypos_new) + width;
    query_region(S, x1, y1, x2, y2, &S->tmp_segs);
    double shortest = 1;
    for (int k = 0; k < S->tmp_segs.n; k++) {
        double t = seg_intersect_with_ray(&S->segs[S->tmp_segs.v[k]], xpos_old, ypos_old, dx, dy, radius);
        if (t == 0) return 0;
        if (t < shortest) shortest = t;
    }
    return shortest;
}
/* physics.py:131-180.  segs==NULL: gather first (death balls). */
#ifdef OSIM_STATS
int osim_last_winner;
long osim_stat_pred[4];
long osim_stat_arc[3];    /* calls 2..31: [0] an arc in the gathered list, [1] an arc inside the query box, [2] the winner is an arc */   /* calls 2..31 of loops that reach call 3+: [0] total, [1] winner == winner two calls back, [2] == previous winner, [3] neither */
#endif
static int get_single_closest_point(OSim *S, double xpos, double ypos, double radius, IVec *segs, double *pa, double *pb) {
    if (!segs) {
        query_region(S, xpos - radius, ypos - radius, xpos + radius, ypos + radius, &S->tmp_segs);
        segs = &S->tmp_segs;
    }
    double shortest = INFINITY; int result = 0;
#ifdef OSIM_STATS
    extern int osim_last_winner; osim_last_winner = -1;
#endif
    double qx0 = xpos - radius, qy0 = ypos - radius, qx1 = xpos + radius, qy1 = ypos + radius;
    for (int k = 0; k < segs->n; k++) {
        const Seg *s = &S->segs[segs->v[k]];
        if (s->b[2] < qx0 || s->b[0] > qx1 || s->b[3] < qy0 || s->b[1] > qy1) continue;
        int back; double a, b;
        seg_closest_point(s, xpos, ypos, &back, &a, &b);
        double d2 = SQ(xpos - a) + SQ(ypos - b);
        if (!back) d2 -= 0.1;
        if (d2 < shortest) { shortest = d2; *pa = a; *pb = b; result = back ? -1 : 1;
#ifdef OSIM_STATS
            osim_last_winner = k;
#endif
        }
    }
    return result;
}
/* physics.py:183-201; returns 1 and (nx,ny,pen,other) on hit */
static int penetration_square_vs_point(double sx, double sy, double px, double py, double semi, double *nx, double *ny, double *pen, double *other) {
    double dx = px - sx, dy = py - sy;
    double penx = semi - fabs(dx), peny = semi - fabs(dy);
    if (penx > 0 && peny > 0) {
        if (peny <= penx) { *nx = 0; *ny = dy < 0 ? -1 : 1; *pen = peny; *other = penx; }
        else { *nx = dx < 0 ? -1 : 1; *ny = 0; *pen = penx; *other = peny; }
        return 1;
    }
    return 0;
}
/* physics.py:204-207 */
static int overlap_circle_vs_circle(double x1, double y1, double r1, double x2, double y2, double r2) {
    double dist = sqrt(SQ(x1 - x2) + SQ(y1 - y2));
    return dist < r1 + r2;
}
/* physics.py:457-471 */
static int overlap_circle_vs_segment(double xpos, double ypos, double radius, double px1, double py1, double px2, double py2) {
    double px = px2 - px1, py = py2 - py1;
    double dx = xpos - px1, dy = ypos - py1;
    double seg_lensq = SQ(px) + SQ(py);
    double u = (dx * px + dy * py) / seg_lensq;
    if (0 > u) u = 0;
    if (1 < u) u = 1;
    double a = px1 + u * px, b = py1 + u * py;
    return SQ(xpos - a) + SQ(ypos - b) < SQ(radius);
}
/* physics.py:210-235 with clamp_half_cell 16-18 */
static int is_empty_row(OSim *S, int x1, int x2, int y, int dir) {
    for (int x = x1; x <= x2; x++) {
        int yy = dir == 1 ? y + 1 : y;
        if (S->hor_edge[clampi(x, 0, 87)][clampi(yy, 0, 49)]) return 0;
    }
    return 1;
}
static int is_empty_column(OSim *S, int x, int y1, int y2, int dir) {
    for (int y = y1; y <= y2; y++) {
        int xx = dir == 1 ? x + 1 : x;
        if (S->ver_edge[clampi(xx, 0, 87)][clampi(y, 0, 49)]) return 0;
    }
    return 1;
}
/* physics.py:317-332; returns 0 on KeyError */
static int orientation_to_vector(double o, double *vx, double *vy) {
    double diag = sqrt(2.0) / 2;
    if (o == 0) { *vx = 1; *vy = 0; } else if (o == 1) { *vx = diag; *vy = diag; }
    else if (o == 2) { *vx = 0; *vy = 1; } else if (o == 3) { *vx = -diag; *vy = diag; }
    else if (o == 4) { *vx = -1; *vy = 0; } else if (o == 5) { *vx = -diag; *vy = -diag; }
    else if (o == 6) { *vx = 0; *vy = -1; } else if (o == 7) { *vx = diag; *vy = -diag; }
    else return 0;
    return 1;
}

/* ------------------------------------------------------------------------- */
/* ninja: ninja.py                                                             */
/* ------------------------------------------------------------------------- */
static int ninja_valid_target(const Ninja *n) { return !(n->state == 6 || n->state == 8 || n->state == 9); } /* ninja.py:1272-1274 */
static void ninja_win(Ninja *n) {                    /* ninja.py:1246-1251 */
    if (n->state < 6) { if (n->state == 3) n->applied_gravity = GRAVITY_FALL; n->state = 8; }
}
static void ninja_kill(Ninja *n, double x, double y, double xs, double ys, int cause) {   /* ninja.py:1253-1270 */
    if (n->state < 6) {
        n->death_xpos = x; n->death_ypos = y; n->death_xspeed = xs; n->death_yspeed = ys;
        n->death_cause = cause;
        if (n->state == 3) n->applied_gravity = GRAVITY_FALL;
        n->state = 7;
    }
}
static void ninja_init(OSim *S, int fresh) {         /* ninja.py:80-196 / reset_state 1288-1394 */
    Ninja *n = &S->ninja;
    IVec keep = n->cached_entities;
    double ix = n->initial_xpos, iy = n->initial_ypos;
    memset(n, 0, sizeof(*n));
    n->cached_entities = keep; n->cached_entities.n = 0;
    if (fresh) { ix = S->map_data[1231] * 6; iy = S->map_data[1232] * 6; }
    n->initial_xpos = ix; n->initial_ypos = iy;
    n->xpos = ix; n->ypos = iy;
    n->applied_gravity = GRAVITY_FALL; n->applied_drag = DRAG_REGULAR; n->applied_friction = FRICTION_GROUND;
    n->jump_buffer = n->floor_buffer = n->wall_buffer = n->launch_pad_buffer = -1;
    n->floor_normalized_y = -1; n->ceiling_normalized_y = 1;
    n->xpos_at_action = n->xpos; n->ypos_at_action = n->ypos;
    S->have_ninja = 1;
}

/* forward decls for entity callbacks */
static int ent_physical_collision(OSim *S, int id, double *nx, double *ny, double *len, double *other);
/* logical_collision result: 0 falsy, 1 launch pad tuple (bx,by), 2 scalar (val) */
static int ent_logical_collision(OSim *S, int id, double *bx, double *by, double *val);

static void ninja_integrate(Ninja *n) {              /* ninja.py:198-206 */
    n->xspeed *= n->applied_drag;
    n->yspeed *= n->applied_drag;
    n->yspeed += n->applied_gravity;
    n->xpos_old = n->xpos; n->ypos_old = n->ypos;
    n->xpos += n->xspeed; n->ypos += n->yspeed;
}
static void ninja_pre_collision(OSim *S) {           /* ninja.py:208-222 */
    Ninja *n = &S->ninja;
    n->xspeed_old = n->xspeed; n->yspeed_old = n->yspeed;
    n->floor_count = n->wall_count = n->ceiling_count = 0;
    n->floor_normal_x = n->floor_normal_y = 0;
    n->ceiling_normal_x = n->ceiling_normal_y = 0;
    n->is_crushable = 0;
    n->x_crush = n->y_crush = 0; n->crush_len = 0;
    gather_entities(S, n->xpos, n->ypos, &n->cached_entities);
}
static void ninja_collide_vs_objects(OSim *S) {      /* ninja.py:224-267 */
    Ninja *n = &S->ninja;
    for (int k = 0; k < n->cached_entities.n; k++) {
        int id = n->cached_entities.v[k];
        Ent *e = &S->ents[id];
        if (!e->physical) continue;
        double depen_x, depen_y, depen_len, other;
        if (!ent_physical_collision(S, id, &depen_x, &depen_y, &depen_len, &other)) continue;
        double pop_x = depen_x * depen_len, pop_y = depen_y * depen_len;
        n->xpos += pop_x; n->ypos += pop_y;
        if (e->type != 17) { n->x_crush += pop_x; n->y_crush += pop_y; n->crush_len += depen_len; }
        if (e->type == 20) n->is_crushable = 1;
        if (e->type == 17 || e->type == 20 || e->type == 28) { n->xspeed += pop_x; n->yspeed += pop_y; }
        if (e->type == 11) {
            double xspeed_new = (n->xspeed * depen_y - n->yspeed * depen_x) * depen_y;
            /* axis-aligned normals are Python ints (physics.py:322-331): -(int 0) is 0, not -0.0 */
            double yspeed_new = (n->xspeed * depen_y - n->yspeed * depen_x) * (depen_x == 0 ? 0.0 : -depen_x);
            n->xspeed = xspeed_new; n->yspeed = yspeed_new;
        }
        if (depen_y >= -0.0001) { n->ceiling_count += 1; n->ceiling_normal_x += depen_x; n->ceiling_normal_y += depen_y; }
        else { n->floor_count += 1; n->floor_normal_x += depen_x; n->floor_normal_y += depen_y; }
    }
}
#ifdef OSIM_STATS   /* tools/iter_stats.py only: depenetration-loop statistics (not thread safe) */
long osim_stat_iters[34], osim_stat_nseg[64], osim_stat_calls;
long osim_stat_period[5][34];
long osim_stat_traj[34];        /* [k]: full-length loops whose position at call k first matched the previous full-length loop's (same env, same call index); 33: none */
static double osim_prev_x[32], osim_prev_y[32]; static int osim_prev_valid;   /* one env at a time (tools/iter_stats.py) */
void osim_stat_new_env(void) { osim_prev_valid = 0; }   /* [p][k]: full-length (32-call) loops whose position first repeated with period p at call k; p = 0: never */
#define STAT_ITERS(k, ns) do { osim_stat_iters[(k)]++; osim_stat_nseg[(ns) < 63 ? (ns) : 63]++; osim_stat_calls += (k); } while (0)
#else
#define STAT_ITERS(k, ns) do { } while (0)
#endif
static void ninja_collide_vs_tiles(OSim *S) {        /* ninja.py:269-379 */
    Ninja *n = &S->ninja;
    double dx = n->xpos - n->xpos_old, dy = n->ypos - n->ypos_old;
    double time = sweep_circle_vs_tiles(S, n->xpos_old, n->ypos_old, dx, dy, NINJA_RADIUS * 0.5);
    n->xpos = n->xpos_old + time * dx;
    n->ypos = n->ypos_old + time * dy;
    static const double rad = NINJA_RADIUS;
    IVec segments = {0, 0, 0};
    query_region(S, n->xpos - rad, n->ypos - rad, n->xpos + rad, n->ypos + rad, &S->tmp_segs);
    /* private copy: nothing else queries inside the loop, but keep the reference's "gather once" shape */
    for (int k = 0; k < S->tmp_segs.n; k++) iv_push(&segments, S->tmp_segs.v[k]);
    double xpos = n->xpos, ypos = n->ypos, xspeed = n->xspeed, yspeed = n->yspeed;
    int cp_calls = 0;
#ifdef OSIM_STATS
    double hx[32], hy[32]; int per_p = 0, per_k = 0;
#endif
    for (int it = 0; it < 32; it++) {
        double a, b;
#ifdef OSIM_STATS
        hx[it] = xpos; hy[it] = ypos;
        static int w1, w2;
        for (int p = 1; p <= 4 && !per_p; p++)
            if (it >= p && memcmp(&hx[it - p], &xpos, 8) == 0 && memcmp(&hy[it - p], &ypos, 8) == 0) { per_p = p; per_k = it; }
#endif
        int result = get_single_closest_point(S, xpos, ypos, NINJA_RADIUS, &segments, &a, &b);
        cp_calls++;
#ifdef OSIM_STATS
        if (it >= 2) {
            int any = 0, inb = 0;
            for (int k = 0; k < segments.n; k++) { const Seg *sg = &S->segs[segments.v[k]]; if (sg->kind) { any = 1;
                if (!(sg->b[2] < xpos - 10 || sg->b[0] > xpos + 10 || sg->b[3] < ypos - 10 || sg->b[1] > ypos + 10)) inb = 1; } }
            osim_stat_arc[0] += any; osim_stat_arc[1] += inb;
            if (osim_last_winner >= 0 && S->segs[segments.v[osim_last_winner]].kind) osim_stat_arc[2]++;
        }
        if (it >= 2) { osim_stat_pred[0]++; if (osim_last_winner == w2) osim_stat_pred[1]++; else if (osim_last_winner == w1) osim_stat_pred[2]++; else osim_stat_pred[3]++; }
        w2 = w1; w1 = osim_last_winner;
#endif
        if (result == 0) break;
        dx = xpos - a; dy = ypos - b;
        if (fabs(dx) <= 0.0000001) {                 /* ninja.py:312-318 band-aid */
            dx = 0;
            if (xpos == 50.51197510492316 || xpos == 49.23232124849253) dx = -ldexp(1.0, -47);
            if (xpos == 49.153536108584795) dx = ldexp(1.0, -47);
        }
        double dx_sq = dx * dx, dy_sq = dy * dy;
        double dist_sq = dx_sq + dy_sq;
        if (dist_sq < 1e-16) break;
        double dist = sqrt(dist_sq);
        double depen_len = NINJA_RADIUS - dist * result;
        if (depen_len < 0.0000001) break;
        double inv_dist = 1.0 / dist;
        double norm_dx = dx * inv_dist, norm_dy = dy * inv_dist;
        double depen_x = norm_dx * depen_len, depen_y = norm_dy * depen_len;
        xpos += depen_x; ypos += depen_y;
        n->x_crush += depen_x; n->y_crush += depen_y; n->crush_len += depen_len;
        double dot_product = xspeed * dx + yspeed * dy;
        if (dot_product < 0) {
            double cross_product = xspeed * dy - yspeed * dx;
            double inv_dist_sq = inv_dist * inv_dist;
            xspeed = cross_product * inv_dist_sq * dy;
            /* dx == 0 only via the band-aid's `dx = 0` (a Python int): -(int 0) is 0, not -0.0 */
            yspeed = cross_product * inv_dist_sq * (dx == 0 ? 0.0 : -dx);
        }
        if (dy >= -0.0001) { n->ceiling_count += 1; n->ceiling_normal_x += norm_dx; n->ceiling_normal_y += norm_dy; }
        else { n->floor_count += 1; n->floor_normal_x += norm_dx; n->floor_normal_y += norm_dy; }
    }
    STAT_ITERS(cp_calls, segments.n);
#ifdef OSIM_STATS
    if (cp_calls == 32) {
        osim_stat_period[per_p][per_k]++;
        int k = 33;
        if (osim_prev_valid) for (int j = 0; j < 32; j++) if (memcmp(&hx[j], &osim_prev_x[j], 8) == 0 && memcmp(&hy[j], &osim_prev_y[j], 8) == 0) { k = j; break; }
        osim_stat_traj[k]++;
        memcpy(osim_prev_x, hx, sizeof hx); memcpy(osim_prev_y, hy, sizeof hy); osim_prev_valid = 1;
    }
#endif
    (void)cp_calls;
    n->xpos = xpos; n->ypos = ypos; n->xspeed = xspeed; n->yspeed = yspeed;
    free(segments.v);
}
static void ninja_post_collision(OSim *S) {          /* ninja.py:381-537 */
    Ninja *n = &S->ninja;
    double wall_normal = 0;
    IVec ents = {0, 0, 0};
    gather_entities(S, n->xpos, n->ypos, &S->tmp_ents);
    for (int k = 0; k < S->tmp_ents.n; k++) iv_push(&ents, S->tmp_ents.v[k]);
    for (int k = 0; k < ents.n; k++) {
        int id = ents.v[k];
        Ent *e = &S->ents[id];
        if (!e->logical) continue;
        double bx, by, val;
        int r = ent_logical_collision(S, id, &bx, &by, &val);
        if (r == 1) {                                 /* launch pad tuple, ninja.py:401-418 */
            double xboost = bx * 2 / 3, yboost = by * 2 / 3;
            n->xpos += xboost; n->ypos += yboost;
            n->xspeed = xboost; n->yspeed = yboost;
            n->floor_count = 0; n->floor_buffer = -1;
            double boost_scalar = sqrt(SQ(xboost) + SQ(yboost));
            n->xlp_boost_normalized = xboost / boost_scalar;
            n->ylp_boost_normalized = yboost / boost_scalar;
            n->launch_pad_buffer = 0;
            if (n->state == 3) n->applied_gravity = GRAVITY_FALL;
            n->state = 4;
        } else if (r == 2) {
            wall_normal += val;
        }
    }
    free(ents.v);
    double rad = NINJA_RADIUS + 0.1;
    query_region(S, n->xpos - rad, n->ypos - rad, n->xpos + rad, n->ypos + rad, &S->tmp_segs);
    for (int k = 0; k < S->tmp_segs.n; k++) {
        int back; double a, b;
        seg_closest_point(&S->segs[S->tmp_segs.v[k]], n->xpos, n->ypos, &back, &a, &b);
        double dx = n->xpos - a, dy = n->ypos - b;
        double dist = sqrt(SQ(dx) + SQ(dy));
        if (fabs(dy) < 0.00001 && 0 < dist && dist <= rad) wall_normal += dx / dist;
    }
    n->airborn_old = n->airborn;
    n->airborn = 1; n->walled = 0;
    if (wall_normal != 0) { n->walled = 1; n->wall_normal = wall_normal / fabs(wall_normal); }
    if (n->floor_count > 0) {
        n->airborn = 0;
        double floor_scalar = sqrt(SQ(n->floor_normal_x) + SQ(n->floor_normal_y));
        if (floor_scalar == 0) { n->floor_normalized_x = 0; n->floor_normalized_y = -1; }
        else { n->floor_normalized_x = n->floor_normal_x / floor_scalar; n->floor_normalized_y = n->floor_normal_y / floor_scalar; }
        if (n->state != 8 && n->airborn_old) {
            double impact_vel = -(n->floor_normalized_x * n->xspeed_old + n->floor_normalized_y * n->yspeed_old);
            if (impact_vel > MAX_SURVIVABLE_IMPACT - 4.0 / 3 * fabs(n->floor_normalized_y)) {
                n->xspeed = n->xspeed_old; n->yspeed = n->yspeed_old;
                ninja_kill(n, n->xpos, n->ypos, n->xspeed * 0.5, n->yspeed * 0.5, 2);
                n->terminal_impact = 1;
            }
        }
    }
    if (n->airborn) n->frames_airborne += 1; else n->frames_airborne = 0;
    if (!n->airborn) n->consecutive_floor_frames += 1; else n->consecutive_floor_frames = 0;
    if (n->walled) n->consecutive_wall_frames += 1; else n->consecutive_wall_frames = 0;
    if (n->ceiling_count > 0) {
        double ceiling_scalar = sqrt(SQ(n->ceiling_normal_x) + SQ(n->ceiling_normal_y));
        if (ceiling_scalar == 0) { n->ceiling_normalized_x = 0; n->ceiling_normalized_y = 1; }
        else { n->ceiling_normalized_x = n->ceiling_normal_x / ceiling_scalar; n->ceiling_normalized_y = n->ceiling_normal_y / ceiling_scalar; }
        if (n->state != 8) {
            double impact_vel = -(n->ceiling_normalized_x * n->xspeed_old + n->ceiling_normalized_y * n->yspeed_old);
            if (impact_vel > MAX_SURVIVABLE_IMPACT - 4.0 / 3 * fabs(n->ceiling_normalized_y)) {
                n->xspeed = n->xspeed_old; n->yspeed = n->yspeed_old;
                ninja_kill(n, n->xpos, n->ypos, n->xspeed * 0.5, n->yspeed * 0.5, 2);
                n->terminal_impact = 1;
            }
        }
    }
    if (n->is_crushable && n->crush_len > 0) {
        if (sqrt(SQ(n->x_crush) + SQ(n->y_crush)) / n->crush_len < MIN_SURVIVABLE_CRUSHING)
            ninja_kill(n, n->xpos, n->ypos, 0, 0, 0);
    }
}
static void ninja_floor_jump(Ninja *n, int was_buffered) {   /* ninja.py:539-579 */
    n->jump_buffer = -1; n->floor_buffer = -1; n->launch_pad_buffer = -1;
    if (was_buffered) n->last_jump_was_buffered = 1;
    n->state = 3;
    n->applied_gravity = GRAVITY_JUMP;
    double jx, jy;
    if (n->floor_normalized_x == 0) { jx = 0; jy = -2; }
    else {
        double dx = n->floor_normalized_x, dy = n->floor_normalized_y;
        if (n->xspeed * dx >= 0) {
            if (n->xspeed * n->hor_input >= 0) { jx = 2.0 / 3 * dx; jy = 2 * dy; }
            else { jx = 0; jy = -1.4; }
        } else {
            if (n->xspeed * n->hor_input > 0) { jx = 0; jy = -1.4; }
            else { n->xspeed = 0; jx = 2.0 / 3 * dx; jy = 2 * dy; }
        }
    }
    if (n->yspeed > 0) n->yspeed = 0;
    n->xspeed += jx; n->yspeed += jy;
    n->xpos += jx; n->ypos += jy;
    n->jump_duration = 0;
}
static void ninja_wall_jump(Ninja *n, int was_buffered) {    /* ninja.py:581-608 */
    double jx, jy;
    if (n->hor_input * n->wall_normal < 0 && n->state == 5) { jx = 2.0 / 3; jy = -1; }
    else { jx = 1; jy = -1.4; }
    if (was_buffered) n->last_jump_was_buffered = 1;
    n->state = 3;
    n->applied_gravity = GRAVITY_JUMP;
    if (n->xspeed * n->wall_normal < 0) n->xspeed = 0;
    if (n->yspeed > 0) n->yspeed = 0;
    n->xspeed += jx * n->wall_normal; n->yspeed += jy;
    n->xpos += jx * n->wall_normal; n->ypos += jy;
    n->jump_buffer = -1; n->wall_buffer = -1; n->launch_pad_buffer = -1;
    n->jump_duration = 0;
}
static void ninja_lp_jump(Ninja *n) {                        /* ninja.py:610-626 */
    n->floor_buffer = -1; n->wall_buffer = -1; n->jump_buffer = -1; n->launch_pad_buffer = -1;
    double boost_scalar = 2 * fabs(n->xlp_boost_normalized) + 2;
    if (boost_scalar == 2) boost_scalar = 1.7;
    n->xspeed += n->xlp_boost_normalized * boost_scalar * (2.0 / 3);
    n->yspeed += n->ylp_boost_normalized * boost_scalar * (2.0 / 3);
}
static void ninja_think(Ninja *n) {                          /* ninja.py:849-1059 */
    n->last_jump_was_buffered = 0;
    if (n->state != n->previous_state) { n->state_change_frame = 0; n->previous_state = n->state; }
    else n->state_change_frame += 1;
    int new_jump_check = n->jump_input ? (n->jump_input_old == 0) : 0;
    n->jump_input_old = n->jump_input;
    if (-1 < n->launch_pad_buffer && n->launch_pad_buffer < 3) n->launch_pad_buffer += 1; else n->launch_pad_buffer = -1;
    int in_lp_buffer = -1 < n->launch_pad_buffer && n->launch_pad_buffer < 4;
    if (-1 < n->jump_buffer && n->jump_buffer < 5) n->jump_buffer += 1; else n->jump_buffer = -1;
    int in_jump_buffer = -1 < n->jump_buffer && n->jump_buffer < 5;
    if (-1 < n->wall_buffer && n->wall_buffer < 5) n->wall_buffer += 1; else n->wall_buffer = -1;
    int in_wall_buffer = -1 < n->wall_buffer && n->wall_buffer < 5;
    if (-1 < n->floor_buffer && n->floor_buffer < 5) n->floor_buffer += 1; else n->floor_buffer = -1;
    int in_floor_buffer = -1 < n->floor_buffer && n->floor_buffer < 5;
    int just = 0;
    if (new_jump_check && n->airborn) { n->jump_buffer = 0; just = 1; }
    if (n->walled) n->wall_buffer = 0;
    if (!n->airborn) n->floor_buffer = 0;
    if (n->jump_input) {
        int has_active;
        if (just) has_active = 1;
        else {
            int cur = -1 < n->jump_buffer && n->jump_buffer < 5;
            has_active = cur || in_floor_buffer || in_wall_buffer || in_lp_buffer;
        }
        n->is_jump_useless = n->airborn && n->state != 3 && !has_active;
    } else n->is_jump_useless = 0;

    if (n->state == 6 || n->state == 9) return;
    if (n->state == 7) { n->state = 6; return; }            /* think_awaiting_death 1061-1063 */
    if (n->state == 8) { n->applied_drag = n->airborn ? DRAG_REGULAR : DRAG_SLOW; return; }

    if (!n->airborn) {
        double xspeed_new = n->xspeed + GROUND_ACCEL * n->hor_input;
        if (fabs(xspeed_new) < MAX_HOR_SPEED) n->xspeed = xspeed_new;
        if (n->state > 2) {
            if (n->xspeed * n->hor_input <= 0) { if (n->state == 3) n->applied_gravity = GRAVITY_FALL; n->state = 2; }
            else { if (n->state == 3) n->applied_gravity = GRAVITY_FALL; n->state = 1; }
        }
        if (!in_jump_buffer && !new_jump_check) {
            if (n->state == 2) {
                double projection = fabs(n->yspeed * n->floor_normalized_x - n->xspeed * n->floor_normalized_y);
                if (n->hor_input * projection * n->xspeed > 0) { n->state = 1; return; }
                if (projection < 0.1 && n->floor_normalized_x == 0) { n->state = 0; return; }
                if (n->yspeed < 0 && n->floor_normalized_x != 0) {
                    double speed_scalar = sqrt(SQ(n->xspeed) + SQ(n->yspeed));
                    double fric_force = fabs(n->xspeed * (1 - FRICTION_GROUND) * n->floor_normalized_y);
                    double fric_force2 = speed_scalar - fric_force * SQ(n->floor_normalized_y);
                    n->xspeed = n->xspeed / speed_scalar * fric_force2;
                    n->yspeed = n->yspeed / speed_scalar * fric_force2;
                    return;
                }
                n->xspeed *= FRICTION_GROUND;
                return;
            }
            if (n->state == 1) {
                double projection = fabs(n->yspeed * n->floor_normalized_x - n->xspeed * n->floor_normalized_y);
                if (n->hor_input * projection * n->xspeed > 0) {
                    if (n->hor_input * n->floor_normalized_x >= 0) return;
                    if (fabs(xspeed_new) < MAX_HOR_SPEED) {
                        double boost = GROUND_ACCEL / 2 * n->hor_input;
                        double xboost = boost * n->floor_normalized_y * n->floor_normalized_y;
                        double yboost = boost * n->floor_normalized_y * -n->floor_normalized_x;
                        n->xspeed += xboost; n->yspeed += yboost;
                    }
                    return;
                }
                n->state = 2;
            } else {
                if (n->hor_input) { n->state = 1; return; }
                double projection = fabs(n->yspeed * n->floor_normalized_x - n->xspeed * n->floor_normalized_y);
                if (projection < 0.1) { n->xspeed *= FRICTION_GROUND_SLOW; return; }
                n->state = 2;
            }
            return;
        }
        ninja_floor_jump(n, 0);
        return;
    } else {
        double xspeed_new = n->xspeed + AIR_ACCEL * n->hor_input;
        if (fabs(xspeed_new) < MAX_HOR_SPEED) n->xspeed = xspeed_new;
        if (n->state < 3) { n->state = 4; return; }
        if (n->state == 3) {
            n->jump_duration += 1;
            if (!n->jump_input || n->jump_duration > MAX_JUMP_DURATION) { n->applied_gravity = GRAVITY_FALL; n->state = 4; return; }
        }
        if (in_jump_buffer || new_jump_check) {
            if (n->walled || in_wall_buffer) { ninja_wall_jump(n, in_wall_buffer); return; }
            if (in_floor_buffer) { ninja_floor_jump(n, 1); return; }
            if (in_lp_buffer && new_jump_check) { ninja_lp_jump(n); return; }
        }
        if (!n->walled) { if (n->state == 5) n->state = 4; }
        else {
            if (n->state == 5) {
                if (n->hor_input * n->wall_normal <= 0) n->yspeed *= FRICTION_WALL;
                else n->state = 4;
            } else {
                if (n->yspeed > 0 && n->hor_input * n->wall_normal < 0) {
                    if (n->state == 3) n->applied_gravity = GRAVITY_FALL;
                    n->state = 5;
                }
            }
        }
    }
}

/* ------------------------------------------------------------------------- */
/* entities: entities.py + entity_classes/                                     */
/* ------------------------------------------------------------------------- */
static void grid_move(OSim *S, int id) {                     /* entities.py:256-264 */
    Ent *e = &S->ents[id];
    int cx = clampi(ifloor(e->xpos / 24), 0, 43), cy = clampi(ifloor(e->ypos / 24), 0, 24);
    if (cx != e->cellx || cy != e->celly) {
        iv_remove(&S->grid[e->cellx][e->celly], id);
        e->cellx = cx; e->celly = cy;
        iv_push(&S->grid[cx][cy], id);
    }
}
static void mine_set_state(Ent *e, int state) {              /* entity_toggle_mine.py:130-135 */
    static const double RADII[3] = {4.0, 3.5, 4.5};
    if (state >= 0 && state <= 2) { e->state = state; e->radius = RADII[state]; }
}
static void door_change_state(OSim *S, Ent *e, int closed) { /* entity_door_base.py:99-108 */
    e->closed = closed;
    for (int k = 0; k < 2; k++) {
        if (e->is_vertical) S->ver_edge[e->edge[k][0]][e->edge[k][1]] += closed ? 1 : -1;
        else S->hor_edge[e->edge[k][0]][e->edge[k][1]] += closed ? 1 : -1;
    }
}
/* drones: entity_drone_base.py:66-164 */
static const int DIR_TO_VEC[4][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}};
static const int DIR_LIST[4][4] = {{1, 0, 3, 2}, {3, 0, 1, 2}, {0, 1, 3, 2}, {0, 3, 1, 2}};
static int drone_test_dir(OSim *S, Ent *e, int dir) {
    int xdir = DIR_TO_VEC[dir][0], ydir = DIR_TO_VEC[dir][1];
    double xtarget = e->xpos + e->grid_width * xdir;
    double ytarget = e->ypos + e->grid_width * ydir;
    if (!ydir) {
        int cell_x = ifloor((e->xpos + xdir * e->radius) / 12);
        int cell_xtarget = ifloor((xtarget + xdir * e->radius) / 12);
        int cell_y1 = ifloor((e->ypos - e->radius) / 12);
        int cell_y2 = ifloor((e->ypos + e->radius) / 12);
        while (cell_x != cell_xtarget) {
            if (!is_empty_column(S, cell_x, cell_y1, cell_y2, xdir)) return 0;
            cell_x += xdir;
        }
    } else {
        int cell_y = ifloor((e->ypos + ydir * e->radius) / 12);
        int cell_ytarget = ifloor((ytarget + ydir * e->radius) / 12);
        int cell_x1 = ifloor((e->xpos - e->radius) / 12);
        int cell_x2 = ifloor((e->xpos + e->radius) / 12);
        while (cell_y != cell_ytarget) {
            if (!is_empty_row(S, cell_x1, cell_x2, cell_y, ydir)) return 0;
            cell_y += ydir;
        }
    }
    e->xtarget = xtarget; e->ytarget = ytarget;
    return 1;
}
static void drone_move(OSim *S, int id) {
    Ent *e = &S->ents[id];
    double xspeed = e->speed * DIR_TO_VEC[e->dir][0];
    double yspeed = e->speed * DIR_TO_VEC[e->dir][1];
    double dx = e->xtarget - e->xpos, dy = e->ytarget - e->ypos;
    double dist = sqrt(SQ(dx) + SQ(dy));
    if (dist < 0.000001 || (dx * (e->xtarget - (e->xpos + xspeed)) + dy * (e->ytarget - (e->ypos + yspeed))) < 0) {
        e->xpos = e->xtarget; e->ypos = e->ytarget;
        int can_move = 0;
        for (int i = 0; i < 4; i++) {
            int new_dir = (e->dir + DIR_LIST[e->mode][i]) % 4;
            if (drone_test_dir(S, e, new_dir)) { e->dir = new_dir; can_move = 1; break; }
        }
        if (can_move) {
            double disp = e->speed - dist;
            e->xpos += disp * DIR_TO_VEC[e->dir][0];
            e->ypos += disp * DIR_TO_VEC[e->dir][1];
        }
    } else {
        e->xpos += xspeed; e->ypos += yspeed;
        grid_move(S, id);
    }
}
/* thwump: entity_thwump.py:109-206 */
static void thwump_move(OSim *S, int id) {
    Ent *e = &S->ents[id];
    if (!e->state) return;
    double speed = e->state == 1 ? 20.0 / 7 : 8.0 / 7;
    int speed_dir = e->direction * e->state;
    if (!e->is_horizontal) {
        double ypos_new = e->ypos + speed * speed_dir;
        if (e->state == -1 && (ypos_new - e->yorigin) * (e->ypos - e->yorigin) < 0) { e->ypos = e->yorigin; e->state = 0; return; }
        int cell_y = ifloor((e->ypos + speed_dir * 11) / 12);
        int cell_y_new = ifloor((ypos_new + speed_dir * 11) / 12);
        if (cell_y != cell_y_new) {
            int cell_x1 = ifloor((e->xpos - 11) / 12), cell_x2 = ifloor((e->xpos + 11) / 12);
            if (!is_empty_row(S, cell_x1, cell_x2, cell_y, speed_dir)) { e->state = -1; return; }
        }
        e->ypos = ypos_new;
    } else {
        double xpos_new = e->xpos + speed * speed_dir;
        if (e->state == -1 && (xpos_new - e->xorigin) * (e->xpos - e->xorigin) < 0) { e->xpos = e->xorigin; e->state = 0; return; }
        int cell_x = ifloor((e->xpos + speed_dir * 11) / 12);
        int cell_x_new = ifloor((xpos_new + speed_dir * 11) / 12);
        if (cell_x != cell_x_new) {
            int cell_y1 = ifloor((e->ypos - 11) / 12), cell_y2 = ifloor((e->ypos + 11) / 12);
            if (!is_empty_column(S, cell_x, cell_y1, cell_y2, speed_dir)) { e->state = -1; return; }
        }
        e->xpos = xpos_new;
    }
    grid_move(S, id);
}
static void thwump_think(OSim *S, int id) {
    Ent *e = &S->ents[id];
    Ninja *n = &S->ninja;
    if (e->state || !ninja_valid_target(n)) return;
    double activation_range = 2 * (9 + NINJA_RADIUS);
    if (!e->is_horizontal) {
        if (fabs(e->xpos - n->xpos) < activation_range) {
            int ninja_ycell = ifloor(n->ypos / 12);
            int thwump_ycell = ifloor((e->ypos - e->direction * 11) / 12);
            int x1 = ifloor((e->xpos - 11) / 12), x2 = ifloor((e->xpos + 11) / 12);
            int dy = ninja_ycell - thwump_ycell;
            if (dy * e->direction >= 0) {
                int i;
                for (i = 0; i < 100; i++) {
                    if (!is_empty_row(S, x1, x2, thwump_ycell, e->direction)) { dy = ninja_ycell - thwump_ycell; break; }
                    thwump_ycell += e->direction;
                }
                if (i == 100) i = 99;                /* Python leaves i at the last value */
                if (i > 0 && dy * e->direction <= 0) e->state = 1;
            }
        }
    } else {
        if (fabs(e->ypos - n->ypos) < activation_range) {
            int ninja_xcell = ifloor(n->xpos / 12);
            int thwump_xcell = ifloor((e->xpos - e->direction * 11) / 12);
            int y1 = ifloor((e->ypos - 11) / 12), y2 = ifloor((e->ypos + 11) / 12);
            int dx = ninja_xcell - thwump_xcell;
            if (dx * e->direction >= 0) {
                int i;
                for (i = 0; i < 100; i++) {
                    if (!is_empty_column(S, thwump_xcell, y1, y2, e->direction)) { dx = ninja_xcell - thwump_xcell; break; }
                    thwump_xcell += e->direction;
                }
                if (i == 100) i = 99;
                if (i > 0 && dx * e->direction <= 0) e->state = 1;
            }
        }
    }
}
/* shove thwump: entity_shove_thwump.py:109-153 */
static void shwump_move_if_possible(OSim *S, int id, double xdir, double ydir, double speed) {
    Ent *e = &S->ents[id];
    if (e->ydir == 0) {
        double xpos_new = e->xpos + xdir * speed;
        int cell_x = ifloor(e->xpos / 12), cell_x_new = ifloor(xpos_new / 12);
        if (cell_x != cell_x_new) {
            int y1 = ifloor((e->ypos - 8) / 12), y2 = ifloor((e->ypos + 8) / 12);
            if (!is_empty_column(S, cell_x, y1, y2, (int)xdir)) { e->state = 3; return; }
        }
        e->xpos = xpos_new;
    } else {
        double ypos_new = e->ypos + ydir * speed;
        int cell_y = ifloor(e->ypos / 12), cell_y_new = ifloor(ypos_new / 12);
        if (cell_y != cell_y_new) {
            int x1 = ifloor((e->xpos - 8) / 12), x2 = ifloor((e->xpos + 8) / 12);
            if (!is_empty_row(S, x1, x2, cell_y, (int)ydir)) { e->state = 3; return; }
        }
        e->ypos = ypos_new;
    }
    grid_move(S, id);
}
static void shwump_think(OSim *S, int id) {
    Ent *e = &S->ents[id];
    if (e->state == 1) {
        if (e->activated) { e->activated = 0; return; }
        e->state = 2;
    }
    if (e->state == 3) {
        double origin_dist = fabs(e->xpos - e->xorigin) + fabs(e->ypos - e->yorigin);
        if (origin_dist >= 1) shwump_move_if_possible(S, id, e->xdir, e->ydir, 1);
        else { e->xpos = e->xorigin; e->ypos = e->yorigin; e->state = 0; }
    } else if (e->state == 2) {
        shwump_move_if_possible(S, id, -e->xdir, -e->ydir, 4);
    }
}
/* death ball: entity_death_ball.py:74-167 */
static void death_ball_think(OSim *S, int id) {
    Ent *e = &S->ents[id];
    Ninja *n = &S->ninja;
    if (!ninja_valid_target(n)) { e->xspeed *= 0.95; e->yspeed *= 0.95; }
    else {
        double dx = n->xpos - e->xpos, dy = n->ypos - e->ypos;
        double dist = sqrt(SQ(dx) + SQ(dy));
        if (dist > 0) { dx /= dist; dy /= dist; }
        e->xspeed += dx * 0.04; e->yspeed += dy * 0.04;
        double speed = sqrt(SQ(e->xspeed) + SQ(e->yspeed));
        if (speed > 0.85) {
            double new_speed = (speed - 0.85) * 0.9;
            if (new_speed <= 0.01) new_speed = 0;
            new_speed += 0.85;
            e->xspeed = e->xspeed / speed * new_speed;
            e->yspeed = e->yspeed / speed * new_speed;
        }
    }
    double xpos_old = e->xpos, ypos_old = e->ypos;
    e->xpos += e->xspeed; e->ypos += e->yspeed;
    double time = sweep_circle_vs_tiles(S, xpos_old, ypos_old, e->xspeed, e->yspeed, 8 * 0.5);
    e->xpos = xpos_old + time * e->xspeed;
    e->ypos = ypos_old + time * e->yspeed;
    double xnormal = 0, ynormal = 0;
    for (int it = 0; it < 16; it++) {
        double a, b;
        int result = get_single_closest_point(S, e->xpos, e->ypos, 8, NULL, &a, &b);
        if (result == 0) break;
        double dx = e->xpos - a, dy = e->ypos - b;
        double dist = sqrt(SQ(dx) + SQ(dy));
        double depen_len = 8 - dist * result;
        if (depen_len < 0.0000001) break;
        if (dist == 0) return;
        double xnorm = dx / dist, ynorm = dy / dist;
        e->xpos += xnorm * depen_len; e->ypos += ynorm * depen_len;
        xnormal += xnorm; ynormal += ynorm;
    }
    double normal_len = sqrt(SQ(xnormal) + SQ(ynormal));
    if (normal_len > 0) {
        double dx = xnormal / normal_len, dy = ynormal / normal_len;
        double dot_product = e->xspeed * dx + e->yspeed * dy;
        if (dot_product < 0) {
            double speed = sqrt(SQ(e->xspeed) + SQ(e->yspeed));
            double bounce_strength = speed <= 1.35 ? 1 : 2;
            e->xspeed -= dx * dot_product * bounce_strength;
            e->yspeed -= dy * dot_product * bounce_strength;
        }
    }
    double db_count = S->map_data[1200];
    if (e->index + 1 < db_count) {
        /* entity_dic[25][index+1:] : later death balls in list order */
        int pos = 0;
        for (int k = 0; k < S->ndic; k++) {
            Ent *t = &S->ents[S->dic_order[k]];
            if (t->dic_type != e->dic_type) continue;
            if (pos++ < e->index + 1) continue;
            double dx = e->xpos - t->xpos, dy = e->ypos - t->ypos;
            double dist = sqrt(SQ(dx) + SQ(dy));
            if (dist < 16) {
                dx = dx / dist * 4; dy = dy / dist * 4;
                e->xspeed += dx; e->yspeed += dy;
                t->xspeed -= dx; t->yspeed -= dy;
            }
        }
    }
    grid_move(S, id);
}

static void ent_move(OSim *S, int id) {
    Ent *e = &S->ents[id];
    Ninja *n = &S->ninja;
    switch (e->kind) {
    case K_DRONE_ZAP: case K_MINI_DRONE: drone_move(S, id); break;
    case K_BOUNCE: {                                           /* entity_bounce_block.py:107-125 */
        e->xspeed *= 0.98; e->yspeed *= 0.98;
        e->xpos += e->xspeed; e->ypos += e->yspeed;
        double xforce = 0.02222222222222222 * (e->xorigin - e->xpos);
        double yforce = 0.02222222222222222 * (e->yorigin - e->ypos);
        e->xpos += xforce; e->ypos += yforce;
        e->xspeed += xforce; e->yspeed += yforce;
        grid_move(S, id);
        break; }
    case K_THWUMP: thwump_move(S, id); break;
    case K_BOOST: {                                            /* entity_boost_pad.py:46-64 */
        if (!ninja_valid_target(n)) { e->touching = 0; break; }
        if (overlap_circle_vs_circle(e->xpos, e->ypos, 6, n->xpos, n->ypos, NINJA_RADIUS)) {
            if (!e->touching) {
                double vel_norm = sqrt(SQ(n->xspeed) + SQ(n->yspeed));
                if (vel_norm > 0) {
                    double x_boost = 2 * n->xspeed / vel_norm, y_boost = 2 * n->yspeed / vel_norm;
                    n->xspeed += x_boost; n->yspeed += y_boost;
                }
                e->touching = 1;
            }
        } else e->touching = 0;
        break; }
    default: break;
    }
}
static void ent_think(OSim *S, int id) {
    Ent *e = &S->ents[id];
    Ninja *n = &S->ninja;
    switch (e->kind) {
    case K_MINE:                                               /* entity_toggle_mine.py:90-118 */
        if (ninja_valid_target(n)) {
            if (e->state == 1) {
                if (overlap_circle_vs_circle(e->xpos, e->ypos, e->radius, n->xpos, n->ypos, NINJA_RADIUS)) mine_set_state(e, 2);
            } else if (e->state == 2) {
                if (!overlap_circle_vs_circle(e->xpos, e->ypos, e->radius, n->xpos, n->ypos, NINJA_RADIUS)) mine_set_state(e, 0);
            }
        } else if (e->state == 2 && n->state == 6) mine_set_state(e, 1);
        break;
    case K_DOOR_REGULAR:                                       /* entity_door_regular.py:46-53 */
        if (!e->closed) { e->open_timer += 1; if (e->open_timer > 5) door_change_state(S, e, 1); }
        break;
    case K_THWUMP: thwump_think(S, id); break;
    case K_DEATH_BALL: death_ball_think(S, id); break;
    case K_SHWUMP: shwump_think(S, id); break;
    default: break;
    }
}
static int one_way_depen(const Ent *e, const Ninja *n, double *len) {   /* entity_one_way_platform.py:79-103 */
    double dx = n->xpos - e->xpos, dy = n->ypos - e->ypos;
    double lateral_dist = dy * e->normal_x - dx * e->normal_y;
    double direction = (n->yspeed * e->normal_x - n->xspeed * e->normal_y) * lateral_dist;
    double radius_scalar = direction < 0 ? 0.91 : 0.51;
    if (fabs(lateral_dist) < radius_scalar * NINJA_RADIUS + 12) {
        double normal_dist = dx * e->normal_x + dy * e->normal_y;
        if (0 < normal_dist && normal_dist <= NINJA_RADIUS) {
            double normal_proj = n->xspeed * e->normal_x + n->yspeed * e->normal_y;
            if (normal_proj <= 0) {
                double dx_old = n->xpos_old - e->xpos, dy_old = n->ypos_old - e->ypos;
                double normal_dist_old = dx_old * e->normal_x + dy_old * e->normal_y;
                if (NINJA_RADIUS - normal_dist_old <= 1.1) { *len = NINJA_RADIUS - normal_dist; return 1; }
            }
        }
    }
    return 0;
}
static int ent_physical_collision(OSim *S, int id, double *nx, double *ny, double *len, double *other) {
    Ent *e = &S->ents[id];
    Ninja *n = &S->ninja;
    switch (e->kind) {
    case K_BOUNCE: {                                           /* entity_bounce_block.py:127-145 */
        double dl, ot;
        if (!penetration_square_vs_point(e->xpos, e->ypos, n->xpos, n->ypos, 9 + NINJA_RADIUS, nx, ny, &dl, &ot)) return 0;
        e->xpos -= *nx * dl * (1 - 0.2); e->ypos -= *ny * dl * (1 - 0.2);
        e->xspeed -= *nx * dl * (1 - 0.2); e->yspeed -= *ny * dl * (1 - 0.2);
        *len = dl * 0.2; *other = ot;
        return 1; }
    case K_THWUMP:                                             /* entity_thwump.py:208-213 */
        return penetration_square_vs_point(e->xpos, e->ypos, n->xpos, n->ypos, 9 + NINJA_RADIUS, nx, ny, len, other);
    case K_ONE_WAY:
        if (one_way_depen(e, n, len)) { *nx = e->normal_x; *ny = e->normal_y; *other = 0; return 1; }
        return 0;
    case K_SHWUMP:                                             /* entity_shove_thwump.py:155-171 */
        if (e->state <= 1) {
            if (penetration_square_vs_point(e->xpos, e->ypos, n->xpos, n->ypos, 12 + NINJA_RADIUS, nx, ny, len, other)) {
                if (e->state == 0 || e->xdir * *nx + e->ydir * *ny >= 0.01) return 1;
            }
        }
        return 0;
    default: return 0;
    }
}
static int ent_logical_collision(OSim *S, int id, double *bx, double *by, double *val) {
    Ent *e = &S->ents[id];
    Ninja *n = &S->ninja;
    double nx, ny, pen, other;
    switch (e->kind) {
    case K_MINE:                                               /* entity_toggle_mine.py:120-128 */
        if (ninja_valid_target(n) && e->state == 0)
            if (overlap_circle_vs_circle(e->xpos, e->ypos, e->radius, n->xpos, n->ypos, NINJA_RADIUS)) {
                mine_set_state(e, 1);
                ninja_kill(n, 0, 0, 0, 0, 1);
            }
        return 0;
    case K_GOLD:                                               /* entity_gold.py:63-72 */
        if (n->state != 8)
            if (overlap_circle_vs_circle(e->xpos, e->ypos, 6, n->xpos, n->ypos, NINJA_RADIUS)) { n->gold_collected += 1; e->active = 0; }
        return 0;
    case K_EXIT:                                               /* entity_exit.py:66-74 */
        if (overlap_circle_vs_circle(e->xpos, e->ypos, 12, n->xpos, n->ypos, NINJA_RADIUS)) ninja_win(n);
        return 0;
    case K_EXIT_SWITCH:                                        /* entity_exit_switch.py:67-129 */
        if (overlap_circle_vs_circle(e->xpos, e->ypos, 6, n->xpos, n->ypos, NINJA_RADIUS)) {
            e->active = 0;
            Ent *p = &S->ents[e->parent];
            iv_push(&S->grid[p->cellx][p->celly], e->parent);
            p->switch_hit = 1;
            S->events |= 1;
        }
        return 0;
    case K_DOOR_REGULAR:                                       /* entity_door_regular.py:55-63 */
        if (overlap_circle_vs_circle(e->xpos, e->ypos, 10, n->xpos, n->ypos, NINJA_RADIUS)) { door_change_state(S, e, 0); e->open_timer = 0; }
        return 0;
    case K_DOOR_LOCKED:                                        /* entity_door_locked.py:54-67 */
        if (overlap_circle_vs_circle(e->xpos, e->ypos, 5, n->xpos, n->ypos, NINJA_RADIUS)) {
            n->doors_opened += 1; door_change_state(S, e, 0); e->active = 0; S->events |= 2;
        }
        return 0;
    case K_DOOR_TRAP:                                          /* entity_door_trap.py:60-67 */
        if (overlap_circle_vs_circle(e->xpos, e->ypos, 5, n->xpos, n->ypos, NINJA_RADIUS)) { door_change_state(S, e, 1); e->active = 0; }
        return 0;
    case K_LAUNCH_PAD:                                         /* entity_launch_pad.py:82-100 */
        if (ninja_valid_target(n))
            if (overlap_circle_vs_circle(e->xpos, e->ypos, 6, n->xpos, n->ypos, NINJA_RADIUS))
                if ((e->xpos - (n->xpos - NINJA_RADIUS * e->normal_x)) * e->normal_x
                    + (e->ypos - (n->ypos - NINJA_RADIUS * e->normal_y)) * e->normal_y >= -0.1) {
                    double yboost_scale = 1;
                    if (e->normal_y < 0) yboost_scale = 1 - e->normal_y;
                    *bx = e->normal_x * (36.0 / 7);
                    *by = e->normal_y * (36.0 / 7) * yboost_scale;
                    return 1;
                }
        return 0;
    case K_ONE_WAY: {                                          /* entity_one_way_platform.py:111-116 */
        double len;
        if (one_way_depen(e, n, &len) && fabs(e->normal_x) == 1) { *val = e->normal_x; return 2; }
        return 0; }
    case K_DRONE_ZAP: case K_MINI_DRONE:                       /* entity_drone_zap.py:57-64 */
        if (ninja_valid_target(n))
            if (overlap_circle_vs_circle(e->xpos, e->ypos, e->radius, n->xpos, n->ypos, NINJA_RADIUS)) ninja_kill(n, 0, 0, 0, 0, 0);
        return 0;
    case K_BOUNCE:                                             /* entity_bounce_block.py:147-158 */
        if (penetration_square_vs_point(e->xpos, e->ypos, n->xpos, n->ypos, 9 + NINJA_RADIUS + 0.1, &nx, &ny, &pen, &other)) {
            if (nx != 0) { *val = nx; return 2; }
        }
        return 0;
    case K_THWUMP:                                             /* entity_thwump.py:215-243 */
        if (ninja_valid_target(n))
            if (penetration_square_vs_point(e->xpos, e->ypos, n->xpos, n->ypos, 9 + NINJA_RADIUS + 0.1, &nx, &ny, &pen, &other)) {
                double px1, py1, px2, py2;
                if (e->is_horizontal) {
                    double dx = (9 + 2) * e->direction, dy = 9 - 2;
                    px1 = e->xpos + dx; py1 = e->ypos - dy; px2 = e->xpos + dx; py2 = e->ypos + dy;
                } else {
                    double dx = 9 - 2, dy = (9 + 2) * e->direction;
                    px1 = e->xpos - dx; py1 = e->ypos + dy; px2 = e->xpos + dx; py2 = e->ypos + dy;
                }
                if (overlap_circle_vs_segment(n->xpos, n->ypos, NINJA_RADIUS + 2, px1, py1, px2, py2)) ninja_kill(n, 0, 0, 0, 0, 0);
                if (nx != 0) { *val = nx; return 2; }
            }
        return 0;
    case K_DEATH_BALL:                                         /* entity_death_ball.py:169-181 */
        if (ninja_valid_target(n))
            if (overlap_circle_vs_circle(e->xpos, e->ypos, 5, n->xpos, n->ypos, NINJA_RADIUS)) {
                double dx = e->xpos - n->xpos, dy = e->ypos - n->ypos;
                double dist = sqrt(SQ(dx) + SQ(dy));
                e->xspeed += dx / dist * 10; e->yspeed += dy / dist * 10;
                ninja_kill(n, 0, 0, 0, 0, 0);
            }
        return 0;
    case K_SHWUMP: {                                           /* entity_shove_thwump.py:173-202 */
        int hit = penetration_square_vs_point(e->xpos, e->ypos, n->xpos, n->ypos, 12 + NINJA_RADIUS + 0.1, &nx, &ny, &pen, &other);
        if (hit && e->state <= 1) {
            if (e->state == 0) {
                e->activated = 1;
                if (other > 0.2) { e->xdir = nx; e->ydir = ny; e->state = 1; }
            } else if (e->state == 1) {
                if (e->xdir * nx + e->ydir * ny >= 0.01) e->activated = 1;
                else return 0;
            }
            if (nx != 0) { *val = nx; return 2; }
            return 0;
        }
        if (overlap_circle_vs_circle(n->xpos, n->ypos, NINJA_RADIUS, e->xpos, e->ypos, 8)) ninja_kill(n, 0, 0, 0, 0, 0);
        return 0; }
    default: return 0;
    }
}

/* ------------------------------------------------------------------------- */
/* load: nsim.py:51-76,142-219  map_loader.py  utils/tile_segment_factory.py   */
/* ------------------------------------------------------------------------- */
static Seg *new_seg(OSim *S) {
    if (S->nsegs == S->segcap) { S->segcap = S->segcap ? 2 * S->segcap : 1024; S->segs = (Seg *)realloc(S->segs, sizeof(Seg) * S->segcap); }
    Seg *s = &S->segs[S->nsegs++];
    memset(s, 0, sizeof(*s));
    return s;
}
static int add_linear(OSim *S, double x1, double y1, double x2, double y2) {   /* entities.py:12-41 */
    Seg *s = new_seg(S);
    s->kind = 0; s->x1 = x1; s->y1 = y1; s->x2 = x2; s->y2 = y2; s->oriented = 1;
    s->px = x2 - x1; s->py = y2 - y1;
    s->seg_lensq = s->px * s->px + s->py * s->py;         /* ints in the reference: exact */
    if (s->seg_lensq == 0) s->seg_lensq = 1e-9;
    s->b[0] = x1 < x2 ? x1 : x2; s->b[1] = y1 < y2 ? y1 : y2;
    s->b[2] = x1 > x2 ? x1 : x2; s->b[3] = y1 > y2 ? y1 : y2;
    return S->nsegs - 1;
}
static int add_circular(OSim *S, double cx, double cy, double hor, double ver, int convex) {  /* entities.py:102-125 */
    Seg *s = new_seg(S);
    s->kind = 1; s->cx = cx; s->cy = cy; s->hor = hor; s->ver = ver; s->radius = 24; s->convex = convex;
    s->phx = cx + 24 * hor; s->phy = cy; s->pvx = cx; s->pvy = cy + 24 * ver;
    s->b[0] = cx < s->phx ? cx : s->phx; s->b[2] = cx > s->phx ? cx : s->phx;
    s->b[1] = cy < s->pvy ? cy : s->pvy; s->b[3] = cy > s->pvy ? cy : s->pvy;
    return S->nsegs - 1;
}
static void process_tile(OSim *S, double tile_id, int x, int y) {   /* tile_segment_factory.py:264-312,169-215 */
    if (tile_id >= 34) return;
    if (!(tile_id >= 0 && tile_id == floor(tile_id))) return;    /* `tile_id in MAP` is False for non-integers / negatives */
    int t = (int)tile_id;
    for (int yl = 0; yl < 3; yl++) for (int xl = 0; xl < 2; xl++) {
        int *p = &S->hor_edge[2 * x + xl][2 * y + yl];
        *p = (*p + GRID_EDGE[t][2 * yl + xl]) % 2;
    }
    for (int xl = 0; xl < 3; xl++) for (int yl = 0; yl < 2; yl++) {
        int *p = &S->ver_edge[2 * x + xl][2 * y + yl];
        *p = (*p + GRID_EDGE[t][2 * xl + yl + 6]) % 2;
    }
    for (int yl = 0; yl < 3; yl++) for (int xl = 0; xl < 2; xl++)
        S->hor_segd[2 * x + xl][2 * y + yl] += SEG_ORTHO[t][2 * yl + xl];
    for (int xl = 0; xl < 3; xl++) for (int yl = 0; yl < 2; yl++)
        S->ver_segd[2 * x + xl][2 * y + yl] += SEG_ORTHO[t][2 * xl + yl + 6];
    int xtl = x * 24, ytl = y * 24;                               /* tile_segment_cache.py:35-62,98-112 */
    if (SEG_DIAG[t][0] >= 0)
        iv_push(&S->cells[x][y].segs, add_linear(S, xtl + SEG_DIAG[t][0], ytl + SEG_DIAG[t][1], xtl + SEG_DIAG[t][2], ytl + SEG_DIAG[t][3]));
    if (t >= 10 && t <= 17) {
        const signed char *c = SEG_CIRC[t - 10];
        iv_push(&S->cells[x][y].segs, add_circular(S, xtl + c[0], ytl + c[1], c[2], c[3], c[4]));
    }
}
static int load_tiles(OSim *S) {
    /* nsim.py:194-219 */
    S->nsegs = 0;
    for (int x = 0; x < NCX; x++) for (int y = 0; y < NCY; y++) { S->cells[x][y].segs.n = 0; S->cells[x][y].present = 0; }
    for (int x = 0; x < NHX; x++) for (int y = 0; y < NHY; y++) {
        S->hor_edge[x][y] = (y == 0 || y == 50) ? 1 : 0;
        S->ver_edge[x][y] = (x == 0 || x == 88) ? 1 : 0;
        S->hor_segd[x][y] = 0; S->ver_segd[x][y] = 0;
    }
    /* map_loader.py:18-41.  tile_dic insertion order: inner tiles x-major, then the border keys */
    for (int x = 0; x < NCX; x++) for (int y = 0; y < NCY; y++) S->tile[x][y] = 0;
    for (int x = 0; x < 42; x++) for (int y = 0; y < 23; y++) S->tile[x + 1][y + 1] = S->map_data[184 + x + y * 42];
    for (int x = 0; x < 44; x++) { S->tile[x][0] = 1; S->tile[x][24] = 1; }
    for (int y = 0; y < 25; y++) { S->tile[0][y] = 1; S->tile[43][y] = 1; }
    /* Every effect of one tile is either commutative (edge parity, ortho sums) or private to its
       own cell (diag/circular append), so any visiting order gives the reference's result. */
    for (int x = 0; x < NCX; x++) for (int y = 0; y < NCY; y++) if (S->tile[x][y] != 0) process_tile(S, S->tile[x][y], x, y);
    /* tile_segment_factory.py:217-262: dict order = x-major over 89x51 */
    for (int x = 0; x < NHX; x++) for (int y = 0; y < NHY; y++) {
        int state = S->hor_segd[x][y];
        if (state == 0) continue;
        int cx = ifloor(x / 2.0), cy = ifloor((y - 0.1 * state) / 2);
        if (cx < 0 || cx >= NCX || cy < 0 || cy >= NCY) return -1;    /* KeyError in the reference */
        double p1x = 12 * x, p1y = 12 * y, p2x = 12 * x + 12, p2y = 12 * y;
        if (state == -1) { double tx = p1x, ty = p1y; p1x = p2x; p1y = p2y; p2x = tx; p2y = ty; }
        iv_push(&S->cells[cx][cy].segs, add_linear(S, p1x, p1y, p2x, p2y));
    }
    for (int x = 0; x < NHX; x++) for (int y = 0; y < NHY; y++) {
        int state = S->ver_segd[x][y];
        if (state == 0) continue;
        int cx = ifloor((x - 0.1 * state) / 2), cy = ifloor(y / 2.0);
        if (cx < 0 || cx >= NCX || cy < 0 || cy >= NCY) return -1;
        double p1x = 12 * x, p1y = 12 * y + 12, p2x = 12 * x, p2y = 12 * y;
        if (state == -1) { double tx = p1x, ty = p1y; p1x = p2x; p1y = p2y; p2x = tx; p2y = ty; }
        iv_push(&S->cells[cx][cy].segs, add_linear(S, p1x, p1y, p2x, p2y));
    }
    /* spatial_segment_index.py:69-110 */
    for (int x = 0; x < NCX; x++) for (int y = 0; y < NCY; y++) {
        Cell *c = &S->cells[x][y];
        if (c->segs.n == 0) continue;
        c->present = 1;
        c->b[0] = c->b[1] = INFINITY; c->b[2] = c->b[3] = -INFINITY;
        for (int k = 0; k < c->segs.n; k++) {
            const double *b = S->segs[c->segs.v[k]].b;
            if (b[0] < c->b[0]) c->b[0] = b[0];
            if (b[1] < c->b[1]) c->b[1] = b[1];
            if (b[2] > c->b[2]) c->b[2] = b[2];
            if (b[3] > c->b[3]) c->b[3] = b[3];
        }
    }
    return 0;
}

static Ent *new_ent(OSim *S, int kind, int type, int dic_type, double xcoord, double ycoord) {  /* entities.py:209-236 */
    if (S->nents >= MAX_ENTS) return NULL;
    Ent *e = &S->ents[S->nents++];
    memset(e, 0, sizeof(*e));
    e->kind = kind; e->type = type; e->dic_type = dic_type;
    e->index = S->entity_counts[type]; S->entity_counts[type] += 1;
    e->xpos = xcoord * 6; e->ypos = ycoord * 6;
    e->active = 1;
    e->cellx = clampi(ifloor(e->xpos / 24), 0, 43); e->celly = clampi(ifloor(e->ypos / 24), 0, 24);
    return e;
}
static void remember_initial(Ent *e) { e->ix = e->xpos; e->iy = e->ypos; e->icx = e->cellx; e->icy = e->celly; e->istate = e->state; }

static int door_init(OSim *S, Ent *e, double orientation, double sw_x, double sw_y) {   /* entity_door_base.py:52-97 */
    e->logical = 1; e->closed = 1;
    double sw_xpos = 6 * sw_x, sw_ypos = 6 * sw_y;
    e->is_vertical = (orientation == 0 || orientation == 4);
    double vx, vy;
    if (!orientation_to_vector(orientation, &vx, &vy)) return -1;
    int dcx = clampi(ifloor((e->xpos - 12 * vx) / 24), 0, 43);
    int dcy = clampi(ifloor((e->ypos - 12 * vy) / 24), 0, 24);
    int hx = 2 * (dcx + 1), hy = 2 * (dcy + 1);
    if (e->is_vertical) {
        e->edge[0][0] = hx; e->edge[0][1] = hy - 2; e->edge[1][0] = hx; e->edge[1][1] = hy - 1;
        for (int k = 0; k < 2; k++) { if (e->edge[k][0] >= NHX || e->edge[k][1] >= NHY) return -1; S->ver_edge[e->edge[k][0]][e->edge[k][1]] += 1; }
    } else {
        e->edge[0][0] = hx - 2; e->edge[0][1] = hy; e->edge[1][0] = hx - 1; e->edge[1][1] = hy;
        for (int k = 0; k < 2; k++) { if (e->edge[k][0] >= NHX || e->edge[k][1] >= NHY) return -1; S->hor_edge[e->edge[k][0]][e->edge[k][1]] += 1; }
    }
    /* the door's own segment goes to segment_dic only; the spatial index was built before
       entities existed (map_loader.py:62-63), so no query ever returns it.  Not modelled. */
    e->xpos = sw_xpos; e->ypos = sw_ypos;
    e->cellx = clampi(ifloor(e->xpos / 24), 0, 43); e->celly = clampi(ifloor(e->ypos / 24), 0, 24);
    return 0;
}

static void clear_grid(OSim *S) { for (int x = 0; x < NCX; x++) for (int y = 0; y < NCY; y++) S->grid[x][y].n = 0; }

static int load_entities(OSim *S) {                          /* map_loader.py:69-145, utils/entity_factory.py:145-233 */
    const double *m = S->map_data; int len = S->map_len;
    if (len < 1235) return -2;
    ninja_init(S, 1);
    int index = 1230;
    int exit_door_count = (int)m[1156];
    /* per-type lists in creation order; dic_order is assembled afterwards in key order 1..28 */
    while (index < len) {
        if (index + 4 >= len) break;
        double type = m[index], xc = m[index + 1], yc = m[index + 2], orientation = m[index + 3], mode = m[index + 4];
        int t = (type == floor(type) && type >= 0 && type < 40) ? (int)type : -1;
        Ent *e = NULL; int in_grid = 1;
        switch (t) {
        case 1: case 21:
            if (!(e = new_ent(S, K_MINE, t, t, xc, yc))) return -3;
            e->thinkable = 1; e->logical = 1; mine_set_state(e, t == 1 ? 0 : 1); remember_initial(e); break;
        case 2:
            if (!(e = new_ent(S, K_GOLD, t, t, xc, yc))) return -3;
            e->logical = 1; remember_initial(e); break;
        case 3: {
            Ent *p = new_ent(S, K_EXIT, 3, 3, xc, yc);
            if (!p) return -3;
            p->logical = 1; remember_initial(p);
            int pid = (int)(p - S->ents);
            int ci = index + 5 * exit_door_count + 1;
            if (ci + 1 >= len) return -4;                     /* IndexError */
            if (!(e = new_ent(S, K_EXIT_SWITCH, 4, 3, m[ci], m[ci + 1]))) return -3;
            e->logical = 1; e->parent = pid; remember_initial(e);
            break; }
        case 5:
            if (!(e = new_ent(S, K_DOOR_REGULAR, t, t, xc, yc))) return -3;
            if (door_init(S, e, orientation, xc, yc)) return -5;
            e->thinkable = 1; e->open_timer = 0; break;
        case 6: case 8:
            if (index + 7 >= len) return -4;
            if (!(e = new_ent(S, t == 6 ? K_DOOR_LOCKED : K_DOOR_TRAP, t, t, xc, yc))) return -3;
            if (door_init(S, e, orientation, m[index + 6], m[index + 7])) return -5;
            if (t == 8) door_change_state(S, e, 0);
            remember_initial(e); break;
        case 10: case 11:
            if (!(e = new_ent(S, t == 10 ? K_LAUNCH_PAD : K_ONE_WAY, t, t, xc, yc))) return -3;
            e->logical = 1; if (t == 11) e->physical = 1;
            if (!orientation_to_vector(orientation, &e->normal_x, &e->normal_y)) return -5;
            break;
        case 14: case 26:
            if (!(e = new_ent(S, t == 14 ? K_DRONE_ZAP : K_MINI_DRONE, t, t, xc, yc))) return -3;
            e->movable = 1; e->logical = 1;
            e->speed = t == 14 ? 8.0 / 7 : 1.3;
            e->radius = t == 14 ? 7.5 : 4.0;
            e->grid_width = t == 14 ? 24 : 12;
            e->dir = (int)floor(orientation / 2);            /* orientation // 2 */
            if (e->dir < 0 || e->dir > 3 || !(mode == 0 || mode == 1 || mode == 2 || mode == 3)) return -5;
            e->mode = (int)mode;
            e->xtarget = e->xpos; e->ytarget = e->ypos;
            break;
        case 17:
            if (!(e = new_ent(S, K_BOUNCE, t, t, xc, yc))) return -3;
            e->physical = 1; e->logical = 1; e->movable = 1;
            e->xorigin = e->xpos; e->yorigin = e->ypos; break;
        case 20:
            if (!(e = new_ent(S, K_THWUMP, t, t, xc, yc))) return -3;
            e->movable = 1; e->thinkable = 1; e->logical = 1; e->physical = 1;
            e->is_horizontal = (orientation == 0 || orientation == 4);
            e->direction = (orientation == 0 || orientation == 2) ? 1 : -1;
            e->xorigin = e->xpos; e->yorigin = e->ypos; e->state = 0; break;
        case 24:
            if (!(e = new_ent(S, K_BOOST, t, t, xc, yc))) return -3;
            e->movable = 1; break;
        case 25:
            if (!(e = new_ent(S, K_DEATH_BALL, t, t, xc, yc))) return -3;
            e->thinkable = 1; e->logical = 1; break;
        case 28:
            if (!(e = new_ent(S, K_SHWUMP, t, t, xc, yc))) return -3;
            e->thinkable = 1; e->logical = 1; e->physical = 1;
            e->xorigin = e->xpos; e->yorigin = e->ypos; break;
        default: break;
        }
        if (e && in_grid) iv_push(&S->grid[e->cellx][e->celly], (int)(e - S->ents));
        if (t == 6 || t == 8) {                               /* map_loader.py:126-141 */
            if (index + 9 < len) {
                if (m[index + 7] != 0 && m[index + 8] == 0 && m[index + 9] == 0) index += 10; else index += 9;
            } else index += 9;
        } else index += 5;
    }
    /* entity_dic = {1..28: []} in key order, each list in creation order (nsim.py:190-192) */
    S->ndic = 0;
    for (int key = 1; key <= 28; key++)
        for (int k = 0; k < S->nents; k++) if (S->ents[k].dic_type == key) S->dic_order[S->ndic++] = k;
    return 0;
}

/* ------------------------------------------------------------------------- */
/* public C API (ctypes)                                                       */
/* ------------------------------------------------------------------------- */
OSim *osim_create(void) {
    OSim *S = (OSim *)calloc(1, sizeof(OSim));
    S->ents = (Ent *)calloc(MAX_ENTS, sizeof(Ent));
    return S;
}
void osim_destroy(OSim *S) {
    if (!S) return;
    free(S->map_data); free(S->segs); free(S->ents);
    for (int x = 0; x < NCX; x++) for (int y = 0; y < NCY; y++) { free(S->cells[x][y].segs.v); free(S->grid[x][y].v); }
    free(S->ninja.cached_entities.v); free(S->tmp_segs.v); free(S->tmp_ents.v); free(S->movable.v); free(S->thinkable.v);
    free(S);
}
/* Simulator.reset: nsim.py:62-76 (+ reset_map_entity_data 142-192) */
int osim_reset(OSim *S) {
    S->frame = 0;
    /* door grid edges are decremented with max(0, v-1) regardless of the door's state (nsim.py:174-185) */
    for (int k = 0; k < S->nents; k++) {
        Ent *e = &S->ents[k];
        if (e->kind == K_DOOR_REGULAR || e->kind == K_DOOR_LOCKED || e->kind == K_DOOR_TRAP)
            for (int j = 0; j < 2; j++) {
                int *p = e->is_vertical ? &S->ver_edge[e->edge[j][0]][e->edge[j][1]] : &S->hor_edge[e->edge[j][0]][e->edge[j][1]];
                *p = (*p - 1 > 0) ? *p - 1 : 0;
            }
    }
    clear_grid(S);
    S->nents = 0; S->ndic = 0; S->events = 0;
    return load_entities(S);
}
/* Simulator.load: nsim.py:51-56.  A fresh Simulator per load is assumed for entity_counts
   (entities.py:213-216 keeps the counters for the life of the Simulator object; pass
   keep_counts=1 to osim_load_ex to model a re-used Simulator). */
int osim_load_ex(OSim *S, const double *map_data, int len, int keep_counts) {
    if (len < 1235) return -2;
    free(S->map_data);
    S->map_data = (double *)malloc(sizeof(double) * len);
    memcpy(S->map_data, map_data, sizeof(double) * len);
    S->map_len = len;
    if (!keep_counts) memset(S->entity_counts, 0, sizeof(S->entity_counts));
    /* load() calls reset_map_tile_data + load_map_tiles, then reset(); reset() first undoes the
       previous level's door edges, which load_tiles has just overwritten -- skip that part by
       dropping the old entities here (their edges no longer exist in the fresh tables). */
    S->nents = 0;
    int rc = load_tiles(S);
    if (rc) return rc;
    return osim_reset(S);
}
int osim_load(OSim *S, const double *map_data, int len) { return osim_load_ex(S, map_data, len, 0); }

/* Simulator.fast_reset: nsim.py:78-140 */
int osim_fast_reset(OSim *S) {
    S->frame = 0; S->events = 0;
    if (!S->have_ninja) return load_entities(S);
    ninja_init(S, 0);
    clear_grid(S);
    for (int k = 0; k < S->ndic; k++) {
        int id = S->dic_order[k];
        Ent *e = &S->ents[id];
        switch (e->kind) {       /* classes that define reset_state() */
        case K_MINE: e->xpos = e->ix; e->ypos = e->iy; e->cellx = e->icx; e->celly = e->icy; e->active = 1; mine_set_state(e, e->istate); break;
        case K_GOLD: case K_EXIT_SWITCH: e->xpos = e->ix; e->ypos = e->iy; e->cellx = e->icx; e->celly = e->icy; e->active = 1; break;
        case K_EXIT: e->xpos = e->ix; e->ypos = e->iy; e->cellx = e->icx; e->celly = e->icy; e->switch_hit = 0; break;
        case K_DOOR_LOCKED: e->xpos = e->ix; e->ypos = e->iy; e->cellx = e->icx; e->celly = e->icy; e->active = 1; if (!e->closed) door_change_state(S, e, 1); break;
        case K_DOOR_TRAP: e->xpos = e->ix; e->ypos = e->iy; e->cellx = e->icx; e->celly = e->icy; e->active = 1; if (e->closed) door_change_state(S, e, 0); break;
        default: e->active = 1; break;   /* no reset_state(): only `active` is restored (nsim.py:131-135) */
        }
        if (e->kind != K_EXIT) iv_push(&S->grid[e->cellx][e->celly], id);
    }
    return 0;
}

/* IntermediateGoalManager.apply_to_simulator
   (gym_environment/reward_calculation/intermediate_goal_manager.py:698-960): the curriculum moves the LAST
   EntityExitSwitch of entity_dic[3] (reversed() scan, :725-729) and its parent door.  Returns -1 when the
   level has no exit switch (the reference raises RuntimeError, :768-781). */
int osim_apply_curriculum(OSim *S, double sx, double sy, double dx, double dy) {
    int sw = -1;
    for (int k = S->ndic - 1; k >= 0; k--) if (S->ents[S->dic_order[k]].kind == K_EXIT_SWITCH) { sw = S->dic_order[k]; break; }
    if (sw < 0) return -1;
    Ent *e = &S->ents[sw];
    e->xpos = sx; e->ypos = sy;                                  /* :810-811 */
    if (S->frame <= 10) e->active = 1;                           /* :843-845 */
    int cx = clampi(ifloor(sx / 24), 0, 43), cy = clampi(ifloor(sy / 24), 0, 24);   /* :877-880 */
    if (cx != e->cellx || cy != e->celly) {                      /* :881-894 */
        iv_remove(&S->grid[e->cellx][e->celly], sw);
        e->cellx = cx; e->celly = cy;
        iv_push(&S->grid[cx][cy], sw);
    }
    Ent *d = &S->ents[e->parent];                                /* :928-960: the door is re-celled but never added */
    d->xpos = dx; d->ypos = dy;
    cx = clampi(ifloor(dx / 24), 0, 43); cy = clampi(ifloor(dy / 24), 0, 24);
    if (cx != d->cellx || cy != d->celly) {
        iv_remove(&S->grid[d->cellx][d->celly], e->parent);
        iv_remove(&S->grid[cx][cy], e->parent);
        d->cellx = cx; d->celly = cy;
    }
    return 0;
}

/* Simulator.tick: nsim.py:221-292 */
void osim_tick(OSim *S, int hor_input, int jump_input) {
    Ninja *n = &S->ninja;
    S->frame += 1;
    n->hor_input = hor_input; n->jump_input = jump_input;
    n->xpos_at_action = n->xpos; n->ypos_at_action = n->ypos;    /* ninja.py:841-847 */
    S->movable.n = 0; S->thinkable.n = 0;
    for (int k = 0; k < S->ndic; k++) {
        Ent *e = &S->ents[S->dic_order[k]];
        if (e->active) { if (e->movable) iv_push(&S->movable, S->dic_order[k]); if (e->thinkable) iv_push(&S->thinkable, S->dic_order[k]); }
    }
    for (int k = 0; k < S->movable.n; k++) ent_move(S, S->movable.v[k]);
    for (int k = 0; k < S->thinkable.n; k++) ent_think(S, S->thinkable.v[k]);
    if (n->state != 9) {
        if (n->state != 6) {          /* ragdoll is None when enable_anim=False (ninja.py:122-128) */
            ninja_integrate(n);
            ninja_pre_collision(S);
            for (int i = 0; i < 4; i++) { ninja_collide_vs_objects(S); ninja_collide_vs_tiles(S); }
            ninja_post_collision(S);
        }
        ninja_think(n);
    }
}

double osim_sq_probe(double x) { return SQ(x); }
int osim_num_entities(const OSim *S) { return S->ndic; }
int osim_num_segments(const OSim *S) { return S->nsegs; }
int osim_events(OSim *S) { int e = S->events; S->events = 0; return e; }

/* ------------------------------------------------------------------------- */
/* observation primitives                                 */
/* ------------------------------------------------------------------------- */
static double dmin(double a, double b) { return b < a ? b : a; }      /* Python min(a, b): first on ties */
static double dmax(double a, double b) { return b > a ? b : a; }
/* NPlayHeadless.get_ninja_state: nplay_headless.py:735-917 (+ Ninja.get_temporal_features ninja.py:1217-1244).
   `x ** 0.5` is libm pow(x, 0.5); math.sqrt / math.atan2 are libm sqrt / atan2. */
void osim_ninja_state(const OSim *S, double *o) {
    const Ninja *n = &S->ninja;
    int k = 0;
    double velocity_magnitude = pow(SQ(n->xspeed) + SQ(n->yspeed), 0.5);
    double max_velocity = MAX_HOR_SPEED * 2;
    o[k++] = dmin(velocity_magnitude / max_velocity, 1.0) * 2 - 1;
    if (velocity_magnitude > 1e-6) { o[k++] = n->xspeed / velocity_magnitude; o[k++] = n->yspeed / velocity_magnitude; }
    else { o[k++] = 0.0; o[k++] = 0.0; }
    o[k++] = (n->state >= 0 && n->state <= 2) ? 1.0 : -1.0;
    o[k++] = (n->state == 3 || n->state == 4) ? 1.0 : -1.0;
    o[k++] = n->state == 5 ? 1.0 : -1.0;
    o[k++] = (n->state >= 6 && n->state <= 9) ? 1.0 : -1.0;
    o[k++] = n->airborn ? 1.0 : -1.0;
    o[k++] = (double)n->hor_input;
    o[k++] = n->jump_input ? 1.0 : -1.0;
    o[k++] = ((n->jump_buffer > 0 ? n->jump_buffer : 0) / 5.0) * 2 - 1;
    o[k++] = ((n->floor_buffer > 0 ? n->floor_buffer : 0) / 5.0) * 2 - 1;
    o[k++] = ((n->wall_buffer > 0 ? n->wall_buffer : 0) / 5.0) * 2 - 1;
    o[k++] = (double)((n->floor_count < 1 ? n->floor_count : 1) * 2 - 1);
    o[k++] = (double)((n->wall_count < 1 ? n->wall_count : 1) * 2 - 1);
    o[k++] = (double)((n->ceiling_count < 1 ? n->ceiling_count : 1) * 2 - 1);
    o[k++] = pow(SQ(n->floor_normalized_x) + SQ(n->floor_normalized_y), 0.5) * 2 - 1;
    o[k++] = n->wall_count > 0 ? n->wall_normal : 0.0;
    o[k++] = n->floor_normalized_y;
    {
        double gravity_range = GRAVITY_FALL - GRAVITY_JUMP;
        o[k++] = gravity_range > 1e-6 ? (n->applied_gravity - GRAVITY_JUMP) / gravity_range * 2 - 1 : 0.0;
    }
    o[k++] = n->walled ? 1.0 : -1.0;
    o[k++] = n->floor_normalized_x;
    o[k++] = n->ceiling_normalized_x;
    o[k++] = n->ceiling_normalized_y;
    {
        double drag_range = DRAG_REGULAR - DRAG_SLOW;
        o[k++] = drag_range > 1e-6 ? (n->applied_drag - DRAG_SLOW) / drag_range * 2 - 1 : 0.0;
        double friction_range = FRICTION_GROUND - FRICTION_GROUND_SLOW;
        o[k++] = friction_range > 1e-6 ? (n->applied_friction - FRICTION_GROUND_SLOW) / friction_range * 2 - 1 : 0.0;
    }
    /* temporal features */
    o[k++] = dmax(-1.0, dmin(1.0, (n->xspeed - n->xspeed_old) / MAX_HOR_SPEED));
    o[k++] = dmax(-1.0, dmin(1.0, (n->yspeed - n->yspeed_old) / MAX_HOR_SPEED));
    o[k++] = dmin(sqrt(SQ(n->xspeed) + SQ(n->yspeed)) / (MAX_HOR_SPEED * 1.5), 1.0) * 2 - 1;
    o[k++] = dmin(n->frames_airborne / 60.0, 1.0) * 2 - 1;
    o[k++] = dmin((double)n->jump_duration / MAX_JUMP_DURATION, 1.0) * 2 - 1;
    o[k++] = dmin(n->state_change_frame / 30.0, 1.0) * 2 - 1;
    /* physics features */
    double kinetic_energy = 0.5 * (SQ(n->xspeed) + SQ(n->yspeed));
    o[k++] = dmin(kinetic_energy / SQ(MAX_HOR_SPEED), 1.0) * 2 - 1;
    o[k++] = (n->ypos / 600.0) * 2 - 1;
    {
        double force_magnitude = sqrt(SQ(n->applied_gravity) + SQ(!n->airborn ? GROUND_ACCEL : AIR_ACCEL));
        o[k++] = dmin(force_magnitude / 0.1, 1.0) * 2 - 1;
    }
    {
        double prev_kinetic = 0.5 * (SQ(n->xspeed_old) + SQ(n->yspeed_old));
        double energy_change_rate = (kinetic_energy - prev_kinetic) / dmax(kinetic_energy + 0.01, 0.01);
        o[k++] = dmax(-1.0, dmin(1.0, energy_change_rate));
    }
    o[k++] = dmin(n->floor_count / 5.0, 1.0) * 2 - 1;
    o[k++] = dmin(n->wall_count / 3.0, 1.0) * 2 - 1;
    o[k++] = atan2(n->floor_normalized_y, n->floor_normalized_x) / 3.141592653589793;   /* math.pi */
    o[k++] = n->walled ? n->wall_normal : 0.0;
}
/* Ninja.get_valid_action_mask: ninja.py:628-839, masking mode "hard" with straightness masking enabled (the
   defaults); path_dir: 0 None, 1 "left", 2 "right", 3 "down" (sim._path_straightness_direction). */
void osim_action_mask(const OSim *S, int path_dir, uint8_t *mask) {
    const Ninja *n = &S->ninja;
    for (int i = 0; i < 6; i++) mask[i] = 1;
    int has_active_buffer = (-1 < n->jump_buffer && n->jump_buffer < 5) || (-1 < n->floor_buffer && n->floor_buffer < 5) ||
                            (-1 < n->wall_buffer && n->wall_buffer < 5) || (-1 < n->launch_pad_buffer && n->launch_pad_buffer < 4);
    int jump_useless = 0;
    if (n->airborn && n->state != 3) jump_useless = n->jump_input != 0 ? !has_active_buffer : 0;
    if (jump_useless) { mask[3] = 0; mask[4] = 0; mask[5] = 0; }
    if (n->walled) {
        if (!n->airborn || n->state == 5) {
            if (n->wall_normal > 0) mask[1] = 0; else if (n->wall_normal < 0) mask[2] = 0;
        } else {
            int would_trigger_wall_slide = n->yspeed >= 0;
            if (n->wall_normal > 0) { if (!would_trigger_wall_slide) mask[1] = 0; }
            else if (n->wall_normal < 0) { if (!would_trigger_wall_slide) mask[2] = 0; }
        }
    }
    if (path_dir == 1) { mask[2] = 0; mask[3] = 0; mask[4] = 0; mask[5] = 0; }
    else if (path_dir == 2) { mask[1] = 0; mask[3] = 0; mask[4] = 0; mask[5] = 0; }
    else if (path_dir == 3) { mask[3] = 0; mask[4] = 0; mask[5] = 0; }
    int any = 0;
    for (int i = 0; i < 6; i++) any |= mask[i];
    if (!any) mask[0] = 1;
}

/* canonical record of oracle/schema.py: nin_f[31], nin_i[27], ent_f[E][4], ent_i[E][8] */
void osim_get_state(const OSim *S, double *nf, int64_t *ni, double *ef, int64_t *ei) {
    const Ninja *n = &S->ninja;
    const double f[31] = { n->xpos, n->ypos, n->xspeed, n->yspeed, n->xpos_old, n->ypos_old, n->xspeed_old, n->yspeed_old,
        n->applied_gravity, n->applied_drag, n->applied_friction, n->floor_normalized_x, n->floor_normalized_y,
        n->ceiling_normalized_x, n->ceiling_normalized_y, n->wall_normal, n->xlp_boost_normalized, n->ylp_boost_normalized,
        n->floor_normal_x, n->floor_normal_y, n->ceiling_normal_x, n->ceiling_normal_y, n->x_crush, n->y_crush, n->crush_len,
        n->death_xpos, n->death_ypos, n->death_xspeed, n->death_yspeed, n->xpos_at_action, n->ypos_at_action };
    const int64_t i[27] = { n->state, n->airborn, n->airborn_old, n->walled, n->jump_input_old, n->jump_duration,
        n->jump_buffer, n->floor_buffer, n->wall_buffer, n->launch_pad_buffer, n->is_jump_useless, n->last_jump_was_buffered,
        n->floor_count, n->ceiling_count, n->is_crushable, n->gold_collected, n->doors_opened, n->terminal_impact, n->death_cause,
        n->frames_airborne, n->state_change_frame, n->previous_state, n->consecutive_floor_frames, n->consecutive_wall_frames,
        n->hor_input, n->jump_input, S->frame };
    if (nf) memcpy(nf, f, sizeof(f));
    if (ni) memcpy(ni, i, sizeof(i));
    if (!ef || !ei) return;
    for (int k = 0; k < S->ndic; k++) {
        int id = S->dic_order[k];
        const Ent *e = &S->ents[id];
        int64_t s0 = 0, s1 = 0; double a0 = 0, a1 = 0;
        switch (e->kind) {
        case K_MINE: case K_THWUMP: s0 = e->state; break;
        case K_EXIT: s0 = e->switch_hit; break;
        case K_DOOR_REGULAR: s0 = e->closed; s1 = e->open_timer; break;
        case K_DOOR_LOCKED: case K_DOOR_TRAP: s0 = e->closed; break;
        case K_BOUNCE: case K_DEATH_BALL: a0 = e->xspeed; a1 = e->yspeed; break;
        case K_DRONE_ZAP: case K_MINI_DRONE: s0 = e->dir; a0 = e->xtarget; a1 = e->ytarget; break;
        case K_BOOST: s0 = e->touching; break;
        case K_SHWUMP: s0 = e->state; s1 = e->activated; a0 = e->xdir; a1 = e->ydir; break;
        default: break;
        }
        int rank = -1;
        const IVec *l = &S->grid[e->cellx][e->celly];
        for (int r = 0; r < l->n; r++) if (l->v[r] == id) { rank = r; break; }
        ef[4 * k + 0] = e->xpos; ef[4 * k + 1] = e->ypos; ef[4 * k + 2] = a0; ef[4 * k + 3] = a1;
        int64_t *o = ei + 8 * k;
        o[0] = e->type; o[1] = e->index; o[2] = e->active; o[3] = e->cellx; o[4] = e->celly; o[5] = s0; o[6] = s1; o[7] = rank;
    }
}

/* T ticks with per-frame records: nf[T][31], ni[T][27], ef[T][E][4], ei[T][E][8] (ef/ei may be NULL) */
void osim_run(OSim *S, const int8_t *hor, const int8_t *jump, int T, double *nf, int64_t *ni, double *ef, int64_t *ei) {
    int E = S->ndic;
    for (int t = 0; t < T; t++) {
        osim_tick(S, hor[t], jump[t]);
        osim_get_state(S, nf ? nf + 31 * (size_t)t : NULL, ni ? ni + 27 * (size_t)t : NULL,
                       ef ? ef + 4 * (size_t)E * t : NULL, ei ? ei + 8 * (size_t)E * t : NULL);
    }
}

/* static tables, for checking the product's map compiler: per cell segment counts, and the
   segment list in cell order (x-major), 8 doubles per segment:
   kind, then linear: x1,y1,x2,y2,0,0,0 | circular: cx,cy,hor,ver,convex,0,0 */
void osim_get_cell_counts(const OSim *S, int32_t *counts /*[44*25] x-major*/, int32_t *present) {
    for (int x = 0; x < NCX; x++) for (int y = 0; y < NCY; y++) {
        counts[x * NCY + y] = S->cells[x][y].segs.n;
        present[x * NCY + y] = S->cells[x][y].present;
    }
}
void osim_get_segments(const OSim *S, double *out) {
    int k = 0;
    for (int x = 0; x < NCX; x++) for (int y = 0; y < NCY; y++) {
        const IVec *l = &S->cells[x][y].segs;
        for (int j = 0; j < l->n; j++, k++) {
            const Seg *s = &S->segs[l->v[j]];
            double *o = out + 8 * k;
            memset(o, 0, 8 * sizeof(double));
            o[0] = s->kind;
            if (s->kind == 0) { o[1] = s->x1; o[2] = s->y1; o[3] = s->x2; o[4] = s->y2; }
            else { o[1] = s->cx; o[2] = s->cy; o[3] = s->hor; o[4] = s->ver; o[5] = s->convex; }
        }
    }
}
void osim_get_grid_edges(const OSim *S, int32_t *hor /*[89*51]*/, int32_t *ver) {
    for (int x = 0; x < NHX; x++) for (int y = 0; y < NHY; y++) { hor[x * NHY + y] = S->hor_edge[x][y]; ver[x * NHY + y] = S->ver_edge[x][y]; }
}
