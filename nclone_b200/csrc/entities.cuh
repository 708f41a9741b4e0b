// entities.cuh -- entity update (move/think) and collision logic, plus the ordered
// neighbourhood gather.  Reference: /root/reference/nclone/entities.py and entity_classes/*.
//
// Ordering is reproduced with per-env insertion stamps instead of Python
// lists: an entity's place inside grid_entity[cell] is its stamp; grid_move / the exit door's
// late insertion take the next value of a per-env counter (append semantics,
// entities.py:256-264, entity_exit_switch.py:120).  Entities that never change cell live in a
// per-map static cell->ids table; the few that roam are found by scanning their slots.
// Included once per kernel variant (sim_variant.inc) with NPP_KINDS = the entity kinds that variant supports:
// HAS(kind) is a compile-time constant, so the code of absent kinds (and whole phases: roaming-entity
// move/think, physical collisions) disappears from the specialised kernel.

struct EntS { double x, y; int kind, p0, p1, dyn, aux, index, load_order; };
DEV EntS ld_ent(const Ctx& c, int id) {
    const NppStaticEnt* e = c.ents + id;
    double2 p = __ldg(reinterpret_cast<const double2*>(e));
    uint4 m = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(e) + 16));
    EntS s;
    s.x = p.x; s.y = p.y;
    s.kind = m.x & 255; s.p0 = (m.x >> 16) & 255; s.p1 = m.x >> 24;
    s.dyn = (int)(int16_t)(m.y & 0xffff); s.aux = (int)(int16_t)(m.y >> 16);
    s.index = m.z >> 16; s.load_order = m.w & 0xffff;
    return s;
}
DEV int ent_kind(const Ctx& c, int id) { return __ldg(reinterpret_cast<const uint8_t*>(c.ents + id) + 16); }

DEV bool ninja_valid_target(const Nin& n) { int s = n.i[I_STATE]; return !(s == 6 || s == 8 || s == 9); }   // ninja.py:1272-1274
DEV void ninja_win(Nin& n, Ctx& c) {                                   // ninja.py:1246-1251
    if (n.i[I_STATE] < 6) {
        if (n.i[I_STATE] == 3) n.f[F_GRAVITY] = GRAVITY_FALL;
        n.i[I_STATE] = 8;
        c.events |= NPPSIM_EV_WIN;
    }
}
DEV void ninja_kill(Nin& n, Ctx& c, double x, double y, double xs, double ys, int cause) {   // ninja.py:1253-1270
    if (n.i[I_STATE] < 6) {
        GF(c, F_DEATH_X) = x; GF(c, F_DEATH_Y) = y; GF(c, F_DEATH_XS) = xs; GF(c, F_DEATH_YS) = ys;
        GI(c, I_DEATH_CAUSE) = cause;
        if (n.i[I_STATE] == 3) n.f[F_GRAVITY] = GRAVITY_FALL;
        n.i[I_STATE] = 7;
        c.events |= NPPSIM_EV_DEATH;
    }
}

DEV int cell_index(double x, double y) {
    return clampi(floordiv24(x), 0, 43) * NPP_NCY + clampi(floordiv24(y), 0, 24);
}
// entities.py:256-264 for a roaming entity in dynamic slot `slot`
DEV void grid_move(Ctx& c, int slot, double x, double y) {
    int cell = cell_index(x, y);
    if (cell != DYNI(c, slot, 0)) { DYNI(c, slot, 0) = cell; DYNI(c, slot, 1) = c.stamp++; }
}
DEV double mine_radius(int state) { return state == 0 ? 4.0 : (state == 1 ? 3.5 : 4.5); }   // physics_constants.py:48

// ---- doors ------------------------------------------------------------------------------------------
// entity_door_base.py:99-108 for locked / trap doors: the edge contribution follows FL_CLOSED.
DEV void door_set_closed(Ctx& c, int id, bool closed) {
    uint8_t f = FLAG(c, id);
    FLAG(c, id) = closed ? (f | FL_CLOSED) : (f & ~FL_CLOSED);
}
// regular door: aux = open_timer (low 8 bits) | signed edge counter << 8
DEV void regular_door_change_state(Ctx& c, int id, int slot, bool closed) {
    door_set_closed(c, id, closed);
    DYNI(c, slot, 2) += closed ? 256 : -256;
}

// ---- drones: entity_drone_base.py:66-164 ------------------------------------------------------------
DEV bool drone_test_dir(Ctx& c, int dir, double xpos, double ypos, double radius, double grid_width, double& xt, double& yt) {
    int xdir = dir == 0 ? 1 : (dir == 2 ? -1 : 0);
    int ydir = dir == 1 ? 1 : (dir == 3 ? -1 : 0);
    double xtarget = xpos + grid_width * xdir, ytarget = ypos + grid_width * ydir;
    if (!ydir) {
        int cell_x = floordiv12(xpos + xdir * radius);
        int cell_xtarget = floordiv12(xtarget + xdir * radius);
        int cell_y1 = floordiv12(ypos - radius), cell_y2 = floordiv12(ypos + radius);
        #pragma unroll 1
        while (cell_x != cell_xtarget) {
            if (!is_empty_column(c, cell_x, cell_y1, cell_y2, xdir)) return false;
            cell_x += xdir;
        }
    } else {
        int cell_y = floordiv12(ypos + ydir * radius);
        int cell_ytarget = floordiv12(ytarget + ydir * radius);
        int cell_x1 = floordiv12(xpos - radius), cell_x2 = floordiv12(xpos + radius);
        #pragma unroll 1
        while (cell_y != cell_ytarget) {
            if (!is_empty_row(c, cell_x1, cell_x2, cell_y, ydir)) return false;
            cell_y += ydir;
        }
    }
    xt = xtarget; yt = ytarget;
    return true;
}
// Rare entity updates are out of line; they take the context BY VALUE (so the caller's copy and the
// ninja record stay in registers) and return the updated insertion-stamp counter.
__device__ __noinline__ int drone_move(Ctx c, int slot, int id, int mode, bool mini) {
    const double speed = mini ? 1.3 : 8.0 / 7, radius = mini ? 4.0 : 7.5, grid_width = mini ? 12.0 : 24.0;
    uint8_t fl = FLAG(c, id);
    int dir = (fl & FL_ST) >> FL_ST_SHIFT;
    double xpos = DYNF(c, slot, 0), ypos = DYNF(c, slot, 1), xtarget = DYNF(c, slot, 2), ytarget = DYNF(c, slot, 3);
    int vx = dir == 0 ? 1 : (dir == 2 ? -1 : 0), vy = dir == 1 ? 1 : (dir == 3 ? -1 : 0);
    double xspeed = speed * vx, yspeed = speed * vy;
    double dx = xtarget - xpos, dy = ytarget - ypos;
    double dist = sqrt(SQ(dx) + SQ(dy));
    if (dist < 0.000001 || (dx * (xtarget - (xpos + xspeed)) + dy * (ytarget - (ypos + yspeed))) < 0) {
        xpos = xtarget; ypos = ytarget;
        bool can_move = false;
        // DIR_LIST rows (entity_drone_base.py:72): mode -> direction offsets tried in order
        const int dl[4] = {0x2301, 0x2103, 0x2310, 0x2130};     // nibbles, first tried in the low nibble
        #pragma unroll 1
        for (int k = 0; k < 4; k++) {
            int new_dir = (dir + ((dl[mode] >> (4 * k)) & 15)) & 3;
            if (drone_test_dir(c, new_dir, xpos, ypos, radius, grid_width, xtarget, ytarget)) { dir = new_dir; can_move = true; break; }
        }
        if (can_move) {
            double disp = speed - dist;
            int wx = dir == 0 ? 1 : (dir == 2 ? -1 : 0), wy = dir == 1 ? 1 : (dir == 3 ? -1 : 0);
            xpos += disp * wx; ypos += disp * wy;
        }
        DYNF(c, slot, 0) = xpos; DYNF(c, slot, 1) = ypos; DYNF(c, slot, 2) = xtarget; DYNF(c, slot, 3) = ytarget;
        FLAG(c, id) = (fl & ~FL_ST) | (dir << FL_ST_SHIFT);
        // no grid_move on this branch (entity_drone_base.py:104-112)
    } else {
        xpos += xspeed; ypos += yspeed;
        DYNF(c, slot, 0) = xpos; DYNF(c, slot, 1) = ypos;
        grid_move(c, slot, xpos, ypos);
    }
    return c.stamp;
}

// ---- thwump: entity_thwump.py:109-206 -----------------------------------------------------------------
DEV int thwump_state(uint8_t fl) { int s = (fl & FL_ST) >> FL_ST_SHIFT; return s == 3 ? -1 : s; }
DEV void thwump_set_state(Ctx& c, int id, int s) { uint8_t fl = FLAG(c, id); FLAG(c, id) = (fl & ~FL_ST) | ((s & 3) << FL_ST_SHIFT); }
__device__ __noinline__ int thwump_move(Ctx c, int slot, int id, int p0) {
    int state = thwump_state(FLAG(c, id));
    if (!state) return c.stamp;
    bool horizontal = p0 & 1; int direction = (p0 & 2) ? -1 : 1;
    double speed = state == 1 ? 20.0 / 7 : 8.0 / 7;
    int speed_dir = direction * state;
    double xpos = DYNF(c, slot, 0), ypos = DYNF(c, slot, 1);
    EntS e = ld_ent(c, id);                                   // origin = initial position
    if (!horizontal) {
        double ypos_new = ypos + speed * speed_dir;
        if (state == -1 && (ypos_new - e.y) * (ypos - e.y) < 0) { DYNF(c, slot, 1) = e.y; thwump_set_state(c, id, 0); return c.stamp; }
        int cell_y = floordiv12(ypos + speed_dir * 11), cell_y_new = floordiv12(ypos_new + speed_dir * 11);
        if (cell_y != cell_y_new) {
            int x1 = floordiv12(xpos - 11), x2 = floordiv12(xpos + 11);
            if (!is_empty_row(c, x1, x2, cell_y, speed_dir)) { thwump_set_state(c, id, -1); return c.stamp; }
        }
        ypos = ypos_new; DYNF(c, slot, 1) = ypos;
    } else {
        double xpos_new = xpos + speed * speed_dir;
        if (state == -1 && (xpos_new - e.x) * (xpos - e.x) < 0) { DYNF(c, slot, 0) = e.x; thwump_set_state(c, id, 0); return c.stamp; }
        int cell_x = floordiv12(xpos + speed_dir * 11), cell_x_new = floordiv12(xpos_new + speed_dir * 11);
        if (cell_x != cell_x_new) {
            int y1 = floordiv12(ypos - 11), y2 = floordiv12(ypos + 11);
            if (!is_empty_column(c, cell_x, y1, y2, speed_dir)) { thwump_set_state(c, id, -1); return c.stamp; }
        }
        xpos = xpos_new; DYNF(c, slot, 0) = xpos;
    }
    grid_move(c, slot, xpos, ypos);
    return c.stamp;
}
__device__ __noinline__ void thwump_think(Ctx c, double ninja_x, double ninja_y, bool ninja_valid, int slot, int id, int p0) {
    if (thwump_state(FLAG(c, id)) || !ninja_valid) return;
    bool horizontal = p0 & 1; int direction = (p0 & 2) ? -1 : 1;
    double xpos = DYNF(c, slot, 0), ypos = DYNF(c, slot, 1);
    const double activation_range = 2 * (9 + NINJA_RADIUS);
    if (!horizontal) {
        if (fabs(xpos - ninja_x) < activation_range) {
            int ninja_ycell = floordiv12(ninja_y);
            int ty = floordiv12(ypos - direction * 11);
            int x1 = floordiv12(xpos - 11), x2 = floordiv12(xpos + 11);
            int dy = ninja_ycell - ty;
            if (dy * direction >= 0) {
                int i;
                #pragma unroll 1
                for (i = 0; i < 100; i++) {
                    if (!is_empty_row(c, x1, x2, ty, direction)) { dy = ninja_ycell - ty; break; }
                    ty += direction;
                }
                if (i == 100) i = 99;                       // Python's loop variable keeps its last value
                if (i > 0 && dy * direction <= 0) thwump_set_state(c, id, 1);
            }
        }
    } else {
        if (fabs(ypos - ninja_y) < activation_range) {
            int ninja_xcell = floordiv12(ninja_x);
            int tx = floordiv12(xpos - direction * 11);
            int y1 = floordiv12(ypos - 11), y2 = floordiv12(ypos + 11);
            int dx = ninja_xcell - tx;
            if (dx * direction >= 0) {
                int i;
                #pragma unroll 1
                for (i = 0; i < 100; i++) {
                    if (!is_empty_column(c, tx, y1, y2, direction)) { dx = ninja_xcell - tx; break; }
                    tx += direction;
                }
                if (i == 100) i = 99;
                if (i > 0 && dx * direction <= 0) thwump_set_state(c, id, 1);
            }
        }
    }
}

// ---- shove thwump: entity_shove_thwump.py:109-153 --------------------------------------------------------
DEV void shwump_set_state(Ctx& c, int id, int s) { uint8_t fl = FLAG(c, id); FLAG(c, id) = (fl & ~FL_ST) | (s << FL_ST_SHIFT); }
DEV void shwump_move_if_possible(Ctx& c, int slot, int id, double xdir, double ydir, double speed) {
    double xpos = DYNF(c, slot, 0), ypos = DYNF(c, slot, 1);
    if (DYNF(c, slot, 3) == 0) {                              // self.ydir == 0
        double xpos_new = xpos + xdir * speed;
        int cell_x = floordiv12(xpos), cell_x_new = floordiv12(xpos_new);
        if (cell_x != cell_x_new) {
            int y1 = floordiv12(ypos - 8), y2 = floordiv12(ypos + 8);
            if (!is_empty_column(c, cell_x, y1, y2, (int)xdir)) { shwump_set_state(c, id, 3); return; }
        }
        xpos = xpos_new; DYNF(c, slot, 0) = xpos;
    } else {
        double ypos_new = ypos + ydir * speed;
        int cell_y = floordiv12(ypos), cell_y_new = floordiv12(ypos_new);
        if (cell_y != cell_y_new) {
            int x1 = floordiv12(xpos - 8), x2 = floordiv12(xpos + 8);
            if (!is_empty_row(c, x1, x2, cell_y, (int)ydir)) { shwump_set_state(c, id, 3); return; }
        }
        ypos = ypos_new; DYNF(c, slot, 1) = ypos;
    }
    grid_move(c, slot, xpos, ypos);
}
__device__ __noinline__ int shwump_think(Ctx c, int slot, int id) {
    uint8_t fl = FLAG(c, id);
    int state = (fl & FL_ST) >> FL_ST_SHIFT;
    if (state == 1) {
        if (fl & FL_ACTIVATED) { FLAG(c, id) = fl & ~FL_ACTIVATED; return c.stamp; }
        state = 2; shwump_set_state(c, id, 2);
    }
    double xdir = DYNF(c, slot, 2), ydir = DYNF(c, slot, 3);
    if (state == 3) {
        EntS e = ld_ent(c, id);
        double origin_dist = fabs(DYNF(c, slot, 0) - e.x) + fabs(DYNF(c, slot, 1) - e.y);
        if (origin_dist >= 1) shwump_move_if_possible(c, slot, id, xdir, ydir, 1);
        else { DYNF(c, slot, 0) = e.x; DYNF(c, slot, 1) = e.y; shwump_set_state(c, id, 0); }
    } else if (state == 2) {
        // -xdir of a Python int 0 is 0; the product with speed is only ever added to a position
        shwump_move_if_possible(c, slot, id, xdir == 0 ? 0.0 : -xdir, ydir == 0 ? 0.0 : -ydir, 4);
    }
    return c.stamp;
}

// ---- death ball: entity_death_ball.py:74-167 -----------------------------------------------------------
__device__ __noinline__ int death_ball_think(Ctx c, double ninja_x, double ninja_y, bool ninja_valid, int slot, int id) {
    double xpos = DYNF(c, slot, 0), ypos = DYNF(c, slot, 1), xspeed = DYNF(c, slot, 2), yspeed = DYNF(c, slot, 3);
    if (!ninja_valid) { xspeed *= 0.95; yspeed *= 0.95; }
    else {
        double dx = ninja_x - xpos, dy = ninja_y - ypos;
        double dist = sqrt(SQ(dx) + SQ(dy));
        if (dist > 0) { dx /= dist; dy /= dist; }
        xspeed += dx * 0.04; yspeed += dy * 0.04;
        double speed = sqrt(SQ(xspeed) + SQ(yspeed));
        if (speed > 0.85) {
            double new_speed = (speed - 0.85) * 0.9;
            if (new_speed <= 0.01) new_speed = 0;
            new_speed += 0.85;
            xspeed = xspeed / speed * new_speed; yspeed = yspeed / speed * new_speed;
        }
    }
    double xpos_old = xpos, ypos_old = ypos;
    xpos += xspeed; ypos += yspeed;
    double time = sweep_circle_vs_tiles(c, xpos_old, ypos_old, xspeed, yspeed, 8 * 0.5);
    xpos = xpos_old + time * xspeed; ypos = ypos_old + time * yspeed;
    double xnormal = 0, ynormal = 0;
    bool bail = false;
    #pragma unroll 1
    for (int it = 0; it < 16; it++) {
        double a, b;
        int result = get_single_closest_point(c, xpos - 8, ypos - 8, xpos + 8, ypos + 8, xpos, ypos, 8, a, b);
        if (result == 0) break;
        double dx = xpos - a, dy = ypos - b;
        double dist = sqrt(SQ(dx) + SQ(dy));
        double depen_len = 8 - dist * result;
        if (depen_len < 0.0000001) break;
        if (dist == 0) { bail = true; break; }                 // `return` in the reference: skips bounce, repulsion, grid_move
        double xnorm = dx / dist, ynorm = dy / dist;
        xpos += xnorm * depen_len; ypos += ynorm * depen_len;
        xnormal += xnorm; ynormal += ynorm;
    }
    if (!bail) {
        double normal_len = sqrt(SQ(xnormal) + SQ(ynormal));
        if (normal_len > 0) {
            double dx = xnormal / normal_len, dy = ynormal / normal_len;
            double dot_product = xspeed * dx + yspeed * dy;
            if (dot_product < 0) {
                double speed = sqrt(SQ(xspeed) + SQ(yspeed));
                double bounce_strength = speed <= 1.35 ? 1 : 2;
                xspeed -= dx * dot_product * bounce_strength; yspeed -= dy * dot_product * bounce_strength;
            }
        }
        int index = (int)(__ldg(reinterpret_cast<const uint32_t*>(c.ents + id) + 6) >> 16);
        if ((double)(index + 1) < c.db_count) {
            // entity_dic[25][index+1:]: the later death balls (slots are in entity_dic order)
            #pragma unroll 1
            for (int t = slot + 1; t < c.n_dyn; t++) {
                if (((__ldg(c.dyn_ent + t) >> 16) & 255) != NPP_K_DEATH_BALL) continue;
                double dx = xpos - DYNF(c, t, 0), dy = ypos - DYNF(c, t, 1);
                double dist = sqrt(SQ(dx) + SQ(dy));
                if (dist < 16) {
                    dx = dx / dist * 4; dy = dy / dist * 4;
                    xspeed += dx; yspeed += dy;
                    DYNF(c, t, 2) -= dx; DYNF(c, t, 3) -= dy;
                }
            }
        }
    }
    DYNF(c, slot, 0) = xpos; DYNF(c, slot, 1) = ypos; DYNF(c, slot, 2) = xspeed; DYNF(c, slot, 3) = yspeed;
    if (!bail) grid_move(c, slot, xpos, ypos);
    return c.stamp;
}

// ---- move / think for the dynamic slots (entity_dic order; nsim.py:235-251) --------------------------------
#ifdef NPP_PROFILE
#define KIND_T0 long long _kt = clock64();
#define KIND_T1(ph, kind) { long long _k1 = clock64(); if ((threadIdx.x & 31) == 0) { atomicAdd(&g_kind_cycles[ph][(kind) & 31][0], (unsigned long long)(_k1 - _kt)); atomicAdd(&g_kind_cycles[ph][(kind) & 31][1], 1ULL); } }
#else
#define KIND_T0
#define KIND_T1(ph, kind)
#endif
DEV void dyn_move_phase(Ctx& c) {
    if (!HAS_ANY(NPP_MOVE_MASK)) return;
    #pragma unroll 1
    for (int s = 0; s < c.n_dyn; s++) {
        KIND_T0
        uint32_t de = __ldg(c.dyn_ent + s);
        int kind = (de >> 16) & 255, id = de & 0xffff, p = de >> 24;
        if (kind == NPP_K_DRONE_ZAP || kind == NPP_K_MINI_DRONE) c.stamp = drone_move(c, s, id, p, kind == NPP_K_MINI_DRONE);
        else if (kind == NPP_K_BOUNCE) {                         // entity_bounce_block.py:107-125
            EntS e = ld_ent(c, id);
            double xpos = DYNF(c, s, 0), ypos = DYNF(c, s, 1), xspeed = DYNF(c, s, 2), yspeed = DYNF(c, s, 3);
            xspeed *= 0.98; yspeed *= 0.98;
            xpos += xspeed; ypos += yspeed;
            double xforce = 0.02222222222222222 * (e.x - xpos), yforce = 0.02222222222222222 * (e.y - ypos);
            xpos += xforce; ypos += yforce; xspeed += xforce; yspeed += yforce;
            DYNF(c, s, 0) = xpos; DYNF(c, s, 1) = ypos; DYNF(c, s, 2) = xspeed; DYNF(c, s, 3) = yspeed;
            grid_move(c, s, xpos, ypos);
        } else if (kind == NPP_K_THWUMP) c.stamp = thwump_move(c, s, id, p);
        KIND_T1(0, kind)
    }
}
DEV void dyn_think_phase(const Nin& n, Ctx& c) {
    if (!HAS_ANY(NPP_THINK_MASK)) return;
    #pragma unroll 1
    for (int s = 0; s < c.n_dyn; s++) {
        KIND_T0
        uint32_t de = __ldg(c.dyn_ent + s);
        int kind = (de >> 16) & 255, id = de & 0xffff, p = de >> 24;
        if (kind == NPP_K_DOOR_REGULAR) {                         // entity_door_regular.py:46-53
            if (!(FLAG(c, id) & FL_CLOSED)) {
                int aux = DYNI(c, s, 2) + 1;                     // open_timer += 1 (low byte)
                DYNI(c, s, 2) = aux;
                if ((aux & 255) > 5) regular_door_change_state(c, id, s, true);
            }
        } else if (kind == NPP_K_THWUMP) thwump_think(c, n.f[F_XPOS], n.f[F_YPOS], ninja_valid_target(n), s, id, p);
        else if (kind == NPP_K_DEATH_BALL) c.stamp = death_ball_think(c, n.f[F_XPOS], n.f[F_YPOS], ninja_valid_target(n), s, id);
        else if (kind == NPP_K_SHWUMP) c.stamp = shwump_think(c, s, id);
        KIND_T1(1, kind)
    }
}

// ---- toggle mines (think) and boost pads (move): static entities ------------------------------------------
// Both only change when the ninja overlaps them or stops doing so.  Overlap needs a centre
// distance < 16 px, i.e. the entity's cell is within one cell of the ninja's, so the scan is
// limited to the 3x3 static cell lists around the ninja plus the (counted) "hot" entities --
// mines in state 2 and boost pads currently touched -- which must notice the ninja leaving.
DEV void mine_think(const Nin& n, Ctx& c, int id, bool valid) {     // entity_toggle_mine.py:90-118
    uint8_t fl = FLAG(c, id);
    int state = fl & FL_MINE;
    if (valid) {
        if (state == 1 || state == 2) {
            EntS e = ld_ent(c, id);
            bool ov = overlap_circle_vs_circle(e.x, e.y, mine_radius(state), n.f[F_XPOS], n.f[F_YPOS], NINJA_RADIUS);
            if (state == 1 && ov) { FLAG(c, id) = (fl & ~FL_MINE) | 2; c.hot++; }
            else if (state == 2 && !ov) { FLAG(c, id) = (fl & ~FL_MINE) | 0; c.hot--; }
        }
    } else if (state == 2 && n.i[I_STATE] == 6) { FLAG(c, id) = (fl & ~FL_MINE) | 1; c.hot--; }
}
DEV void boost_move(Nin& n, Ctx& c, int id, bool valid) {           // entity_boost_pad.py:46-64
    uint8_t fl = FLAG(c, id);
    if (!valid) { if (fl & FL_TOUCH) { FLAG(c, id) = fl & ~FL_TOUCH; c.hot--; } return; }
    EntS e = ld_ent(c, id);
    if (overlap_circle_vs_circle(e.x, e.y, 6, n.f[F_XPOS], n.f[F_YPOS], NINJA_RADIUS)) {
        if (!(fl & FL_TOUCH)) {
            double vel_norm = sqrt(SQ(n.f[F_XSPEED]) + SQ(n.f[F_YSPEED]));
            if (vel_norm > 0) {
                double x_boost = 2 * n.f[F_XSPEED] / vel_norm, y_boost = 2 * n.f[F_YSPEED] / vel_norm;
                n.f[F_XSPEED] += x_boost; n.f[F_YSPEED] += y_boost;
            }
            FLAG(c, id) = fl | FL_TOUCH; c.hot++;
        }
    } else if (fl & FL_TOUCH) { FLAG(c, id) = fl & ~FL_TOUCH; c.hot--; }
}
DEV void static_update_phase(Nin& n, Ctx& c) {
    bool valid = ninja_valid_target(n);
    // pass B first: hot entities anywhere on the map (they may be leaving contact)
    if (c.hot > 0) {
        int seen = 0, want = c.hot;
        #pragma unroll 1
        for (int id = 0; id < c.n_ent && seen < want; id++) {
            uint8_t fl = FLAG(c, id);
            if ((fl & FL_MINE) == 2) { seen++; mine_think(n, c, id, valid); }
            else if (HAS(NPP_K_BOOST) && (fl & FL_TOUCH)) { seen++; boost_move(n, c, id, valid); }
        }
    }
    if (!valid) return;
    double x = n.f[F_XPOS], y = n.f[F_YPOS];
    bool interior = (x >= 24.0 && x < 1032.0 && y >= 24.0 && y < 576.0);
    if (interior) {
        int cx = floordiv24(x), cy = floordiv24(y);      // 1..42, 1..23: the 3x3 block needs no clamping
        int nboost = 0;
        #pragma unroll 1
        for (int xc = cx - 1; xc <= cx + 1; xc++) {
            int a = __ldg(c.ecell_off + xc * NPP_NCY + cy - 1), b = __ldg(c.ecell_off + xc * NPP_NCY + cy + 2);
            #pragma unroll 1
            for (int k = a; k < b; k++) {
                int id = __ldg(c.ecell_ids + k);
                uint8_t fl = FLAG(c, id);
                if ((fl & FL_MINE) == 1) mine_think(n, c, id, true);               // only mines use these bits
                else if (HAS(NPP_K_BOOST) && (fl & (FL_IS_BOOST | FL_TOUCH)) == FL_IS_BOOST) nboost++;
            }
        }
        // boost pads change the ninja's velocity, so they run in entity_dic (= id) order
        int last = -1;
        #pragma unroll 1
        for (int r = 0; HAS(NPP_K_BOOST) && r < nboost; r++) {
            int best = 0x7fffffff;
            #pragma unroll 1
            for (int xc = cx - 1; xc <= cx + 1; xc++) {
                int a = __ldg(c.ecell_off + xc * NPP_NCY + cy - 1), b = __ldg(c.ecell_off + xc * NPP_NCY + cy + 2);
                #pragma unroll 1
                for (int k = a; k < b; k++) {
                    int id = __ldg(c.ecell_ids + k);
                    if (id > last && id < best && (FLAG(c, id) & (FL_IS_BOOST | FL_TOUCH)) == FL_IS_BOOST) best = id;
                }
            }
            if (best == 0x7fffffff) break;
            boost_move(n, c, best, true); last = best;
        }
    } else {
        // ninja in the border ring or outside the map: test everything (exact, never taken in practice)
        #pragma unroll 1
        for (int id = 0; id < c.n_ent; id++) {
            int kind = ent_kind(c, id);
            uint8_t fl = FLAG(c, id);
            if (kind == NPP_K_MINE && (fl & FL_MINE) == 1) mine_think(n, c, id, true);
            else if (HAS(NPP_K_BOOST) && kind == NPP_K_BOOST && !(fl & FL_TOUCH)) boost_move(n, c, id, true);
        }
    }
}

// ---- ordered 3x3 neighbourhood gather: physics.py:79-101 ------------------------------------------------------
// f(id) is called for every active entity of the nine cells, x-major over cells, list order inside a cell.
#define NPP_ROAM_REG 4        // roaming entities of one neighbourhood kept in registers; more take the generic merge
template <class F>
DEV void gather_entities(Ctx& c, double xpos, double ypos, F&& f) {
    int cx = clampi(floordiv24(xpos), 0, 43), cy = clampi(floordiv24(ypos), 0, 24);
    int x0 = max(cx - 1, 0), x1 = min(cx + 1, 43), y0 = max(cy - 1, 0), y1 = min(cy + 1, 24);
    // roaming entities inside the neighbourhood: (cell, stamp, id) of the first few in registers
    int nroam = 0;
    int rc0 = -1, rc1 = -1, rc2 = -1, rc3 = -1, rs0 = 0, rs1 = 0, rs2 = 0, rs3 = 0, ri0 = 0, ri1 = 0, ri2 = 0, ri3 = 0;
    #pragma unroll 1
    for (int s = 0; s < c.n_dyn; s++) {
        int cell = DYNI(c, s, 0);
        if (cell < 0) continue;
        int ex = cell / NPP_NCY, ey = cell - ex * NPP_NCY;
        if (ex < x0 || ex > x1 || ey < y0 || ey > y1) continue;
        if (nroam < NPP_ROAM_REG) {
            int st = DYNI(c, s, 1), id = (int)(__ldg(c.dyn_ent + s) & 0xffff);
            if (nroam == 0) { rc0 = cell; rs0 = st; ri0 = id; } else if (nroam == 1) { rc1 = cell; rs1 = st; ri1 = id; }
            else if (nroam == 2) { rc2 = cell; rs2 = st; ri2 = id; } else { rc3 = cell; rs3 = st; ri3 = id; }
        }
        nroam++;
    }
    if (nroam == 0) {
        #pragma unroll 1
        for (int xc = x0; xc <= x1; xc++) {
            int a = __ldg(c.ecell_off + xc * NPP_NCY + y0), b = __ldg(c.ecell_off + xc * NPP_NCY + y1 + 1);
            #pragma unroll 1
            for (int k = a; k < b; k++) { int id = __ldg(c.ecell_ids + k); if (FLAG(c, id) & FL_ACTIVE) f(id); }
        }
        return;
    }
    #pragma unroll 1
    for (int xc = x0; xc <= x1; xc++) {
        const int cbase = xc * NPP_NCY;
        // a column without roaming entities streams its three cells as one CSR range
        const bool col_has = (nroam > NPP_ROAM_REG) || (rc0 >= cbase + y0 && rc0 <= cbase + y1) || (rc1 >= cbase + y0 && rc1 <= cbase + y1) ||
                             (rc2 >= cbase + y0 && rc2 <= cbase + y1) || (rc3 >= cbase + y0 && rc3 <= cbase + y1);
        if (!col_has) {
            int a = __ldg(c.ecell_off + cbase + y0), b = __ldg(c.ecell_off + cbase + y1 + 1);
            #pragma unroll 1
            for (int k = a; k < b; k++) { int id = __ldg(c.ecell_ids + k); if (FLAG(c, id) & FL_ACTIVE) f(id); }
            continue;
        }
        #pragma unroll 1
        for (int yc = y0; yc <= y1; yc++) {
            int cell = cbase + yc;
            int ka = __ldg(c.ecell_off + cell), kb = __ldg(c.ecell_off + cell + 1);
            const bool cell_has = (nroam > NPP_ROAM_REG) || rc0 == cell || rc1 == cell || rc2 == cell || rc3 == cell;
            if (!cell_has) {
                #pragma unroll 1
                for (int k = ka; k < kb; k++) { int id = __ldg(c.ecell_ids + k); if (FLAG(c, id) & FL_ACTIVE) f(id); }
                continue;
            }
            int last = -1;
            #pragma unroll 1
            for (;;) {
                int sid = -1, sst = 0x7fffffff;
                if (ka < kb) {
                    sid = __ldg(c.ecell_ids + ka);
                    sst = c.order_mode ? sid : (int)(__ldg(reinterpret_cast<const uint32_t*>(c.ents + sid) + 7) & 0xffff);
                }
                int rid = -1, rst = 0x7fffffff;
                if (nroam <= NPP_ROAM_REG) {
                    if (rc0 == cell && rs0 > last && rs0 < rst) { rst = rs0; rid = ri0; }
                    if (rc1 == cell && rs1 > last && rs1 < rst) { rst = rs1; rid = ri1; }
                    if (rc2 == cell && rs2 > last && rs2 < rst) { rst = rs2; rid = ri2; }
                    if (rc3 == cell && rs3 > last && rs3 < rst) { rst = rs3; rid = ri3; }
                } else {
                    #pragma unroll 1
                    for (int s = 0; s < c.n_dyn; s++)
                        if (DYNI(c, s, 0) == cell) { int st = DYNI(c, s, 1); if (st > last && st < rst) { rst = st; rid = __ldg(c.dyn_ent + s) & 0xffff; } }
                }
                if (sid < 0 && rid < 0) break;
                int id;
                if (sst < rst) { id = sid; last = sst; ka++; } else { id = rid; last = rst; }
                if (FLAG(c, id) & FL_ACTIVE) f(id);
            }
        }
    }
}
// after a logical-collision pass: exit doors whose switch was taken join the grid now, so the
// pass that took the switch never sees them (the reference gathers its list before looping).
DEV void commit_pending_doors(Ctx& c) {
    if (!c.pending_doors) return;
    #pragma unroll 1
    for (int s = 0; s < c.n_dyn; s++) { int cell = DYNI(c, s, 0); if (cell <= -2) DYNI(c, s, 0) = -(cell + 2); }
    c.pending_doors = false;
}

// ---- physical collisions (ninja.py:224-267 callee side) ------------------------------------------------------
DEV bool is_physical_kind(int k) { return k == NPP_K_ONE_WAY || k == NPP_K_BOUNCE || k == NPP_K_THWUMP || k == NPP_K_SHWUMP; }
// entity_one_way_platform.py:79-103
DEV bool one_way_depen(const Nin& n, double ex, double ey, double normal_x, double normal_y, double& len) {
    double dx = n.f[F_XPOS] - ex, dy = n.f[F_YPOS] - ey;
    double lateral_dist = dy * normal_x - dx * normal_y;
    double direction = (n.f[F_YSPEED] * normal_x - n.f[F_XSPEED] * normal_y) * lateral_dist;
    double radius_scalar = direction < 0 ? 0.91 : 0.51;
    if (fabs(lateral_dist) < radius_scalar * NINJA_RADIUS + 12) {
        double normal_dist = dx * normal_x + dy * normal_y;
        if (0 < normal_dist && normal_dist <= NINJA_RADIUS) {
            double normal_proj = n.f[F_XSPEED] * normal_x + n.f[F_YSPEED] * normal_y;
            if (normal_proj <= 0) {
                double dx_old = n.f[F_XPOS_OLD] - ex, dy_old = n.f[F_YPOS_OLD] - ey;
                double normal_dist_old = dx_old * normal_x + dy_old * normal_y;
                if (NINJA_RADIUS - normal_dist_old <= 1.1) { len = NINJA_RADIUS - normal_dist; return true; }
            }
        }
    }
    return false;
}
// returns true with (nx, ny, len) when the entity pushes the ninja; `type` is entity.type
DEV bool ent_physical_collision(const Nin& n, Ctx& c, int id, const EntS& e, double& nx, double& ny, double& len) {
    double other;
    switch (e.kind) {
    case NPP_K_BOUNCE: {                                        // entity_bounce_block.py:127-145
        double bx = DYNF(c, e.dyn, 0), by = DYNF(c, e.dyn, 1), dl;
        if (!penetration_square_vs_point(bx, by, n.f[F_XPOS], n.f[F_YPOS], 9 + NINJA_RADIUS, nx, ny, dl, other)) return false;
        DYNF(c, e.dyn, 0) = bx - nx * dl * (1 - 0.2); DYNF(c, e.dyn, 1) = by - ny * dl * (1 - 0.2);
        DYNF(c, e.dyn, 2) -= nx * dl * (1 - 0.2); DYNF(c, e.dyn, 3) -= ny * dl * (1 - 0.2);
        len = dl * 0.2;
        return true; }
    case NPP_K_THWUMP:                                          // entity_thwump.py:208-213
        return penetration_square_vs_point(DYNF(c, e.dyn, 0), DYNF(c, e.dyn, 1), n.f[F_XPOS], n.f[F_YPOS], 9 + NINJA_RADIUS, nx, ny, len, other);
    case NPP_K_ONE_WAY: {
        double vx, vy; orientation_to_vector(e.p0, vx, vy);
        if (one_way_depen(n, e.x, e.y, vx, vy, len)) { nx = vx; ny = vy; return true; }
        return false; }
    case NPP_K_SHWUMP: {                                        // entity_shove_thwump.py:155-171
        int state = (FLAG(c, id) & FL_ST) >> FL_ST_SHIFT;
        if (state <= 1 && penetration_square_vs_point(DYNF(c, e.dyn, 0), DYNF(c, e.dyn, 1), n.f[F_XPOS], n.f[F_YPOS], 12 + NINJA_RADIUS, nx, ny, len, other))
            if (state == 0 || DYNF(c, e.dyn, 2) * nx + DYNF(c, e.dyn, 3) * ny >= 0.01) return true;
        return false; }
    default: return false;
    }
}

// ---- logical collisions (ninja.py:397-420 callee side) ----------------------------------------------------------
// Applies the entity's logical_collision(); launch-pad boosts are applied to the ninja here
// (ninja.py:401-418), any other truthy result is added to wall_normal (ninja.py:419-420).
DEV void ent_logical_collision(Nin& n, Ctx& c, int id, double& wall_normal) {
    EntS e = ld_ent(c, id);
    double nx, ny, pen, other;
    const double px = n.f[F_XPOS], py = n.f[F_YPOS];
    switch (e.kind) {
    case NPP_K_MINE: if (!HAS(NPP_K_MINE)) break; {                                          // entity_toggle_mine.py:120-128
        uint8_t fl = FLAG(c, id);
        if (ninja_valid_target(n) && (fl & FL_MINE) == 0 && overlap_circle_vs_circle(e.x, e.y, 4.0, px, py, NINJA_RADIUS)) {
            FLAG(c, id) = (fl & ~FL_MINE) | 1;
            ninja_kill(n, c, 0, 0, 0, 0, 1);
        }
        break; }
    case NPP_K_GOLD: if (!HAS(NPP_K_GOLD)) break; // entity_gold.py:63-72
        if (n.i[I_STATE] != 8 && overlap_circle_vs_circle(e.x, e.y, 6, px, py, NINJA_RADIUS)) { GI(c, I_GOLD) += 1; FLAG(c, id) &= ~FL_ACTIVE; }
        break;
    case NPP_K_EXIT: if (!HAS(NPP_K_EXIT)) break; // entity_exit.py:66-74
        if (overlap_circle_vs_circle(DYNF(c, e.dyn, 0), DYNF(c, e.dyn, 1), 12, px, py, NINJA_RADIUS)) ninja_win(n, c);
        break;
    case NPP_K_EXIT_SWITCH: if (!HAS(NPP_K_EXIT_SWITCH)) break; // entity_exit_switch.py:67-129
        if (overlap_circle_vs_circle(DYNF(c, e.dyn, 0), DYNF(c, e.dyn, 1), 6, px, py, NINJA_RADIUS)) {
            FLAG(c, id) &= ~FL_ACTIVE;
            EntS p = ld_ent(c, e.aux);
            // grid_entity[parent.cell].append(parent): next stamp; visible to gathers from the next pass on.
            // parent.cell is kept in the door slot's aux word while the door is outside the grid.
            // (a door already listed -- only after the curriculum re-armed a taken switch -- stays listed once)
            if (DYNI(c, p.dyn, 0) == -1) { DYNI(c, p.dyn, 0) = -(DYNI(c, p.dyn, 2) + 2); DYNI(c, p.dyn, 1) = c.stamp++; c.pending_doors = true; }
            FLAG(c, e.aux) |= FL_CLOSED;                         // parent.switch_hit = True
            c.events |= NPPSIM_EV_EXIT_SWITCH;
            if (c.sw_frame < 0) c.sw_frame = n.i[I_FRAME];
        }
        break;
    case NPP_K_DOOR_REGULAR: if (!HAS(NPP_K_DOOR_REGULAR)) break; // entity_door_regular.py:55-63
        if (overlap_circle_vs_circle(e.x, e.y, 10, px, py, NINJA_RADIUS)) {
            regular_door_change_state(c, id, e.dyn, false);
            DYNI(c, e.dyn, 2) &= ~255;                           // open_timer = 0
        }
        break;
    case NPP_K_DOOR_LOCKED: if (!HAS(NPP_K_DOOR_LOCKED)) break; // entity_door_locked.py:54-67
        if (overlap_circle_vs_circle(e.x, e.y, 5, px, py, NINJA_RADIUS)) {
            GI(c, I_DOORS_OPENED) += 1; door_set_closed(c, id, false); FLAG(c, id) &= ~FL_ACTIVE; c.events |= NPPSIM_EV_LOCKED_DOOR;
        }
        break;
    case NPP_K_DOOR_TRAP: if (!HAS(NPP_K_DOOR_TRAP)) break; // entity_door_trap.py:60-67
        if (overlap_circle_vs_circle(e.x, e.y, 5, px, py, NINJA_RADIUS)) { door_set_closed(c, id, true); FLAG(c, id) &= ~FL_ACTIVE; }
        break;
    case NPP_K_LAUNCH_PAD: if (!HAS(NPP_K_LAUNCH_PAD)) break; // entity_launch_pad.py:82-100
        if (ninja_valid_target(n) && overlap_circle_vs_circle(e.x, e.y, 6, px, py, NINJA_RADIUS)) {
            double vx, vy; orientation_to_vector(e.p0, vx, vy);
            if ((e.x - (px - NINJA_RADIUS * vx)) * vx + (e.y - (py - NINJA_RADIUS * vy)) * vy >= -0.1) {
                double yboost_scale = 1;
                if (vy < 0) yboost_scale = 1 - vy;
                double bx = vx * (36.0 / 7), by = vy * (36.0 / 7) * yboost_scale;
                double xboost = bx * 2 / 3, yboost = by * 2 / 3;                 // ninja.py:404-418
                n.f[F_XPOS] += xboost; n.f[F_YPOS] += yboost;
                n.f[F_XSPEED] = xboost; n.f[F_YSPEED] = yboost;
                n.i[I_FLOOR_COUNT] = 0; n.i[I_FLOOR_BUFFER] = -1;
                double boost_scalar = sqrt(SQ(xboost) + SQ(yboost));
                GF(c, F_XLP) = xboost / boost_scalar; GF(c, F_YLP) = yboost / boost_scalar;
                n.i[I_LP_BUFFER] = 0;
                if (n.i[I_STATE] == 3) n.f[F_GRAVITY] = GRAVITY_FALL;
                n.i[I_STATE] = 4;
            }
        }
        break;
    case NPP_K_ONE_WAY: if (!HAS(NPP_K_ONE_WAY)) break; {                                       // entity_one_way_platform.py:111-116
        double vx, vy, len; orientation_to_vector(e.p0, vx, vy);
        if (one_way_depen(n, e.x, e.y, vx, vy, len) && fabs(vx) == 1) wall_normal += vx;
        break; }
    case NPP_K_DRONE_ZAP: case NPP_K_MINI_DRONE: if (!HAS_ANY((1u << NPP_K_DRONE_ZAP) | (1u << NPP_K_MINI_DRONE))) break;                // entity_drone_zap.py:57-64, entity_mini_drone.py:60-67
        if (ninja_valid_target(n) && overlap_circle_vs_circle(DYNF(c, e.dyn, 0), DYNF(c, e.dyn, 1), e.kind == NPP_K_DRONE_ZAP ? 7.5 : 4.0, px, py, NINJA_RADIUS))
            ninja_kill(n, c, 0, 0, 0, 0, 0);
        break;
    case NPP_K_BOUNCE: if (!HAS(NPP_K_BOUNCE)) break; // entity_bounce_block.py:147-158
        if (penetration_square_vs_point(DYNF(c, e.dyn, 0), DYNF(c, e.dyn, 1), px, py, 9 + NINJA_RADIUS + 0.1, nx, ny, pen, other))
            if (nx != 0) wall_normal += nx;
        break;
    case NPP_K_THWUMP: if (!HAS(NPP_K_THWUMP)) break; // entity_thwump.py:215-243
        if (ninja_valid_target(n)) {
            double tx = DYNF(c, e.dyn, 0), ty = DYNF(c, e.dyn, 1);
            if (penetration_square_vs_point(tx, ty, px, py, 9 + NINJA_RADIUS + 0.1, nx, ny, pen, other)) {
                bool horizontal = e.p0 & 1; int direction = (e.p0 & 2) ? -1 : 1;
                double px1, py1, px2, py2;
                if (horizontal) { double dx = (9 + 2) * direction, dy = 9 - 2; px1 = tx + dx; py1 = ty - dy; px2 = tx + dx; py2 = ty + dy; }
                else { double dx = 9 - 2, dy = (9 + 2) * direction; px1 = tx - dx; py1 = ty + dy; px2 = tx + dx; py2 = ty + dy; }
                if (overlap_circle_vs_segment(px, py, NINJA_RADIUS + 2, px1, py1, px2, py2)) ninja_kill(n, c, 0, 0, 0, 0, 0);
                if (nx != 0) wall_normal += nx;
            }
        }
        break;
    case NPP_K_DEATH_BALL: if (!HAS(NPP_K_DEATH_BALL)) break; // entity_death_ball.py:169-181
        if (ninja_valid_target(n)) {
            double bx = DYNF(c, e.dyn, 0), by = DYNF(c, e.dyn, 1);
            if (overlap_circle_vs_circle(bx, by, 5, px, py, NINJA_RADIUS)) {
                double dx = bx - px, dy = by - py;
                double dist = sqrt(SQ(dx) + SQ(dy));
                DYNF(c, e.dyn, 2) += dx / dist * 10; DYNF(c, e.dyn, 3) += dy / dist * 10;
                ninja_kill(n, c, 0, 0, 0, 0, 0);
            }
        }
        break;
    case NPP_K_SHWUMP: if (!HAS(NPP_K_SHWUMP)) break; {                                        // entity_shove_thwump.py:173-202
        double sx = DYNF(c, e.dyn, 0), sy = DYNF(c, e.dyn, 1);
        uint8_t fl = FLAG(c, id);
        int state = (fl & FL_ST) >> FL_ST_SHIFT;
        bool hit = penetration_square_vs_point(sx, sy, px, py, 12 + NINJA_RADIUS + 0.1, nx, ny, pen, other);
        if (hit && state <= 1) {
            if (state == 0) {
                fl |= FL_ACTIVATED;
                if (other > 0.2) { DYNF(c, e.dyn, 2) = nx; DYNF(c, e.dyn, 3) = ny; fl = (fl & ~FL_ST) | (1 << FL_ST_SHIFT); }
                FLAG(c, id) = fl;
            } else {
                if (DYNF(c, e.dyn, 2) * nx + DYNF(c, e.dyn, 3) * ny >= 0.01) FLAG(c, id) = fl | FL_ACTIVATED;
                else break;                                      // `return` (None)
            }
            if (nx != 0) wall_normal += nx;
            break;
        }
        if (overlap_circle_vs_circle(px, py, NINJA_RADIUS, sx, sy, 8)) ninja_kill(n, c, 0, 0, 0, 0, 0);
        break; }
    default: break;
    }
}
