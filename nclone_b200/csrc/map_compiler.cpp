// map_compiler.cpp -- host side: one level's map_data -> one blob (map_blob.h).
//
// Replaces, for the batched engine, what Simulator.load does once per level in the
// reference (paths under /root/reference/nclone/):
//   MapLoader.load_map_tiles                        map_loader.py:18-41
//   TileSegmentFactory.create_segments_for_simulator utils/tile_segment_factory.py:264-344
//   SpatialSegmentIndex._build_index                utils/spatial_segment_index.py:69-110
//   Simulator.reset_map_tile_data                   nsim.py:194-219
//   MapLoader.load_map_entities + entity ctors      map_loader.py:69-145, utils/entity_factory.py:145-233
// Integer-only work; every ordering the tick path can observe is preserved:
// segment order inside a cell (own diagonal/arc, then horizontal edges in x-major
// half-cell order, then vertical ones) and entity order (entity_dic order for
// move/think, creation order inside grid cells).
#include <cmath>
#include <cstring>
#include <vector>
#include <cstdio>

#include "map_blob.h"
#include "tile_tables.h"
#include "../../include/nppsim.h"

extern thread_local char g_nppsim_err[256];       // nppsim_kernels.cu: the buffer behind nppsim_last_error()
namespace {

struct Box { int x0, y0, x1, y1; };
inline Box seg_box(uint32_t s) {   // entities.py:36-41, 121-125
    Box b;
    if (s & NPP_SEG_ARC) {
        int cx = s & 127, cy = (s >> 7) & 63;
        int px = cx + (((s >> 13) & 1) ? -2 : 2), py = cy + (((s >> 14) & 1) ? -2 : 2);
        b.x0 = cx < px ? cx : px; b.x1 = cx > px ? cx : px; b.y0 = cy < py ? cy : py; b.y1 = cy > py ? cy : py;
    } else {
        int x1 = s & 127, y1 = (s >> 7) & 63, x2 = (s >> 13) & 127, y2 = (s >> 20) & 63;
        b.x0 = x1 < x2 ? x1 : x2; b.x1 = x1 > x2 ? x1 : x2; b.y0 = y1 < y2 ? y1 : y2; b.y1 = y1 > y2 ? y1 : y2;
    }
    return b;
}
// physics.py:317-332: orientation dict has keys 0..7 only
inline bool orientation_vec(double o, double* vx, double* vy) {
    if (!is_int(o) || o < 0 || o > 7) return false;
    const double d = std::sqrt(2.0) / 2;
    const double t[8][2] = {{1,0},{d,d},{0,1},{-d,d},{-1,0},{-d,-d},{0,-1},{d,-d}};
    *vx = t[(int)o][0]; *vy = t[(int)o][1];
    return true;
}

struct Draft {               // entity in creation order
    NppStaticEnt se; int dic_key; bool in_grid; bool roaming;
    NppDoor door; bool has_door;
};

char (&g_err)[256] = g_nppsim_err;
int fail(int code, const char* msg) { std::snprintf(g_err, sizeof(g_err), "%s", msg); return code; }

}  // namespace


extern "C" int64_t nppsim_map_blob_bound(int32_t map_len) {
    int64_t max_ent = map_len > 1230 ? (map_len - 1230) / 5 * 2 + 2 : 2;
    return (int64_t)sizeof(NppMapHeader) + NPP_NCELL * 4 + NPP_NCELL * NPP_SEGS_PER_CELL * 4 + 2 * NPP_EDGE_WORDS * 4
         + max_ent * (int64_t)(sizeof(NppStaticEnt) + 2 * 2 + 4 + sizeof(NppDoor)) + (NPP_NCELL + 1) * 2 + NPP_NCELL + 16 * 12;
}

extern "C" int nppsim_compile_map(const double* m, int32_t len, uint8_t* out, int64_t cap, int64_t* used,
                                  int32_t* n_ent_out, int32_t* n_dyn_out) {
    if (!m || !out || len < 1235) return fail(NPPSIM_E_MAP, "map_data shorter than 1235 values");
    // ---------------------------------------------------------------- tiles
    static thread_local int tile[NPP_NCX][NPP_NCY];
    for (int x = 0; x < NPP_NCX; x++) for (int y = 0; y < NPP_NCY; y++) tile[x][y] = 0;
    for (int x = 0; x < 42; x++) for (int y = 0; y < 23; y++) {
        double v = m[184 + x + 42 * y];
        // `tile_id in TILE_GRID_EDGE_MAP` is False for anything but the integers 0..37, and 34..37
        // return early (tile_segment_factory.py:184-191,278-283): all of those behave as empty.
        tile[x + 1][y + 1] = (is_int(v) && v >= 1 && v <= 33) ? (int)v : 0;
    }
    for (int x = 0; x < NPP_NCX; x++) { tile[x][0] = 1; tile[x][NPP_NCY - 1] = 1; }
    for (int y = 0; y < NPP_NCY; y++) { tile[0][y] = 1; tile[NPP_NCX - 1][y] = 1; }

    static thread_local int8_t hsum[NPP_NHX][NPP_NHY], vsum[NPP_NHX][NPP_NHY];
    static thread_local uint8_t hedge[NPP_NHX][NPP_NHY], vedge[NPP_NHX][NPP_NHY];
    for (int x = 0; x < NPP_NHX; x++) for (int y = 0; y < NPP_NHY; y++) {
        hsum[x][y] = vsum[x][y] = 0;
        hedge[x][y] = (y == 0 || y == NPP_NHY - 1) ? 1 : 0;     // nsim.py:208-215
        vedge[x][y] = (x == 0 || x == NPP_NHX - 1) ? 1 : 0;
    }
    std::vector<uint32_t> cell_segs[NPP_NCELL];
    for (int x = 0; x < NPP_NCX; x++) for (int y = 0; y < NPP_NCY; y++) {
        int t = tile[x][y];
        if (!t) continue;
        for (int r = 0; r < 3; r++) for (int c = 0; c < 2; c++) {            // horizontal halves
            hedge[2 * x + c][2 * y + r] ^= (uint8_t)kEdge[t][2 * r + c];        // (v + e) % 2
            hsum[2 * x + c][2 * y + r] += kOrtho[t][2 * r + c];
        }
        for (int c = 0; c < 3; c++) for (int r = 0; r < 2; r++) {            // vertical halves
            vedge[2 * x + c][2 * y + r] ^= (uint8_t)kEdge[t][2 * c + r + 6];
            vsum[2 * x + c][2 * y + r] += kOrtho[t][2 * c + r + 6];
        }
        const Slanted& s = kSlanted[t];
        if (s.kind == 1) cell_segs[x * NPP_NCY + y].push_back(pack_linear(2 * x + s.a, 2 * y + s.b, 2 * x + s.c, 2 * y + s.d));
        else if (s.kind == 2) cell_segs[x * NPP_NCY + y].push_back(pack_arc(2 * x + s.a, 2 * y + s.b, s.c, s.d, s.convex));
    }
    // tile_segment_factory.py:231-262.  state s != 0 -> one segment in cell floor((coord - 0.1*s)/2);
    // the end points are swapped only for s == -1.
    for (int x = 0; x < NPP_NHX; x++) for (int y = 0; y < NPP_NHY; y++) {
        int s = hsum[x][y];
        if (!s) continue;
        int cx = x >> 1, cy = (int)std::floor((y - 0.1 * s) / 2);
        if (cx >= NPP_NCX || cy < 0 || cy >= NPP_NCY) return fail(NPPSIM_E_MAP, "horizontal edge outside the cell grid");
        cell_segs[cx * NPP_NCY + cy].push_back(s == -1 ? pack_linear(x + 1, y, x, y) : pack_linear(x, y, x + 1, y));
    }
    for (int x = 0; x < NPP_NHX; x++) for (int y = 0; y < NPP_NHY; y++) {
        int s = vsum[x][y];
        if (!s) continue;
        int cx = (int)std::floor((x - 0.1 * s) / 2), cy = y >> 1;
        if (cx < 0 || cx >= NPP_NCX || cy >= NPP_NCY) return fail(NPPSIM_E_MAP, "vertical edge outside the cell grid");
        cell_segs[cx * NPP_NCY + cy].push_back(s == -1 ? pack_linear(x, y, x, y + 1) : pack_linear(x, y + 1, x, y));
    }

    // ---------------------------------------------------------------- entities
    std::vector<Draft> drafts;
    int counts[40] = {0};
    double dexit = m[1156];
    if (!is_int(dexit)) return fail(NPPSIM_E_MAP, "map_data[1156] (exit door count) is not an integer");
    auto make = [&](int kind, int type, int key, double xc, double yc) -> Draft& {
        Draft d; std::memset(&d, 0, sizeof(d));
        d.se.kind = (uint8_t)kind; d.se.type = (uint8_t)type; d.dic_key = key;
        d.se.x = xc * 6; d.se.y = yc * 6;
        d.se.index = (uint16_t)counts[type]++;
        d.se.cell = (uint16_t)cell_of(d.se.x, d.se.y);
        d.se.dyn = -1; d.se.aux = -1; d.in_grid = true; d.roaming = false; d.has_door = false;
        d.se.load_order = (uint16_t)drafts.size();
        drafts.push_back(d);
        return drafts.back();
    };
    auto make_door = [&](Draft& d, double orientation, double swx, double swy) -> int {   // entity_door_base.py:52-97
        double vx, vy;
        if (!orientation_vec(orientation, &vx, &vy)) return fail(NPPSIM_E_MAP, "door orientation not in 0..7");
        bool vertical = (orientation == 0 || orientation == 4);
        int dcx = cell_axis(d.se.x - 12 * vx, 43);
        int dcy = cell_axis(d.se.y - 12 * vy, 24);
        int hx = 2 * (dcx + 1), hy = 2 * (dcy + 1);
        d.has_door = true; d.door.vertical = vertical ? 1 : 0;
        if (vertical) { d.door.x0 = (uint8_t)hx; d.door.y0 = (uint8_t)(hy - 2); d.door.x1 = (uint8_t)hx; d.door.y1 = (uint8_t)(hy - 1); }
        else { d.door.x0 = (uint8_t)(hx - 2); d.door.y0 = (uint8_t)hy; d.door.x1 = (uint8_t)(hx - 1); d.door.y1 = (uint8_t)hy; }
        if (d.door.x0 >= NPP_NHX || d.door.x1 >= NPP_NHX || d.door.y0 >= NPP_NHY || d.door.y1 >= NPP_NHY)
            return fail(NPPSIM_E_MAP, "door grid edge outside the half-cell grid");
        d.se.p0 = vertical ? 1 : 0;
        d.se.x = 6 * swx; d.se.y = 6 * swy;                   // entity moves to its switch
        d.se.cell = (uint16_t)cell_of(d.se.x, d.se.y);
        return 0;
    };
    int index = 1230;
    while (index < len) {                                      // map_loader.py:88-141
        if (index + 4 >= len) break;
        double type = m[index], xc = m[index + 1], yc = m[index + 2], orientation = m[index + 3], mode = m[index + 4];
        int t = (is_int(type) && type >= 0 && type < 40) ? (int)type : -1;
        switch (t) {
        case 1: case 21: { Draft& d = make(NPP_K_MINE, t, t, xc, yc); d.se.p0 = (t == 1) ? 0 : 1; break; }   // entity_factory.py:181-182,230-231
        case 2: make(NPP_K_GOLD, 2, 2, xc, yc); break;
        case 3: {                                               // entity_factory.py:185-197
            Draft& p = make(NPP_K_EXIT, 3, 3, xc, yc);
            p.in_grid = false; p.roaming = true;
            int parent_pos = (int)drafts.size() - 1;
            double cd = (double)index + 5.0 * dexit + 1;         // in the double domain: no overflow on garbage counts
            if (!(cd >= 0 && cd + 1 < len)) return fail(NPPSIM_E_MAP, "exit switch coordinates out of range");
            long ci = (long)cd;
            Draft& s = make(NPP_K_EXIT_SWITCH, 4, 3, m[ci], m[ci + 1]);
            s.roaming = true; s.se.aux = (int16_t)parent_pos;   // creation position; remapped to entity id below
            break; }
        case 5: { Draft& d = make(NPP_K_DOOR_REGULAR, 5, 5, xc, yc); int rc = make_door(d, orientation, xc, yc); if (rc) return rc; break; }
        case 6: case 8: {
            if (index + 7 >= len) return fail(NPPSIM_E_MAP, "door switch coordinates out of range");
            Draft& d = make(t == 6 ? NPP_K_DOOR_LOCKED : NPP_K_DOOR_TRAP, t, t, xc, yc);
            int rc = make_door(d, orientation, m[index + 6], m[index + 7]); if (rc) return rc; break; }
        case 10: case 11: {
            double vx, vy;
            if (!orientation_vec(orientation, &vx, &vy)) return fail(NPPSIM_E_MAP, "launch pad / one-way orientation not in 0..7");
            Draft& d = make(t == 10 ? NPP_K_LAUNCH_PAD : NPP_K_ONE_WAY, t, t, xc, yc); d.se.p0 = (uint8_t)(int)orientation; break; }
        case 14: case 26: {                                     // entity_drone_base.py:66-78
            double dir = std::floor(orientation / 2);
            if (!(dir >= 0 && dir <= 3) || !(mode == 0 || mode == 1 || mode == 2 || mode == 3))
                return fail(NPPSIM_E_MAP, "drone orientation/mode out of range");
            Draft& d = make(t == 14 ? NPP_K_DRONE_ZAP : NPP_K_MINI_DRONE, t, t, xc, yc);
            d.se.p0 = (uint8_t)(int)dir; d.se.p1 = (uint8_t)(int)mode; d.roaming = true; break; }
        case 17: { Draft& d = make(NPP_K_BOUNCE, 17, 17, xc, yc); d.roaming = true; break; }
        case 20: { Draft& d = make(NPP_K_THWUMP, 20, 20, xc, yc); d.roaming = true;   // entity_thwump.py:92-101
            d.se.p0 = (uint8_t)(((orientation == 0 || orientation == 4) ? 1 : 0) | ((orientation == 0 || orientation == 2) ? 0 : 2)); break; }
        case 24: make(NPP_K_BOOST, 24, 24, xc, yc); break;
        case 25: { Draft& d = make(NPP_K_DEATH_BALL, 25, 25, xc, yc); d.roaming = true; break; }
        case 28: { Draft& d = make(NPP_K_SHWUMP, 28, 28, xc, yc); d.roaming = true; break; }
        default: break;                                         // unknown types are skipped (entity_factory.py:69,176-177)
        }
        if (t == 6 || t == 8) {                                 // map_loader.py:126-141
            if (index + 9 < len) index += (m[index + 7] != 0 && m[index + 8] == 0 && m[index + 9] == 0) ? 10 : 9;
            else index += 9;
        } else index += 5;
        // NppStaticEnt.dyn / .aux are int16 and the kernels index ent_flag with 32-bit (id * stride): 32767 entities is the format's limit
        if (drafts.size() > 32767) return fail(NPPSIM_E_CAPACITY, "more than 32767 entities");
    }
    // entity_dic order: keys 1..28 ascending, creation order inside a key (nsim.py:190-192)
    std::vector<int> order;
    for (int key = 1; key <= 28; key++) for (size_t k = 0; k < drafts.size(); k++) if (drafts[k].dic_key == key) order.push_back((int)k);
    std::vector<int> id_of(drafts.size(), -1);
    for (size_t i = 0; i < order.size(); i++) id_of[order[i]] = (int)i;
    std::vector<NppStaticEnt> ents(order.size());
    std::vector<NppDoor> doors;
    std::vector<uint32_t> dyn_ent;
    bool reads_edges = false;
    for (size_t i = 0; i < order.size(); i++) {
        Draft& d = drafts[order[i]];
        NppStaticEnt se = d.se;
        if (se.kind == NPP_K_EXIT_SWITCH) se.aux = (int16_t)id_of[se.aux];
        if (d.has_door) { NppDoor dr = d.door; dr.ent = (uint16_t)i; dr.pad = 0; se.aux = (int16_t)doors.size(); doors.push_back(dr); }
        bool needs_slot = d.roaming || se.kind == NPP_K_DOOR_REGULAR;
        if (needs_slot) {
            se.dyn = (int16_t)dyn_ent.size();
            uint32_t p = (se.kind == NPP_K_DRONE_ZAP || se.kind == NPP_K_MINI_DRONE) ? se.p1 : se.p0;
            dyn_ent.push_back((uint32_t)i | ((uint32_t)se.kind << 16) | (p << 24));
        }
        if (se.kind == NPP_K_DRONE_ZAP || se.kind == NPP_K_MINI_DRONE || se.kind == NPP_K_THWUMP || se.kind == NPP_K_SHWUMP) reads_edges = true;
        ents[i] = se;
    }
    // static per-cell lists (entities that never change cell and are always in the grid), two orders:
    // creation order (Simulator.load / reset) and entity_dic order (fast_reset, nsim.py:126-140)
    std::vector<uint16_t> ecell_off(NPP_NCELL + 1, 0), ids_load, ids_dic;
    {
        std::vector<std::vector<int>> per(NPP_NCELL);
        for (size_t k = 0; k < drafts.size(); k++) if (!drafts[k].roaming) per[drafts[k].se.cell].push_back(id_of[k]);  // creation order
        for (int c = 0; c < NPP_NCELL; c++) {
            ecell_off[c] = (uint16_t)ids_load.size();
            for (int id : per[c]) ids_load.push_back((uint16_t)id);
            std::vector<int> s = per[c];
            for (size_t a = 1; a < s.size(); a++) { int v = s[a]; size_t b = a; while (b > 0 && s[b - 1] > v) { s[b] = s[b - 1]; b--; } s[b] = v; }
            for (int id : s) ids_dic.push_back((uint16_t)id);
        }
        ecell_off[NPP_NCELL] = (uint16_t)ids_load.size();
    }

    // ---------------------------------------------------------------- emit
    auto align16 = [](int64_t v) { return (v + 15) & ~(int64_t)15; };
    NppMapHeader h; std::memset(&h, 0, sizeof(h));
    int64_t off = align16(sizeof(NppMapHeader));
    h.off_cell_hdr = (uint32_t)off; off = align16(off + NPP_NCELL * 4);
    h.off_segs = (uint32_t)off; off = align16(off + NPP_NCELL * NPP_SEGS_PER_CELL * 4);
    h.off_hedge = (uint32_t)off; off = align16(off + NPP_EDGE_WORDS * 4);
    h.off_vedge = (uint32_t)off; off = align16(off + NPP_EDGE_WORDS * 4);
    h.off_ents = (uint32_t)off; off = align16(off + (int64_t)ents.size() * sizeof(NppStaticEnt));
    h.off_ecell_off = (uint32_t)off; off = align16(off + (NPP_NCELL + 1) * 2);
    h.off_ecell_ids = (uint32_t)off; off = align16(off + (int64_t)ids_load.size() * 2);
    h.off_ecell_ids_dic = (uint32_t)off; off = align16(off + (int64_t)ids_dic.size() * 2);
    h.off_dyn_ent = (uint32_t)off; off = align16(off + (int64_t)dyn_ent.size() * 4);
    h.off_doors = (uint32_t)off; off = align16(off + (int64_t)doors.size() * sizeof(NppDoor));
    h.off_tiles = (uint32_t)off; off = align16(off + NPP_NCELL);
    if (off > cap) return fail(NPPSIM_E_CAPACITY, "output buffer too small for the compiled map");
    std::memset(out, 0, (size_t)off);
    h.magic = NPP_BLOB_MAGIC; h.total_bytes = (uint32_t)off;
    h.n_ent = (uint32_t)ents.size(); h.n_dyn = (uint32_t)dyn_ent.size(); h.n_doors = (uint32_t)doors.size();
    h.n_static_cell = (uint32_t)ids_load.size(); h.flags = reads_edges ? 1u : 0u;
    h.kind_mask = 1u << NPP_K_NONE;              // never 0: 0 means "census unknown" to the launcher
    for (size_t k = 0; k < ents.size(); k++) h.kind_mask |= 1u << ents[k].kind;
    h.spawn_x = m[1231] * 6; h.spawn_y = m[1232] * 6; h.db_count = m[1200];
    uint32_t* cell_hdr = (uint32_t*)(out + h.off_cell_hdr);
    uint32_t* segs = (uint32_t*)(out + h.off_segs);
    uint32_t nseg = 0;
    for (int c = 0; c < NPP_NCELL; c++) {
        const std::vector<uint32_t>& v = cell_segs[c];
        if (v.empty()) continue;
        if (v.size() > NPP_SEGS_PER_CELL) return fail(NPPSIM_E_CAPACITY, "more than 8 segments in one cell");
        Box bb = {1 << 20, 1 << 20, -1, -1};
        for (size_t k = 0; k < v.size(); k++) {
            segs[c * NPP_SEGS_PER_CELL + k] = v[k];
            Box b = seg_box(v[k]);
            if (b.x0 < bb.x0) bb.x0 = b.x0; if (b.y0 < bb.y0) bb.y0 = b.y0;
            if (b.x1 > bb.x1) bb.x1 = b.x1; if (b.y1 > bb.y1) bb.y1 = b.y1;
        }
        cell_hdr[c] = (uint32_t)v.size() | ((uint32_t)bb.x0 << 4) | ((uint32_t)bb.y0 << 11) | ((uint32_t)bb.x1 << 17) | ((uint32_t)bb.y1 << 24);
        nseg += (uint32_t)v.size();
    }
    h.n_segments = nseg;
    uint32_t* hb = (uint32_t*)(out + h.off_hedge); uint32_t* vb = (uint32_t*)(out + h.off_vedge);
    for (int x = 0; x < NPP_NHX; x++) for (int y = 0; y < NPP_NHY; y++) {
        int bit = x * NPP_NHY + y;
        if (hedge[x][y]) hb[bit >> 5] |= 1u << (bit & 31);
        if (vedge[x][y]) vb[bit >> 5] |= 1u << (bit & 31);
    }
    if (!ents.empty()) std::memcpy(out + h.off_ents, ents.data(), ents.size() * sizeof(NppStaticEnt));
    std::memcpy(out + h.off_ecell_off, ecell_off.data(), ecell_off.size() * 2);
    if (!ids_load.empty()) { std::memcpy(out + h.off_ecell_ids, ids_load.data(), ids_load.size() * 2); std::memcpy(out + h.off_ecell_ids_dic, ids_dic.data(), ids_dic.size() * 2); }
    if (!dyn_ent.empty()) std::memcpy(out + h.off_dyn_ent, dyn_ent.data(), dyn_ent.size() * 4);
    if (!doors.empty()) std::memcpy(out + h.off_doors, doors.data(), doors.size() * sizeof(NppDoor));
    uint8_t* tl = out + h.off_tiles;
    for (int x = 0; x < NPP_NCX; x++) for (int y = 0; y < NPP_NCY; y++) tl[x * NPP_NCY + y] = (uint8_t)tile[x][y];
    std::memcpy(out, &h, sizeof(h));
    if (used) *used = off;
    if (n_ent_out) *n_ent_out = (int32_t)ents.size();
    if (n_dyn_out) *n_dyn_out = (int32_t)dyn_ent.size();
    g_err[0] = 0;
    return 0;
}
