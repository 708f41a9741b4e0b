// map_blob.h -- layout of one compiled map ("blob") in HBM, shared by the host map
// compiler (map_compiler.cpp) and the kernels (nppsim_kernels.cu).
//
// A blob is everything about a level that never changes while envs run on it: the
// 44x25 tile-segment grid the ninja collides with, the two 89x51 half-cell edge bitmaps
// drones/thwumps navigate by, and the entity table.  Envs that share a level share a blob
// (BatchedSimulator dedups by content), so per-env HBM holds only mutable state.
//
// Reference structures this replaces (paths under /root/reference/nclone/):
//   sim.segment_dic + SpatialSegmentIndex.cell_data   utils/spatial_segment_index.py:69-110
//   sim.hor_grid_edge_dic / ver_grid_edge_dic          nsim.py:206-215
//   sim.entity_dic / sim.grid_entity (initial)         map_loader.py:69-145
#pragma once
#include <stdint.h>

#define NPP_NCX 44
#define NPP_NCY 25
#define NPP_NCELL (NPP_NCX * NPP_NCY)
#define NPP_NHX 89
#define NPP_NHY 51
#define NPP_EDGE_WORDS ((NPP_NHX * NPP_NHY + 31) / 32)   // 142
#define NPP_SEGS_PER_CELL 8                              // provable max: a tile owns <= 8 ortho halves or <= 5 + 1 diagonal

#define NPP_BLOB_MAGIC 0x4E505042u   // "NPPB"

// Entity kinds (one per reference class that utils/entity_factory.py:50-67 instantiates)
enum NppKind {
    NPP_K_NONE = 0, NPP_K_MINE = 1, NPP_K_GOLD = 2, NPP_K_EXIT = 3, NPP_K_EXIT_SWITCH = 4,
    NPP_K_DOOR_REGULAR = 5, NPP_K_DOOR_LOCKED = 6, NPP_K_DOOR_TRAP = 7, NPP_K_LAUNCH_PAD = 8,
    NPP_K_ONE_WAY = 9, NPP_K_DRONE_ZAP = 10, NPP_K_BOUNCE = 11, NPP_K_THWUMP = 12, NPP_K_BOOST = 13,
    NPP_K_DEATH_BALL = 14, NPP_K_MINI_DRONE = 15, NPP_K_SHWUMP = 16
};

// Cell header (u32): bits 0-3 segment count (0 = key absent from cell_data), then the
// cell's AABB over its segments in units of 12 px: x0:7 | y0:6 | x1:7 | y1:6.
// Segment (u32), coordinates in units of 12 px (every tile vertex is a multiple of 12):
//   bit 31 = 0 linear : x1:7 | y1:6 | x2:7 | y2:6          (oriented, entities.py:12-41)
//   bit 31 = 1 arc    : cx:7 | cy:6 | hor<0:1 | ver<0:1 | convex:1   (radius 24, entities.py:102-125)
#define NPP_SEG_ARC 0x80000000u

// Static entity record, 32 bytes, in entity_dic order (type-major; nsim.py:237-243).
typedef struct {
    double x, y;          // initial xpos,ypos (doors: the switch position, entity_door_base.py:94-96)
    uint8_t kind;         // NppKind
    uint8_t type;         // reference entity.type (4 for the exit switch)
    uint8_t p0;           // mine: initial state | drone: initial dir | thwump: bit0 horizontal, bit1 direction<0
                          // launch pad / one-way: orientation 0..7 | door: bit0 vertical
    uint8_t p1;           // drone: mode
    int16_t dyn;          // dynamic slot or -1
    int16_t aux;          // exit switch: parent entity id | doors: door record id
    uint16_t cell;        // initial cell cx*25+cy
    uint16_t index;       // entity.index (per-type counter, entities.py:213-216)
    uint16_t load_order;  // position in creation order == initial stamp inside grid_entity lists
    uint16_t pad;
} NppStaticEnt;

// Door record: the two half-cell edges a door adds to (entity_door_base.py:70-89)
typedef struct {
    uint16_t ent;         // entity id
    uint8_t vertical;     // 1: ver_grid_edge_dic, 0: hor_grid_edge_dic
    uint8_t x0, y0, x1, y1;
    uint8_t pad;
} NppDoor;

typedef struct {
    uint32_t magic, total_bytes;
    uint32_t n_ent;        // entities (entity_dic order)
    uint32_t n_dyn;        // dynamic slots
    uint32_t n_doors;
    uint32_t n_static_cell;   // entries of ecell_ids
    uint32_t n_segments;
    uint32_t flags;        // bit0: has entities that read grid edges
    double spawn_x, spawn_y;  // map_data[1231]*6, map_data[1232]*6   (ninja.py:88-89)
    double db_count;          // map_data[1200]  (entity_death_ball.py:149)
    // byte offsets from the blob start (16-byte aligned)
    uint32_t off_cell_hdr;    // u32[1100]           index cx*25+cy
    uint32_t off_segs;        // u32[1100*8]
    uint32_t off_hedge;       // u32[142] bit (x*51+y) of hor_grid_edge_dic (tile part, 0/1)
    uint32_t off_vedge;       // u32[142]
    uint32_t off_ents;        // NppStaticEnt[n_ent]
    uint32_t off_ecell_off;   // u16[1101] CSR over cells: static (never-moving, always-in-grid) entities
    uint32_t off_ecell_ids;   // u16[n_static_cell] entity ids, creation order inside a cell (load / reset)
    uint32_t off_ecell_ids_dic; // same cells, entity_dic order inside a cell (after fast_reset, nsim.py:126-140)
    uint32_t off_dyn_ent;     // u32[n_dyn], slots in entity_dic order: entity id | kind << 16 | (drone: mode, else p0) << 24
    uint32_t off_doors;       // NppDoor[n_doors]
    uint32_t off_tiles;       // u8[1100] tile ids (cx*25+cy), for tile_dic readers
    uint32_t kind_mask;       // OR of (1 << NppKind) over the level's entities: picks the step-kernel variant
} NppMapHeader;
