// nppsim_kernels.cu -- sm_100a kernels + C ABI launchers (include/nppsim.h).
// Device code lives in geom.cuh / entities.cuh / ninja.cuh; see geom.cuh for the design notes.
#include "pool.cuh"

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>

std::atomic<long long> g_nppsim_launches{0};     // shared with map_compile_device.cu

thread_local char g_nppsim_err[256] = "";         // nppsim_last_error(): one buffer for all three translation units

namespace {

char (&g_err)[256] = g_nppsim_err;
std::atomic<long long>& g_launches = g_nppsim_launches;

int fail(int code, const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg); return code; }
int cuda_fail(cudaError_t e, const char* what) {
    snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    return (int)e > 0 ? (int)e : 1;
}

// ---- reset -------------------------------------------------------------------------------------------
// mode 0: state right after Simulator.reset() on the Simulator that has held the level so far (nsim.py:62-76; Ninja.__init__
//         ninja.py:80-196; entity ctors) -- what NPlayHeadless.reset and ActionReplayer call; the entities are created again and
//         their `index` goes on counting (entities.py:213-216), which env_i[NPPSIM_ENV_RELOADS] records
// mode 2: the same state on a FRESH Simulator (Simulator() + load(), nsim.py:51-56): indices start at 0 again
// mode 1: Simulator.fast_reset (nsim.py:78-140): only classes with reset_state() are restored, every
//         entity (except exit doors) is re-appended to its cell in entity_dic order.
// One lane owns one env for the test and the ninja record; the entity tables of the envs that do reset (about 1 % per
// agent step) are then rewritten by the whole warp, entity id = lane, lane + 32, ... (a level has up to a few hundred
// entities: a per-thread loop over them was 7 % of a step).
__global__ void __launch_bounds__(128) reset_kernel(nppsim_state_t st, const uint8_t* mask, int mode, int auto_terminal, int max_frames,
                                                     unsigned long long* stats) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const size_t N = npp_stride(st);
    const bool fresh = mode == 2;
    if (fresh) mode = 0;
    bool need = i < st.n_envs && !(mask && !mask[npp_ext(st, i)]);
    if (need && auto_terminal) {
        int s = st.nin_i[I_STATE * N + i], fr = st.nin_i[I_FRAME * N + i];
        bool term = (s == 6 || s == 7 || s == 8), trunc = (max_frames > 0 && fr >= max_frames);
        need = term || trunc;
        if (need && stats) {
            atomicAdd(stats + 0, 1ULL);
            if (s == 8) atomicAdd(stats + 1, 1ULL); else if (term) atomicAdd(stats + 2, 1ULL); else atomicAdd(stats + 3, 1ULL);
        }
    }
    unsigned todo = __ballot_sync(0xffffffffu, need);
    if (!todo) return;
    if (need) {
        const NppMapHeader* h = reinterpret_cast<const NppMapHeader*>(st.map_blob + st.map_off[st.env_map[i]]);
        double* nf = st.nin_f + i; int32_t* ni = st.nin_i + i; int32_t* envi = st.env_i + i;
        #pragma unroll 1
        for (int f = 0; f < NPPSIM_NF; f++) nf[f * N] = 0.0;
        #pragma unroll 1
        for (int f = 0; f < NPPSIM_NI; f++) ni[f * N] = 0;
        nf[F_XPOS * N] = h->spawn_x; nf[F_YPOS * N] = h->spawn_y;
        nf[F_X_AT_ACTION * N] = h->spawn_x; nf[F_Y_AT_ACTION * N] = h->spawn_y;
        nf[F_GRAVITY * N] = GRAVITY_FALL; nf[F_DRAG * N] = DRAG_REGULAR; nf[F_FRICTION * N] = FRICTION_GROUND;
        nf[F_FNY * N] = -1; nf[F_CNY * N] = 1;
        ni[I_JUMP_BUFFER * N] = -1; ni[I_FLOOR_BUFFER * N] = -1; ni[I_WALL_BUFFER * N] = -1; ni[I_LP_BUFFER * N] = -1;
        envi[NPPSIM_ENV_STAMP * N] = h->n_ent;
        envi[NPPSIM_ENV_EVENTS * N] = 0;
        envi[NPPSIM_ENV_SWFRAME * N] = -1;
        if (mode == 0) envi[NPPSIM_ENV_RELOADS * N] = fresh ? 0 : envi[NPPSIM_ENV_RELOADS * N] + 1;     // (fast_reset re-creates nothing)
        envi[NPPSIM_ENV_ORDER * N] = mode ? 1 : 0;           // (NPPSIM_ENV_WEDGE stays: it belongs to the handoff queue)
    }
    #pragma unroll 1
    for (; todo; todo &= todo - 1) {
        const int e_i = __shfl_sync(0xffffffffu, i, __ffs(todo) - 1);        // the env whose entity tables the warp rewrites now
        const uint8_t* blob = st.map_blob + st.map_off[st.env_map[e_i]];
        const NppMapHeader* h = reinterpret_cast<const NppMapHeader*>(blob);
        const NppStaticEnt* ents = reinterpret_cast<const NppStaticEnt*>(blob + h->off_ents);
        uint8_t* flag = st.ent_flag + e_i; double* dynf = st.dyn_f + e_i; int32_t* dyni = st.dyn_i + e_i;
        const uint32_t n_ent = h->n_ent;
        int hot = 0;   // fast_reset keeps boost pads' is_touching_ninja (no reset_state())
        #pragma unroll 1
        for (uint32_t id = lane; id < n_ent; id += 32) {
            const NppStaticEnt& e = ents[id];
            int stamp = mode ? (int)id : (int)e.load_order;
            uint8_t fl = FL_ACTIVE;
            switch (e.kind) {
            case NPP_K_MINE: fl |= e.p0; break;
            case NPP_K_GOLD: break;
            case NPP_K_BOOST: fl |= FL_IS_BOOST; if (mode) { fl |= flag[id * N] & FL_TOUCH; if (fl & FL_TOUCH) hot++; } break;
            case NPP_K_DOOR_LOCKED: fl |= FL_CLOSED; break;
            case NPP_K_DOOR_TRAP: break;                         // starts open (entity_door_trap.py:51-53)
            case NPP_K_DOOR_REGULAR:
                if (mode == 0) { fl |= FL_CLOSED; dyni[(e.dyn * 3 + 0) * N] = -1; dyni[(e.dyn * 3 + 1) * N] = stamp; dyni[(e.dyn * 3 + 2) * N] = 256; }
                else fl |= flag[id * N] & FL_CLOSED;
                break;
            case NPP_K_EXIT:
                dynf[(e.dyn * 4 + 0) * N] = e.x; dynf[(e.dyn * 4 + 1) * N] = e.y; dynf[(e.dyn * 4 + 2) * N] = 0; dynf[(e.dyn * 4 + 3) * N] = 0;
                dyni[(e.dyn * 3 + 0) * N] = -1; dyni[(e.dyn * 3 + 1) * N] = stamp; dyni[(e.dyn * 3 + 2) * N] = e.cell;
                break;
            case NPP_K_EXIT_SWITCH:
                dynf[(e.dyn * 4 + 0) * N] = e.x; dynf[(e.dyn * 4 + 1) * N] = e.y; dynf[(e.dyn * 4 + 2) * N] = 0; dynf[(e.dyn * 4 + 3) * N] = 0;
                dyni[(e.dyn * 3 + 0) * N] = e.cell; dyni[(e.dyn * 3 + 1) * N] = stamp; dyni[(e.dyn * 3 + 2) * N] = 0;
                break;
            case NPP_K_DRONE_ZAP: case NPP_K_MINI_DRONE: case NPP_K_BOUNCE: case NPP_K_THWUMP: case NPP_K_DEATH_BALL: case NPP_K_SHWUMP:
                if (mode == 0) {
                    bool drone = (e.kind == NPP_K_DRONE_ZAP || e.kind == NPP_K_MINI_DRONE);
                    if (drone) fl |= (uint8_t)(e.p0 << FL_ST_SHIFT);
                    dynf[(e.dyn * 4 + 0) * N] = e.x; dynf[(e.dyn * 4 + 1) * N] = e.y;
                    dynf[(e.dyn * 4 + 2) * N] = drone ? e.x : 0.0; dynf[(e.dyn * 4 + 3) * N] = drone ? e.y : 0.0;   // xtarget,ytarget | speeds | xdir,ydir
                    dyni[(e.dyn * 3 + 0) * N] = e.cell; dyni[(e.dyn * 3 + 2) * N] = 0;
                } else {
                    fl = flag[id * N] | FL_ACTIVE;      // no reset_state(): position, speed, state and cell all persist
                }
                dyni[(e.dyn * 3 + 1) * N] = stamp;
                break;
            default: break;
            }
            flag[id * N] = fl;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) hot += __shfl_xor_sync(0xffffffffu, hot, d);
        if (lane == 0) st.env_i[NPPSIM_ENV_HOT * N + e_i] = hot;
    }
}

}  // namespace

// ---- SM ranks for the solo SMs of sim_variant.inc ---------------------------------------------------------
// %smid values need not be contiguous, so each device gets a table smid -> rank among the SMs that exist (ascending: the
// two SMs of a TPC stay neighbours), built once by a probe launch; an SM the probe never saw ranks last and is never reserved.
#define NPP_MAX_SMID 1024
__device__ int g_smid_rank[NPP_MAX_SMID];
namespace {
__global__ void smid_probe_kernel(int* seen) {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (threadIdx.x == 0 && smid < NPP_MAX_SMID) seen[smid] = 1;
    const long long t0 = clock64();
    while (clock64() - t0 < 20000) { }                  // hold the slot for ~10 us so that the grid spreads over every SM
}
__global__ void smid_rank_kernel(const int* seen) {
    int r = 0;
    for (int k = 0; k < NPP_MAX_SMID; k++) g_smid_rank[k] = seen[k] ? r++ : (1 << 30);
}
}  // namespace

// ---- step kernel variants ------------------------------------------------------------------------------
#ifndef STEP_BLOCK
#define STEP_BLOCK 128
#endif
// __launch_bounds__ minimum: 1 leaves the register budget to ptxas (254-255 per thread, so TWO 128-thread blocks are resident
// per SM); 3 / 4 cap it at 168 / 128 registers and spill the ninja record (measured -14 % / -20 %)
#ifndef STEP_BLOCKS_PER_SM
#define STEP_BLOCKS_PER_SM 1
#endif
#ifdef NPP_STAGE_TMA
#define NPP_STAGE_BYTES (NPP_NCELL * 4 + NPP_NCELL * NPP_SEGS_PER_CELL * 4)        // cell_hdr + segs: contiguous behind the blob header
#define STEP_SMEM ((STEP_BLOCK / 32) * sizeof(WarpPool) + NPP_STAGE_BYTES + 16)
#else
#define STEP_SMEM ((STEP_BLOCK / 32) * sizeof(WarpPool))
#endif
#define NPP_MOVE_MASK ((1u << NPP_K_DRONE_ZAP) | (1u << NPP_K_MINI_DRONE) | (1u << NPP_K_BOUNCE) | (1u << NPP_K_THWUMP))
#define NPP_THINK_MASK ((1u << NPP_K_DOOR_REGULAR) | (1u << NPP_K_THWUMP) | (1u << NPP_K_DEATH_BALL) | (1u << NPP_K_SHWUMP))
#define NPP_PHYS_MASK ((1u << NPP_K_ONE_WAY) | (1u << NPP_K_BOUNCE) | (1u << NPP_K_THWUMP) | (1u << NPP_K_SHWUMP))
// "basic" levels: nothing moves, nothing pushes (every generated curriculum category of BASELINE.json's configs)
#define NPP_BASIC_MASK ((1u << NPP_K_NONE) | (1u << NPP_K_MINE) | (1u << NPP_K_GOLD) | (1u << NPP_K_EXIT) | (1u << NPP_K_EXIT_SWITCH) | (1u << NPP_K_DOOR_LOCKED) | (1u << NPP_K_DOOR_TRAP))

// NPP_GROUPS: the grouped cooperative service of ninja.cuh (several waiting lanes of a warp served at once).  It pays wherever
// the launch is throughput bound (+7 % config 4, +11 % 131,072 envs of config 3, +45 % the 130 replay levels) and costs 2 % in
// code size / registers where wedged envs leave their warps anyway (handoff: up to two waves of warps), so the basic kernel
// exists with and without it and the launcher picks by whether the state carries a handoff queue (profiles/r02_handoff_ab.md).
#ifdef NPP_COOP_NOGROUPS
#define NPP_GROUPS_DEFAULT 0
#else
#define NPP_GROUPS_DEFAULT 1
#endif
// NPP_SOLO: solo warps + handoff queue (sim_variant.inc).  Only sim_basic_solo carries them: the flag tests at every phase
// barrier and the generalised tick loop cost the other variants 3-7 % (one-level config 2: 0.096 -> 0.104 ms per launch).
#define NPP_VARIANT sim_full
#define NPP_KINDS 0xffffffffu
#define NPP_GROUPS NPP_GROUPS_DEFAULT
#define NPP_SOLO 0
#include "sim_variant.inc"
#undef NPP_VARIANT
#undef NPP_KINDS
#undef NPP_GROUPS
#undef NPP_SOLO
#define NPP_VARIANT sim_basic_solo
#define NPP_KINDS NPP_BASIC_MASK
#ifdef NPP_SOLO_GROUPS
#define NPP_GROUPS NPP_GROUPS_DEFAULT
#else
#define NPP_GROUPS 0
#endif
#define NPP_SOLO 1
#include "sim_variant.inc"
#undef NPP_VARIANT
#undef NPP_KINDS
#undef NPP_GROUPS
#undef NPP_SOLO
#define NPP_VARIANT sim_basic
#define NPP_KINDS NPP_BASIC_MASK
#define NPP_GROUPS NPP_GROUPS_DEFAULT
#define NPP_SOLO 0
#include "sim_variant.inc"
#undef NPP_VARIANT
#undef NPP_KINDS
#undef NPP_GROUPS
#undef NPP_SOLO

namespace {
// Longest-first order for the next launch from the costs of this one (one block; nw <= 2^20 / 32).
// Six classes by cost relative to the mean (sched_class), placed in that order; inside a
// class the order is arbitrary (it only affects scheduling, never results).
#define SCHED_NCLS 6
__device__ __forceinline__ int sched_class(unsigned long long v, unsigned long long mean) {
    // cost relative to the mean: > 24x, > 12x, > 6x, > 3x, > 1.5x, the rest
    return v > 24 * mean ? 0 : (v > 12 * mean ? 1 : (v > 6 * mean ? 2 : (v > 3 * mean ? 3 : (2 * v > 3 * mean ? 4 : 5))));
}
__global__ void __launch_bounds__(1024) sched_kernel(int32_t* __restrict__ order, const uint32_t* __restrict__ cost, int nw, int32_t* __restrict__ handoff,
                                                     int32_t* __restrict__ wedge_mark, int solo_limit, int solo_sm_max, int solo_sm_q8) {
    __shared__ unsigned long long s_sum;
    if (handoff) {
        // Handoff queue between two launches (sim_variant.inc): the envs the finished launch appended to `next` become the
        // initial entries of the following one and get this queue's mark; the previous initial entries lose theirs.
        int32_t* cur = handoff + NPPSIM_HANDOFF_HDR; int32_t* next = cur + NPPSIM_HANDOFF_CAP;
        // (at most solo_limit envs start on solo warps: on levels where a large share of the ninjas is wedged the chains are the
        // bulk of the work, not its tail, and a warp per env would only take slots from the regular warps)
        const int tag = handoff[6], n_old = min(handoff[7], NPPSIM_HANDOFF_CAP), n_new = min(min(handoff[5], NPPSIM_HANDOFF_CAP), solo_limit);
        for (int k = threadIdx.x; k < n_old; k += blockDim.x) { const int v = cur[k]; if (v > 0 && wedge_mark[v - 1] == tag) wedge_mark[v - 1] = 0; }
        __syncthreads();
        for (int k = threadIdx.x; k < NPPSIM_HANDOFF_CAP; k += blockDim.x) {
            const int v = k < n_new ? next[k] : 0;
            cur[k] = v; next[k] = 0;
            if (v > 0) wedge_mark[v - 1] = tag;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            // saturation counters for the caller's policy: launches whose wedged envs did not all fit the queue / launches
            if (handoff[5] > solo_limit) handoff[12]++;
            handoff[13]++;
            handoff[0] = n_new; handoff[1] = n_new; handoff[2] = 0; handoff[5] = 0; handoff[7] = n_new;
            // SMs reserved for solo blocks in the next launch: solo_sm_q8 / 256 warps per initial entry, eight warps per SM, at
            // least one SM (the envs handed over mid-launch need a warp too)
            handoff[8] = 0; handoff[9] = 0; handoff[11] = 0;
            handoff[10] = solo_sm_max > 0 ? max(1, min(solo_sm_max, (n_new * solo_sm_q8 + 8 * 256 - 1) / (8 * 256))) : 0;
        }
    }
    if (!order || !cost) return;
    __shared__ int s_cnt[SCHED_NCLS], s_pos[SCHED_NCLS];
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { s_sum = 0; for (int k = 0; k < SCHED_NCLS; k++) s_cnt[k] = 0; }
    __syncthreads();
    unsigned long long loc = 0;
    for (int w = threadIdx.x; w < nw; w += blockDim.x) loc += cost[w];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) loc += __shfl_xor_sync(0xffffffffu, loc, d);
    if (lane == 0) atomicAdd(&s_sum, loc);
    __syncthreads();
    const unsigned long long mean = s_sum / (unsigned long long)max(nw, 1);
    const int rounds = (nw + blockDim.x - 1) / blockDim.x;
    // warp-aggregated counting / placement: one shared-memory atomic per warp and class
    for (int r = 0; r < rounds; r++) {
        const int w = r * blockDim.x + threadIdx.x;
        const int cls = w < nw ? sched_class(cost[w], mean) : SCHED_NCLS;
#pragma unroll
        for (int k = 0; k < SCHED_NCLS; k++) { unsigned m = __ballot_sync(0xffffffffu, cls == k); if (lane == 0 && m) atomicAdd(&s_cnt[k], __popc(m)); }
    }
    __syncthreads();
    if (threadIdx.x == 0) { int p = 0; for (int k = 0; k < SCHED_NCLS; k++) { s_pos[k] = p; p += s_cnt[k]; } }
    __syncthreads();
    for (int r = 0; r < rounds; r++) {
        const int w = r * blockDim.x + threadIdx.x;
        const int cls = w < nw ? sched_class(cost[w], mean) : SCHED_NCLS;
#pragma unroll
        for (int k = 0; k < SCHED_NCLS; k++) {
            unsigned m = __ballot_sync(0xffffffffu, cls == k);
            int base = 0;
            if (lane == 0 && m) base = atomicAdd(&s_pos[k], __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (cls == k) order[base + __popc(m & ((1u << lane) - 1))] = w;
        }
    }
}

// ---- observation primitives -----------------------------------------------------------
// One thread per env, straight from the SoA state.  ninja_state: NPlayHeadless.get_ninja_state
// (nplay_headless.py:735-917 + Ninja.get_temporal_features ninja.py:1217-1244), computed in double and stored
// as the float32 the env wrapper converts it to; `x ** 0.5` is evaluated as sqrt (differs from libm pow(x, 0.5)
// by at most one double ulp, invisible after the float32 rounding except on a rounding boundary).
// mask: Ninja.get_valid_action_mask (ninja.py:628-839), masking mode "hard"; path_dir[i] in {0 None, 1 "left",
// 2 "right", 3 "down"} is sim._path_straightness_direction (NULL: None everywhere).
DEV double obs_min(double a, double b) { return b < a ? b : a; }      // Python min(a, b) / max(a, b)
DEV double obs_max(double a, double b) { return b > a ? b : a; }
// NPlayHeadless.get_ninja_state for env i -> o[0..39]
DEV void ninja_state_row(const nppsim_state_t& st, int i, float* __restrict__ o) {
    const size_t N = npp_stride(st);
    const double* f = st.nin_f + i; const int32_t* q = st.nin_i + i;
#define NF_(k) f[(size_t)(k) * N]
#define NI_(k) q[(size_t)(k) * N]
    const int state = NI_(I_STATE), airborn = NI_(I_AIRBORN), walled = NI_(I_WALLED), jump_input = NI_(I_JUMP_INPUT);
    const int jump_buffer = NI_(I_JUMP_BUFFER), floor_buffer = NI_(I_FLOOR_BUFFER), wall_buffer = NI_(I_WALL_BUFFER);
    const double wall_normal = NF_(F_WALL_NORMAL), yspeed = NF_(F_YSPEED);
    const double xspeed = NF_(F_XSPEED), xspeed_old = NF_(F_XSPEED_OLD), yspeed_old = NF_(F_YSPEED_OLD);
    const double fnx = NF_(F_FNX), fny = NF_(F_FNY), gravity = NF_(F_GRAVITY);
    const int floor_count = NI_(I_FLOOR_COUNT), ceiling_count = NI_(I_CEILING_COUNT);
    const int wall_count = 0;                                  // never written by the reference
    const double v2 = SQ(xspeed) + SQ(yspeed);
    const double velocity_magnitude = sqrt(v2);
    o[0] = (float)(obs_min(velocity_magnitude / (MAX_HOR_SPEED * 2), 1.0) * 2 - 1);
    o[1] = velocity_magnitude > 1e-6 ? (float)(xspeed / velocity_magnitude) : 0.0f;
    o[2] = velocity_magnitude > 1e-6 ? (float)(yspeed / velocity_magnitude) : 0.0f;
    o[3] = (state >= 0 && state <= 2) ? 1.0f : -1.0f;
    o[4] = (state == 3 || state == 4) ? 1.0f : -1.0f;
    o[5] = state == 5 ? 1.0f : -1.0f;
    o[6] = (state >= 6 && state <= 9) ? 1.0f : -1.0f;
    o[7] = airborn ? 1.0f : -1.0f;
    o[8] = (float)NI_(I_HOR_INPUT);
    o[9] = jump_input ? 1.0f : -1.0f;
    o[10] = (float)((max(jump_buffer, 0) / 5.0) * 2 - 1);
    o[11] = (float)((max(floor_buffer, 0) / 5.0) * 2 - 1);
    o[12] = (float)((max(wall_buffer, 0) / 5.0) * 2 - 1);
    o[13] = (float)(min(floor_count, 1) * 2 - 1);
    o[14] = (float)(min(wall_count, 1) * 2 - 1);
    o[15] = (float)(min(ceiling_count, 1) * 2 - 1);
    o[16] = (float)(sqrt(SQ(fnx) + SQ(fny)) * 2 - 1);
    o[17] = 0.0f;                                              // wall_count > 0 never holds
    o[18] = (float)fny;
    o[19] = (float)((gravity - GRAVITY_JUMP) / (GRAVITY_FALL - GRAVITY_JUMP) * 2 - 1);
    o[20] = walled ? 1.0f : -1.0f;
    o[21] = (float)fnx;
    o[22] = (float)NF_(F_CNX);
    o[23] = (float)NF_(F_CNY);
    o[24] = (float)((NF_(F_DRAG) - DRAG_SLOW) / (DRAG_REGULAR - DRAG_SLOW) * 2 - 1);
    o[25] = (float)((NF_(F_FRICTION) - FRICTION_GROUND_SLOW) / (FRICTION_GROUND - FRICTION_GROUND_SLOW) * 2 - 1);
    o[26] = (float)obs_max(-1.0, obs_min(1.0, (xspeed - xspeed_old) / MAX_HOR_SPEED));
    o[27] = (float)obs_max(-1.0, obs_min(1.0, (yspeed - yspeed_old) / MAX_HOR_SPEED));
    o[28] = (float)(obs_min(velocity_magnitude / (MAX_HOR_SPEED * 1.5), 1.0) * 2 - 1);
    o[29] = (float)(obs_min(NI_(I_FRAMES_AIRBORNE) / 60.0, 1.0) * 2 - 1);
    o[30] = (float)(obs_min((double)NI_(I_JUMP_DURATION) / MAX_JUMP_DURATION, 1.0) * 2 - 1);
    o[31] = (float)(obs_min(NI_(I_STATE_CHANGE_FRAME) / 30.0, 1.0) * 2 - 1);
    const double kinetic_energy = 0.5 * v2;
    o[32] = (float)(obs_min(kinetic_energy / SQ(MAX_HOR_SPEED), 1.0) * 2 - 1);
    o[33] = (float)((NF_(F_YPOS) / 600.0) * 2 - 1);
    o[34] = (float)(obs_min(sqrt(SQ(gravity) + SQ(!airborn ? GROUND_ACCEL : AIR_ACCEL)) / 0.1, 1.0) * 2 - 1);
    const double prev_kinetic = 0.5 * (SQ(xspeed_old) + SQ(yspeed_old));
    o[35] = (float)obs_max(-1.0, obs_min(1.0, (kinetic_energy - prev_kinetic) / obs_max(kinetic_energy + 0.01, 0.01)));
    o[36] = (float)(obs_min(floor_count / 5.0, 1.0) * 2 - 1);
    o[37] = (float)(obs_min(wall_count / 3.0, 1.0) * 2 - 1);
    o[38] = (float)(atan2(fny, fnx) / 3.141592653589793);
    o[39] = walled ? (float)wall_normal : 0.0f;
#undef NF_
#undef NI_
}

__global__ void __launch_bounds__(128) obs_kernel(nppsim_state_t st, const uint8_t* __restrict__ path_dir, uint8_t* __restrict__ mask_out, float* __restrict__ state_out,
                                                  double* __restrict__ pos_vel_out, int32_t* __restrict__ ninja_state_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= st.n_envs) return;
    const int x = npp_ext(st, i);                               // outputs (and path_dir) are in the caller's env order
    const size_t N = npp_stride(st);
    const double* f = st.nin_f + i; const int32_t* q = st.nin_i + i;
#define NF_(k) f[(size_t)(k) * N]
#define NI_(k) q[(size_t)(k) * N]
    const int state = NI_(I_STATE), airborn = NI_(I_AIRBORN), walled = NI_(I_WALLED), jump_input = NI_(I_JUMP_INPUT);
    const int jump_buffer = NI_(I_JUMP_BUFFER), floor_buffer = NI_(I_FLOOR_BUFFER), wall_buffer = NI_(I_WALL_BUFFER), lp_buffer = NI_(I_LP_BUFFER);
    const double wall_normal = NF_(F_WALL_NORMAL), yspeed = NF_(F_YSPEED);
    if (mask_out) {
        bool m[6] = {true, true, true, true, true, true};
        const bool has_active_buffer = (-1 < jump_buffer && jump_buffer < 5) || (-1 < floor_buffer && floor_buffer < 5) ||
                                       (-1 < wall_buffer && wall_buffer < 5) || (-1 < lp_buffer && lp_buffer < 4);
        bool jump_useless = false;
        if (airborn && state != 3) jump_useless = jump_input != 0 ? !has_active_buffer : false;
        if (jump_useless) { m[3] = false; m[4] = false; m[5] = false; }
        if (walled) {
            if (!airborn || state == 5) { if (wall_normal > 0) m[1] = false; else if (wall_normal < 0) m[2] = false; }
            else {
                const bool would_trigger_wall_slide = yspeed >= 0;
                if (wall_normal > 0) { if (!would_trigger_wall_slide) m[1] = false; }
                else if (wall_normal < 0) { if (!would_trigger_wall_slide) m[2] = false; }
            }
        }
        const int pd = path_dir ? path_dir[x] : 0;
        if (pd == 1) { m[2] = false; m[3] = false; m[4] = false; m[5] = false; }
        else if (pd == 2) { m[1] = false; m[3] = false; m[4] = false; m[5] = false; }
        else if (pd == 3) { m[3] = false; m[4] = false; m[5] = false; }
        if (!(m[0] || m[1] || m[2] || m[3] || m[4] || m[5])) m[0] = true;
#pragma unroll
        for (int k = 0; k < 6; k++) mask_out[(size_t)x * 6 + k] = m[k] ? 1 : 0;
    }
    if (state_out) ninja_state_row(st, i, state_out + (size_t)x * 40);
    if (pos_vel_out) {                                          // nppsim_step_host's block: xpos, ypos, xspeed, yspeed as [4][n_envs], then the state
#pragma unroll
        for (int k = 0; k < 4; k++) pos_vel_out[(size_t)k * st.n_envs + x] = NF_(k);
    }
    if (ninja_state_out) ninja_state_out[x] = state;
#undef NF_
#undef NI_
}

// ---- replay -> dataset rows --------------------------------------------------------------
// The simulator-derived columns of ReplayExecutor._get_raw_observation (replay/replay_executor.py:470-600) for the
// observation taken after stream tick `obs_tick` (frame_idx = obs_tick - 1, kept iff (frame_idx + 1) % frame_skip == 0,
// :441), one row per env whose replay is at least that long, at row_off[env] + obs_tick / frame_skip - 1.
__global__ void __launch_bounds__(128) replay_rows_kernel(nppsim_state_t st, nppsim_input_stream_t in, const int64_t* __restrict__ row_off, int obs_tick,
                                                          int frame_skip, float* __restrict__ game_state, double* __restrict__ pos,
                                                          uint8_t* __restrict__ flags, int32_t* __restrict__ frame) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= st.n_envs) return;
    const int x = npp_ext(st, i);
    const long long len = in.offset[x + 1] - in.offset[x];
    if (obs_tick > len) return;
    const size_t N = npp_stride(st);
    const size_t row = (size_t)(row_off[x] + obs_tick / frame_skip - 1);
    const int state = st.nin_i[(size_t)I_STATE * N + i];
    if (game_state) {
        float* o = game_state + row * 41;
        ninja_state_row(st, i, o);
        const double est = (double)len;                                   // len(_current_input_sequence) or 2000: len >= 1 here
        const double tr = (est - (double)st.nin_i[(size_t)I_FRAME * N + i]) / est;
        o[40] = (float)(tr > 0.0 ? tr : 0.0);
    }
    if (pos || flags) {
        const uint8_t* blob = st.map_blob + st.map_off[st.env_map[i]];
        const NppMapHeader* h = reinterpret_cast<const NppMapHeader*>(blob);
        const NppStaticEnt* ents = reinterpret_cast<const NppStaticEnt*>(blob + h->off_ents);
        int sw = -1;                                                      // NPlayHeadless._sim_exit_switch: reversed(entity_dic[3])
        #pragma unroll 1
        for (int id = (int)h->n_ent - 1; id >= 0; id--) if (ents[id].kind == NPP_K_EXIT_SWITCH) { sw = id; break; }
        if (pos) {
            double* p = pos + row * 6;
            p[0] = st.nin_f[(size_t)F_XPOS * N + i]; p[1] = st.nin_f[(size_t)F_YPOS * N + i];
            p[2] = p[3] = p[4] = p[5] = 0.0;                              // (0, 0) without an exit switch (nplay_headless.py:577-616)
            if (sw >= 0) {
                const int s = ents[sw].dyn, d = ents[ents[sw].aux].dyn;
                p[2] = st.dyn_f[(size_t)(s * 4 + 0) * N + i]; p[3] = st.dyn_f[(size_t)(s * 4 + 1) * N + i];
                p[4] = st.dyn_f[(size_t)(d * 4 + 0) * N + i]; p[5] = st.dyn_f[(size_t)(d * 4 + 1) * N + i];
            }
        }
        if (flags) {
            uint8_t* fl = flags + row * 4;
            fl[0] = sw < 0 ? 1 : ((st.ent_flag[(size_t)sw * N + i] & FL_ACTIVE) ? 0 : 1);   // exit_switch_activated (:566-575)
            fl[1] = state == 8;                                           // ninja_has_won
            fl[2] = (state == 6 || state == 7);                           // Ninja.has_died
            const int b = in.data[in.offset[x] + obs_tick - 1];
            const uint8_t to_action[8] = {0, 3, 2, 5, 1, 4, 0, 0};        // map_input_to_action (replay_executor.py:42-58)
            fl[3] = b < 8 ? to_action[b] : 0;
        }
    }
    if (frame) frame[row] = obs_tick - 1;
}

// ---- curriculum repositioning (SURVEY.md 8(b) "mutation from outside tick", 8(f) f2) -----------------------
// IntermediateGoalManager.apply_to_simulator (gym_environment/reward_calculation/intermediate_goal_manager.py:698-960)
// for every masked env: the LAST exit switch of entity_dic[3] (reversed() scan, :725-729) moves to switch_xy and is
// re-armed while sim.frame <= 10 (:843-845); changing cell = list.remove + append, i.e. a fresh insertion stamp
// (:877-894).  Its parent door moves to door_xy; a door that changes cell leaves grid_entity and is NOT re-added
// (:938-953) -- it comes back when the switch is taken (entity_exit_switch.py:120).
__global__ void __launch_bounds__(128) curriculum_kernel(nppsim_state_t st, const uint8_t* __restrict__ mask, const double* __restrict__ switch_xy,
                                                         const double* __restrict__ door_xy, int32_t* __restrict__ status) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= st.n_envs) return;
    const int x = npp_ext(st, i);
    if (mask && !mask[x]) return;
    const size_t N = npp_stride(st);
    const uint8_t* blob = st.map_blob + st.map_off[st.env_map[i]];
    const NppMapHeader* h = reinterpret_cast<const NppMapHeader*>(blob);
    const NppStaticEnt* ents = reinterpret_cast<const NppStaticEnt*>(blob + h->off_ents);
    int sw = -1;
    #pragma unroll 1
    for (int id = (int)h->n_ent - 1; id >= 0; id--) if (ents[id].kind == NPP_K_EXIT_SWITCH) { sw = id; break; }
    if (status) status[x] = sw < 0 ? -1 : 0;
    if (sw < 0) return;                                       // the reference raises RuntimeError (:768-781)
    double* dynf = st.dyn_f + i; int32_t* dyni = st.dyn_i + i;
    const int s = ents[sw].dyn, d = ents[ents[sw].aux].dyn;
    const double sx = switch_xy[x], sy = switch_xy[(size_t)st.n_envs + x];          // caller arrays are dense over this batch
    dynf[(size_t)(s * 4 + 0) * N] = sx; dynf[(size_t)(s * 4 + 1) * N] = sy;
    if (st.nin_i[(size_t)I_FRAME * N + i] <= 10) st.ent_flag[(size_t)sw * N + i] |= FL_ACTIVE;
    int cx = min(max((int)floor(sx / 24), 0), NPP_NCX - 1), cy = min(max((int)floor(sy / 24), 0), NPP_NCY - 1);
    int cell = cx * NPP_NCY + cy;
    if (cell != dyni[(size_t)(s * 3 + 0) * N]) {
        dyni[(size_t)(s * 3 + 0) * N] = cell;
        dyni[(size_t)(s * 3 + 1) * N] = st.env_i[(size_t)NPPSIM_ENV_STAMP * N + i]++;
    }
    const double dx = door_xy[x], dy = door_xy[(size_t)st.n_envs + x];
    dynf[(size_t)(d * 4 + 0) * N] = dx; dynf[(size_t)(d * 4 + 1) * N] = dy;
    cx = min(max((int)floor(dx / 24), 0), NPP_NCX - 1); cy = min(max((int)floor(dy / 24), 0), NPP_NCY - 1);
    cell = cx * NPP_NCY + cy;
    if (cell != dyni[(size_t)(d * 3 + 2) * N]) {              // slot word 2 always holds door.cell; word 0 is -1 outside the grid
        dyni[(size_t)(d * 3 + 0) * N] = -1;
        dyni[(size_t)(d * 3 + 2) * N] = cell;
    }
}

__global__ void div2_kernel(const double* a1, const double* a2, const double* b, unsigned long long* mismatches, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double q1, q2;
    div2(a1[i], a2[i], b[i], q1, q2);
    const double w1 = a1[i] / b[i], w2 = a2[i] / b[i];
    if (__double_as_longlong(q1) != __double_as_longlong(w1) || __double_as_longlong(q2) != __double_as_longlong(w2)) atomicAdd(mismatches, 1ULL);
}
__global__ void sq_kernel(const double* in, double* out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = SQ(in[i]);
}

int check_state(const nppsim_state_t* st) {
    if (!st) return fail(NPPSIM_E_ARG, "state is NULL");
    if (st->n_envs <= 0) return fail(NPPSIM_E_ARG, "n_envs must be positive");
    if (!st->nin_f || !st->nin_i || !st->env_i || !st->env_map || !st->map_blob || !st->map_off)
        return fail(NPPSIM_E_ARG, "a required state pointer is NULL");
    if ((st->ent_cap > 0 && !st->ent_flag) || (st->dyn_cap > 0 && (!st->dyn_f || !st->dyn_i)))
        return fail(NPPSIM_E_ARG, "entity buffers missing for a non-zero capacity");
    if (st->stride != 0 && st->stride < (uint32_t)st->n_envs) return fail(NPPSIM_E_ARG, "stride must be 0 or >= n_envs");
    return 0;
}

int check_stream(const nppsim_input_stream_t* in) {
    if (!in || !in->data || !in->offset) return fail(NPPSIM_E_ARG, "input stream (data / offset) is NULL");
    if (in->repeat < 1) return fail(NPPSIM_E_ARG, "stream repeat must be >= 1");
    if (in->start_tick < 0) return fail(NPPSIM_E_ARG, "stream start_tick must be >= 0");
    if (in->encoding != NPPSIM_STREAM_ACTIONS && in->encoding != NPPSIM_STREAM_REPLAY_BYTES) return fail(NPPSIM_E_ARG, "unknown stream encoding");
    return 0;
}

int launch_step(const nppsim_state_t* st, const int8_t* hor, const int8_t* jump, const uint8_t* actions, int input_ticks, int n_ticks,
                const int32_t* tick_limit, int stop_on_terminal, int32_t* frames_executed, void* stream,
                const nppsim_input_stream_t* in = nullptr, double* host_pv = nullptr, int32_t* host_state = nullptr) {
    int rc = check_state(st);
    if (rc) return rc;
    if (n_ticks <= 0) return fail(NPPSIM_E_ARG, "n_ticks must be positive");
    nppsim_input_stream_t sp;
    memset(&sp, 0, sizeof(sp));
    sp.repeat = 1;
    if (in) sp = *in;
    if (!in && !actions && (!hor || !jump)) return fail(NPPSIM_E_ARG, "inputs are NULL");
    if (!in && !actions && !(input_ticks == 1 || input_ticks == n_ticks)) return fail(NPPSIM_E_ARG, "input_ticks must be 1 or n_ticks");
    const int block = STEP_BLOCK;
    static int sm_count[64] = {0};
    // shared-memory carve-out: room for the two blocks per SM the register file admits and not more -- the rest
    // of the 256 KB stays L1 for the level blobs (max-shared left 28 KB of L1: hit rate 40 % instead of 66 %)
    // (function attributes belong to the current device's context: set once per device, results checked)
    int sms = 148;
    {
        static std::mutex init_mu;
        static bool init_done[64] = {false};
        int dev = 0;
        cudaError_t ie = cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lock(init_mu);
        if (ie == cudaSuccess && (dev < 0 || dev >= 64 || !init_done[dev])) {
            const int resident = STEP_BLOCKS_PER_SM > 2 ? STEP_BLOCKS_PER_SM : 2;
            const int carve = (int)((resident * (STEP_SMEM + 1024) * 100 + 233471) / 233472) + 1;
            ie = cudaFuncSetAttribute(sim_basic::step_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            if (ie == cudaSuccess) ie = cudaFuncSetAttribute(sim_full::step_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            if (ie == cudaSuccess) ie = cudaFuncSetAttribute(sim_basic::step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)STEP_SMEM);
            if (ie == cudaSuccess) ie = cudaFuncSetAttribute(sim_basic_solo::step_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            if (ie == cudaSuccess) ie = cudaFuncSetAttribute(sim_basic_solo::step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)STEP_SMEM);
            if (ie == cudaSuccess) ie = cudaFuncSetAttribute(sim_full::step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)STEP_SMEM);
            if (ie == cudaSuccess) {                                   // g_smid_rank of this device (solo SMs)
                int* seen = nullptr;
                ie = cudaMalloc(&seen, NPP_MAX_SMID * sizeof(int));
                if (ie == cudaSuccess) ie = cudaMemset(seen, 0, NPP_MAX_SMID * sizeof(int));
                if (ie == cudaSuccess) {
                    int n_sm = 148;
                    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
                    smid_probe_kernel<<<n_sm * 16, 128>>>(seen);
                    smid_rank_kernel<<<1, 1>>>(seen);
                    ie = cudaDeviceSynchronize();
                }
                if (seen) cudaFree(seen);
            }
            if (ie == cudaSuccess && dev >= 0 && dev < 64) init_done[dev] = true;
        }
        if (ie != cudaSuccess) return cuda_fail(ie, "step kernel attributes");
        if (dev >= 0 && dev < 64 && !sm_count[dev]) cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
        sms = (dev >= 0 && dev < 64 && sm_count[dev] > 0) ? sm_count[dev] : 148;
    }
    // Solo blocks at the front of the grid for wedged envs (sim_variant.inc): 4 solo warps per 512 envs; all but the first
    // eight blocks exit at once unless the previous launch left them an env.  Solo warps that had work stay resident and
    // poll until the regular blocks are done, so their number is bounded well below the resident block slots (2 per SM):
    // at most sms / 2 blocks can stay, whatever the batch size -- the regular blocks they wait for always find a slot.
    const int stay_cap = std::max(1, sms / 2 - 8) * (block / 32);
    // kind_mask == 0: census unknown (older callers) -> the full kernel.  Solo warps exist in the basic kernel only.
    const bool basic = st->kind_mask != 0 && (st->kind_mask & ~NPP_BASIC_MASK) == 0;
    int32_t* const handoff = basic ? st->handoff : nullptr;
    const int n_regular = (st->n_envs + block - 1) / block;
    // Solo SMs (sim_variant.inc): from one full wave of blocks up, solo blocks are dealt by SM -- up to solo_sm_max SMs run
    // nothing but solo warps for the launch (sched_kernel sizes the reservation from the queue) -- and only two front blocks
    // remain as the guaranteed servers; the grid holds a block for every slot of every SM that may be reserved.  Smaller
    // launches keep solo blocks at the front of the grid.  NPPSIM_SOLO_SM=0|1|2 / NPPSIM_SOLO_SM_MAX / NPPSIM_SOLO_SM_Q8: A/B knobs.
    // (read per launch: the parity tests switch them; 2 = solo SMs at any grid size)
    const char* const ev_sm = getenv("NPPSIM_SOLO_SM"); const char* const ev_max = getenv("NPPSIM_SOLO_SM_MAX"); const char* const ev_q8 = getenv("NPPSIM_SOLO_SM_Q8");
    const int env_sm = ev_sm ? atoi(ev_sm) : 1, env_sm_max = ev_max ? atoi(ev_max) : 0, env_sm_q8 = ev_q8 ? atoi(ev_q8) : 256;
    const bool sm_solo = handoff && env_sm && (n_regular >= 2 * sms || env_sm == 2);
    const int solo_sm_max = sm_solo ? (env_sm_max > 0 ? std::min(env_sm_max, sms / 4) : sms / 8) : 0;
    const int n_solo = !handoff ? 0 : (sm_solo ? 2 : std::min(sms / 2, std::max(2, st->n_envs / 512)));
    const int solo_limit = std::min(stay_cap, std::max(64, st->n_envs / 384));      // envs per launch that start on solo warps; as many again may follow mid-launch
    const int grid = n_regular + n_solo + 2 * solo_sm_max;
    auto kernel = basic ? (handoff ? sim_basic_solo::step_kernel : sim_basic::step_kernel) : sim_full::step_kernel;
    kernel<<<grid, block, STEP_SMEM, (cudaStream_t)stream>>>(*st, hor, jump, actions, input_ticks, n_ticks, tick_limit, stop_on_terminal, frames_executed, sp, n_solo, 2 * solo_limit,
                                                               n_regular, sm_solo ? 1 : 0, host_pv, host_state);
    g_launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "step_kernel launch");
    if ((st->warp_order && st->warp_cost) || handoff) {
        const size_t S = st->stride ? (size_t)st->stride : (size_t)st->n_envs;
        sched_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(st->warp_cost ? st->warp_order : nullptr, st->warp_cost, (st->n_envs + 31) / 32, handoff,
                                                           st->env_i + (size_t)NPPSIM_ENV_WEDGE * S, solo_limit, solo_sm_max, env_sm_q8);
        g_launches++;
        e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(e, "sched_kernel launch");
    }
    return 0;
}

}  // namespace

extern "C" {

int nppsim_version(void) { return NPPSIM_VERSION; }
const char* nppsim_last_error(void) { return g_nppsim_err; }
int64_t nppsim_launch_count(void) { return (int64_t)g_launches.load(); }
int nppsim_sq_mode(void) { return NPP_SQ_MODE; }
#ifdef NPP_PROFILE
int nppsim_debug_warp_cycles(unsigned long long* out, unsigned long long* phases, int n, int reset) {
    cudaDeviceSynchronize();
    if (out) cudaMemcpyFromSymbol(out, g_warp_cycles, sizeof(unsigned long long) * n);
    if (phases) cudaMemcpyFromSymbol(phases, g_warp_phase, sizeof(unsigned long long) * n * 4);
    return 0;
}
int nppsim_debug_warp_span(unsigned long long* out, int n) {
    cudaDeviceSynchronize();
    return (int)cudaMemcpyFromSymbol(out, g_warp_span, sizeof(unsigned long long) * n * 2);
}
int nppsim_debug_solo_log(unsigned long long* out, int max_rows, int reset) {      // -> rows logged since the last reset
    cudaDeviceSynchronize();
    unsigned int n = 0;
    cudaMemcpyFromSymbol(&n, g_solo_n, sizeof(n));
    if (out) cudaMemcpyFromSymbol(out, g_solo_log, sizeof(unsigned long long) * 12 * (size_t)std::min<unsigned>(std::min<unsigned>(n, 8192u), (unsigned)max_rows));
    if (reset) { unsigned int z = 0; cudaMemcpyToSymbol(g_solo_n, &z, sizeof(z)); }
    return (int)n;
}
int nppsim_debug_coop_cycles(unsigned long long* out8, int reset) {
    cudaDeviceSynchronize();
    if (out8) cudaMemcpyFromSymbol(out8, g_coop_cycles, sizeof(unsigned long long) * 8);
    if (reset) { unsigned long long z[8] = {0}; cudaMemcpyToSymbol(g_coop_cycles, z, sizeof(z)); }
    return 0;
}
int nppsim_debug_kind_cycles(unsigned long long* out128, int reset) {      // [move | think][32 kinds][cycles, calls]
    cudaDeviceSynchronize();
    if (out128) cudaMemcpyFromSymbol(out128, g_kind_cycles, sizeof(unsigned long long) * 128);
    if (reset) { unsigned long long z[128] = {0}; cudaMemcpyToSymbol(g_kind_cycles, z, sizeof(z)); }
    return 0;
}
int nppsim_debug_phase_cycles(unsigned long long* out16, int reset) {
    cudaDeviceSynchronize();
    if (out16) cudaMemcpyFromSymbol(out16, g_phase_cycles, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(g_phase_cycles, z, sizeof(z)); }
    return 0;
}
#endif

int nppsim_debug_div2(const double* a1, const double* a2, const double* b, unsigned long long* mismatches, int32_t n, void* stream) {
    g_err[0] = 0;
    if (!a1 || !a2 || !b || !mismatches || n <= 0) return fail(NPPSIM_E_ARG, "bad arguments");
    div2_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(a1, a2, b, mismatches, n);
    g_launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "div2_kernel launch");
    return 0;
}

int nppsim_debug_sq(const double* in, double* out, int32_t n, void* stream) {
    g_err[0] = 0;
    if (!in || !out || n <= 0) return fail(NPPSIM_E_ARG, "bad arguments");
    sq_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(in, out, n);
    g_launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "sq_kernel launch");
    return 0;
}

int nppsim_reset(const nppsim_state_t* st, const uint8_t* mask, int32_t mode, void* stream) {
    g_err[0] = 0;
    int rc = check_state(st);
    if (rc) return rc;
    if (mode < 0 || mode > 2) return fail(NPPSIM_E_ARG, "mode must be 0 (reset), 1 (fast_reset) or 2 (load on a fresh Simulator)");
    const int block = 128;
    reset_kernel<<<(st->n_envs + block - 1) / block, block, 0, (cudaStream_t)stream>>>(*st, mask, mode, 0, 0, nullptr);
    g_launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "reset_kernel launch");
    return 0;
}

int nppsim_reset_terminal(const nppsim_state_t* st, int32_t max_frames, int32_t mode, unsigned long long* stats, void* stream) {
    g_err[0] = 0;
    int rc = check_state(st);
    if (rc) return rc;
    if (mode < 0 || mode > 2) return fail(NPPSIM_E_ARG, "mode must be 0 (reset), 1 (fast_reset) or 2 (load on a fresh Simulator)");
    const int block = 128;
    reset_kernel<<<(st->n_envs + block - 1) / block, block, 0, (cudaStream_t)stream>>>(*st, nullptr, mode, 1, max_frames, stats);
    g_launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "reset_kernel launch");
    return 0;
}

int nppsim_observe(const nppsim_state_t* st, const uint8_t* path_dir, uint8_t* mask_out, float* ninja_state_out, void* stream) {
    g_err[0] = 0;
    int rc = check_state(st);
    if (rc) return rc;
    if (!mask_out && !ninja_state_out) return fail(NPPSIM_E_ARG, "no output buffer");
    obs_kernel<<<(st->n_envs + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*st, path_dir, mask_out, ninja_state_out, nullptr, nullptr);
    g_launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "obs_kernel launch");
    return 0;
}

int nppsim_set_exit_positions(const nppsim_state_t* st, const uint8_t* mask, const double* switch_xy, const double* door_xy,
                              int32_t* status, void* stream) {
    g_err[0] = 0;
    int rc = check_state(st);
    if (rc) return rc;
    if (!switch_xy || !door_xy) return fail(NPPSIM_E_ARG, "switch_xy / door_xy is NULL");
    curriculum_kernel<<<(st->n_envs + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*st, mask, switch_xy, door_xy, status);
    g_launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "curriculum_kernel launch");
    return 0;
}

int nppsim_step(const nppsim_state_t* st, const int8_t* hor, const int8_t* jump, int32_t input_ticks, int32_t n_ticks,
                const int32_t* tick_limit, int32_t stop_on_terminal, int32_t* frames_executed, void* stream) {
    g_err[0] = 0;
    return launch_step(st, hor, jump, nullptr, input_ticks, n_ticks, tick_limit, stop_on_terminal, frames_executed, stream);
}

int nppsim_step_actions(const nppsim_state_t* st, const uint8_t* actions, int32_t n_ticks, int32_t stop_on_terminal,
                        int32_t* frames_executed, void* stream) {
    g_err[0] = 0;
    if (!actions) return fail(NPPSIM_E_ARG, "actions is NULL");
    return launch_step(st, nullptr, nullptr, actions, 1, n_ticks, nullptr, stop_on_terminal, frames_executed, stream);
}

int nppsim_step_stream(const nppsim_state_t* st, const nppsim_input_stream_t* in, int32_t n_ticks, int32_t stop_on_terminal,
                       int32_t* frames_executed, void* stream) {
    g_err[0] = 0;
    int rc = check_stream(in);
    if (rc) return rc;
    return launch_step(st, nullptr, nullptr, nullptr, 1, n_ticks, nullptr, stop_on_terminal, frames_executed, stream, in);
}

int nppsim_replay_rows(const nppsim_state_t* st, const nppsim_input_stream_t* in, const int64_t* row_off, int32_t obs_tick,
                       int32_t frame_skip, float* game_state, double* pos, uint8_t* flags, int32_t* frame, void* stream) {
    g_err[0] = 0;
    int rc = check_state(st);
    if (rc) return rc;
    rc = check_stream(in);
    if (rc) return rc;
    if (in->encoding != NPPSIM_STREAM_REPLAY_BYTES || in->repeat != 1) return fail(NPPSIM_E_ARG, "replay rows need a replay-byte stream with repeat 1");
    if (!row_off || frame_skip < 1 || obs_tick < 1 || obs_tick % frame_skip != 0) return fail(NPPSIM_E_ARG, "bad row_off / frame_skip / obs_tick");
    replay_rows_kernel<<<(st->n_envs + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*st, *in, row_off, obs_tick, frame_skip, game_state, pos, flags, frame);
    g_launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "replay_rows_kernel launch");
    return 0;
}

int64_t nppsim_host_block_bytes(int32_t n_envs, int32_t with_obs) {
    const int64_t n = n_envs > 0 ? n_envs : 0;
    return n * (4 * 8 + 4) + (with_obs ? n * (40 * 4 + 6) : 0);
}

static int step_host_impl(const nppsim_state_t* st, const uint8_t* actions_host, uint8_t* actions_dev, int32_t n_ticks,
                          int32_t stop_on_terminal, uint8_t* staging_dev, uint8_t* out_host, int32_t with_obs, void* stream,
                          bool turnover, int32_t max_frames, int32_t reset_mode, unsigned long long* stats) {
    g_err[0] = 0;
    int rc = check_state(st);
    if (rc) return rc;
    if (!actions_host || !actions_dev || !staging_dev || !out_host) return fail(NPPSIM_E_ARG, "a host/device buffer is NULL");
    if (((uintptr_t)staging_dev & 7) != 0) return fail(NPPSIM_E_ARG, "staging_dev must be 8-byte aligned");
    if (n_ticks <= 0) return fail(NPPSIM_E_ARG, "n_ticks must be positive");
    // (everything is checked before anything is queued: an error return leaves the state as it was)
    if (turnover && (reset_mode < 0 || reset_mode > 2)) return fail(NPPSIM_E_ARG, "mode must be 0 (reset), 1 (fast_reset) or 2 (load on a fresh Simulator)");
    cudaStream_t s = (cudaStream_t)stream;
    const size_t N = (size_t)st->n_envs;
    // Pinned actions are read by the step kernel where they are (one byte per env as the env is loaded: mapped host memory)
    // -- no copy operation in front of the kernel; pageable actions are copied to actions_dev first.
    const uint8_t* actions_in = actions_dev;
    cudaError_t e = cudaSuccess;
    {
        const char* const ev_zc = getenv("NPPSIM_ZEROCOPY");
        cudaPointerAttributes attr;
        if ((!ev_zc || atoi(ev_zc) != 0) && cudaPointerGetAttributes(&attr, actions_host) == cudaSuccess && attr.type == cudaMemoryTypeHost && attr.devicePointer)
            actions_in = reinterpret_cast<const uint8_t*>(attr.devicePointer);
        else {
            cudaGetLastError();
            e = cudaMemcpyAsync(actions_dev, actions_host, N, cudaMemcpyHostToDevice, s);
            if (e != cudaSuccess) return cuda_fail(e, "actions H2D");
        }
    }
    // A pinned result block is mapped into the device's address space (unified addressing): the step kernel then stores
    // position / velocity / state of every env into it as the env finishes, and only the observations -- computed from the
    // final state by a second kernel -- are staged and copied.  A pageable block takes the staged path for everything, and
    // so does a state whose columns are permuted (ext_of): a warp's 32 results then go to 32 different places of the block,
    // 5 PCIe writes of 8 bytes per env instead of 5 of 256 per warp (131,072 envs of config 3: 1.19 against 1.10 ms staged;
    // in the caller's order 65,536 envs 0.720 against 0.768 ms).  NPPSIM_ZEROCOPY=0 / 2: never / also when permuted (A/B).
    double* host_pv = nullptr; int32_t* host_state = nullptr;
    {
        const char* const ev_zc = getenv("NPPSIM_ZEROCOPY");
        const int zc = ev_zc ? atoi(ev_zc) : 1;
        cudaPointerAttributes attr;
        if (zc && (!st->ext_of || zc == 2) && cudaPointerGetAttributes(&attr, out_host) == cudaSuccess && attr.type == cudaMemoryTypeHost && attr.devicePointer &&
            ((uintptr_t)attr.devicePointer & 7) == 0) {          // (f64 stores: an unaligned slice of a pinned buffer takes the staged path)
            host_pv = reinterpret_cast<double*>(attr.devicePointer);
            host_state = reinterpret_cast<int32_t*>(reinterpret_cast<uint8_t*>(attr.devicePointer) + N * 32);
        } else cudaGetLastError();                               // (an unregistered pointer is not an error here)
    }
    rc = launch_step(st, nullptr, nullptr, actions_in, 1, n_ticks, nullptr, stop_on_terminal, nullptr, stream, nullptr, host_pv, host_state);
    if (rc) return rc;
    if (!host_pv || with_obs) {
        // one pass over the batch packs what the caller reads every step, in the caller's env order, into one block: one D2H copy
        double* pv = host_pv ? nullptr : reinterpret_cast<double*>(staging_dev);
        int32_t* stt = host_pv ? nullptr : reinterpret_cast<int32_t*>(staging_dev + N * 32);
        float* obs = with_obs ? reinterpret_cast<float*>(staging_dev + N * 36) : nullptr;
        uint8_t* msk = with_obs ? staging_dev + N * 196 : nullptr;
        obs_kernel<<<(st->n_envs + 127) / 128, 128, 0, s>>>(*st, nullptr, msk, obs, pv, stt);
        g_launches++;
        e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(e, "obs_kernel launch");
        const size_t skip = host_pv ? N * 36 : 0;
        e = cudaMemcpyAsync(out_host + skip, staging_dev + skip, (size_t)nppsim_host_block_bytes(st->n_envs, with_obs) - skip, cudaMemcpyDeviceToHost, s);
        if (e != cudaSuccess) return cuda_fail(e, "result D2H");
    }
    if (turnover) {                                              // the block holds the step's results; the episode turnover queues behind them
        reset_kernel<<<(st->n_envs + 127) / 128, 128, 0, s>>>(*st, nullptr, reset_mode, 1, max_frames, stats);
        g_launches++;
        e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(e, "reset_kernel launch");
    }
    e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return cuda_fail(e, "stream sync");
    return 0;
}

int nppsim_step_host(const nppsim_state_t* st, const uint8_t* actions_host, uint8_t* actions_dev, int32_t n_ticks,
                     int32_t stop_on_terminal, uint8_t* staging_dev, uint8_t* out_host, int32_t with_obs, void* stream) {
    return step_host_impl(st, actions_host, actions_dev, n_ticks, stop_on_terminal, staging_dev, out_host, with_obs, stream, false, 0, 0, nullptr);
}

int nppsim_step_host_reset(const nppsim_state_t* st, const uint8_t* actions_host, uint8_t* actions_dev, int32_t n_ticks,
                           int32_t stop_on_terminal, uint8_t* staging_dev, uint8_t* out_host, int32_t with_obs,
                           int32_t max_frames, int32_t reset_mode, unsigned long long* stats, void* stream) {
    return step_host_impl(st, actions_host, actions_dev, n_ticks, stop_on_terminal, staging_dev, out_host, with_obs, stream, true, max_frames, reset_mode, stats);
}

// ---- nppsim_step_host_static: the host-buffer step for callers that cannot give the library device buffers ---------------
// nppsim_step_host needs two pieces of device memory that are neither input nor output: `actions_dev` (where pageable actions
// land) and `staging_dev` (where the pack kernel assembles the result block before its one copy to the host).  A caller
// that has no allocator to get them from -- see include/nppsim.h -- uses this entry point instead: both live in the
// file-scope array below, which every device that loads the library carries from load to exit.
//   [0, 202 * MAX)            result block of the current call, laid out for the call's own n_envs (nppsim_host_block_bytes)
//   [202 * MAX, 203 * MAX)    the call's actions
// Nothing here is accumulated from call to call: every byte a call copies out it has written itself first, so nothing needs
// clearing -- but the array keeps the last call's data, and two calls in flight on one device would write over each other.
#define STATIC_BLOCK_BYTES ((size_t)NPPSIM_STATIC_MAX_ENVS * 202)
#define STATIC_SCRATCH_BYTES ((size_t)NPPSIM_STATIC_MAX_ENVS * 203)
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

// The array's name means something only to device code and to the runtime's symbol calls; the address a kernel argument
// needs is asked for once per device, with that device current.  The same set-up gives the two step kernels this entry
// point launches the attributes launch_step gives them (same values: whichever runs first, the kernels end up configured
// alike) -- without launch_step's probe launch, which allocates.
static int static_scratch(unsigned char** out) {
    static std::mutex mu;
    static unsigned char* ptr_of[64] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && ptr_of[dev]) { *out = ptr_of[dev]; return 0; }
    unsigned char* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
    if (e != cudaSuccess || !p) return cuda_fail(e != cudaSuccess ? e : cudaErrorInvalidSymbol, "address of the static scratch array");
    const int resident = STEP_BLOCKS_PER_SM > 2 ? STEP_BLOCKS_PER_SM : 2;
    const int carve = (int)((resident * (STEP_SMEM + 1024) * 100 + 233471) / 233472) + 1;
    e = cudaFuncSetAttribute(sim_basic::step_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(sim_full::step_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(sim_basic::step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)STEP_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(sim_full::step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)STEP_SMEM);
    if (e != cudaSuccess) return cuda_fail(e, "step kernel attributes");
    if (dev >= 0 && dev < 64) ptr_of[dev] = p;                  // (a device number past the table is looked up on every call)
    *out = p;
    return 0;
}

int nppsim_step_host_static(const nppsim_state_t* st, const uint8_t* actions_host, int32_t n_ticks, int32_t stop_on_terminal,
                            uint8_t* out_host, int32_t with_obs, void* stream) {
    g_err[0] = 0;
    int rc = check_state(st);
    if (rc) return rc;
    if (!actions_host || !out_host) return fail(NPPSIM_E_ARG, "a host buffer is NULL");
    if (n_ticks <= 0) return fail(NPPSIM_E_ARG, "n_ticks must be positive");
    if (st->n_envs > NPPSIM_STATIC_MAX_ENVS) {
        snprintf(g_err, sizeof(g_err), "n_envs %d is more than the built-in scratch holds (NPPSIM_STATIC_MAX_ENVS = %d): split the batch, or "
                 "give nppsim_step_host buffers of your own", (int)st->n_envs, (int)NPPSIM_STATIC_MAX_ENVS);
        return NPPSIM_E_CAPACITY;
    }
    unsigned char* scratch = nullptr;
    rc = static_scratch(&scratch);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t N = (size_t)st->n_envs;
    uint8_t* const actions = scratch + STATIC_BLOCK_BYTES;
    cudaError_t e = cudaMemcpyAsync(actions, actions_host, N, cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return cuda_fail(e, "actions H2D");
    // every env in its home lane: the solo-warp queue needs launch_step's per-device SM table (results are the same either way)
    nppsim_state_t own = *st;
    own.handoff = nullptr;
    const bool basic = own.kind_mask != 0 && (own.kind_mask & ~NPP_BASIC_MASK) == 0;
    nppsim_input_stream_t sp;
    memset(&sp, 0, sizeof(sp));
    sp.repeat = 1;
    const int n_regular = (own.n_envs + STEP_BLOCK - 1) / STEP_BLOCK;
    auto kernel = basic ? sim_basic::step_kernel : sim_full::step_kernel;
    kernel<<<n_regular, STEP_BLOCK, STEP_SMEM, s>>>(own, nullptr, nullptr, actions, 1, n_ticks, nullptr, stop_on_terminal, nullptr, sp, 0, 0, n_regular, 0, nullptr, nullptr);
    g_launches++;
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "step_kernel launch");
    if (own.warp_order && own.warp_cost) {                      // keep the caller's longest-first order current, as nppsim_step_host does
        sched_kernel<<<1, 1024, 0, s>>>(own.warp_order, own.warp_cost, (own.n_envs + 31) / 32, nullptr, nullptr, 0, 0, 256);
        g_launches++;
        e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(e, "sched_kernel launch");
    }
    obs_kernel<<<(own.n_envs + 127) / 128, 128, 0, s>>>(own, nullptr, with_obs ? scratch + N * 196 : nullptr, with_obs ? reinterpret_cast<float*>(scratch + N * 36) : nullptr,
                                                         reinterpret_cast<double*>(scratch), reinterpret_cast<int32_t*>(scratch + N * 32));
    g_launches++;
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "obs_kernel launch");
    e = cudaMemcpyAsync(out_host, scratch, (size_t)nppsim_host_block_bytes(own.n_envs, with_obs), cudaMemcpyDeviceToHost, s);
    if (e != cudaSuccess) return cuda_fail(e, "result D2H");
    e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return cuda_fail(e, "stream sync");
    return 0;
}

}  // extern "C"
