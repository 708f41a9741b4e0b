/*
 * nppsim.h -- C ABI of libnppsim.so, the B200 batched N++ physics engine.
 *
 * Drop-in boundary for the reference's Simulator hot path (paths are under
 * /root/reference/nclone/).  The reference has no FFI layer: the seam is the Python
 * class nsim.Simulator behind the facade nplay_headless.NPlayHeadless, so each entry
 * point below names the Simulator method it replaces for a whole batch of envs.
 * The Python binding a maintainer would add is a ctypes stub.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types cross this boundary;
 *   - all buffers are caller-owned (the library never allocates device memory);
 *   - every function returns 0 on success, <0 for an argument / map error (NPPSIM_E_*),
 *     >0 for a CUDA runtime error code; nppsim_last_error() describes the last failure
 *     on the calling thread;
 *   - launches are asynchronous on the given cudaStream_t (passed as void*); the library keeps
 *     no simulation state of its own -- the only process-wide variables are the atomic launch
 *     counter behind nppsim_launch_count() and per-device one-time kernel attributes -- so one
 *     instance per device / stream is safe.
 */
#ifndef NPPSIM_H
#define NPPSIM_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NPPSIM_VERSION 102

#define NPPSIM_E_ARG      (-1)
#define NPPSIM_E_MAP      (-2)   /* malformed map_data (the reference raises KeyError/IndexError/ValueError) */
#define NPPSIM_E_CAPACITY (-3)   /* entity / slot capacity exceeded: an error at load, never truncation */
#define NPPSIM_E_NOGPU    (-4)

/* Canonical per-env ninja record: NPPSIM_NF doubles + NPPSIM_NI int32, structure-of-arrays
 * with stride n_envs: field f of env i lives at nin_f[f * n_envs + i].  Field order =
 * Ninja.__init__ (ninja.py:80-196); names in nclone_b200/fields.py. */
#define NPPSIM_NF 31
#define NPPSIM_NI 27
/* per-env bookkeeping ints (env_i[k * n_envs + i]) */
#define NPPSIM_ENV_STAMP    0   /* next grid_entity insertion stamp (entities.py:256-264 order) */
#define NPPSIM_ENV_HOT      1   /* count of toggling mines + touched boost pads */
#define NPPSIM_ENV_EVENTS   2   /* NPPSIM_EV_* bits accumulated since the caller last cleared them */
#define NPPSIM_ENV_SWFRAME  3   /* sim.frame at which the exit switch was taken, -1 if not yet */
#define NPPSIM_ENV_ORDER    4   /* 0: grid lists in creation order (load/reset); 1: entity_dic order (fast_reset) */
#define NPPSIM_ENV_TICKS   5   /* ticks executed since load_maps (wraps at 2^31), for throughput accounting */
#define NPPSIM_ENV_WEDGE   6   /* == the tag of a handoff queue (nppsim_state_t.handoff): the env is among that queue's initial
                                  entries and starts the next launch on a solo warp; written only between launches */
#define NPPSIM_ENV_RELOADS 7   /* Simulator.reset() calls since the level was loaded on a fresh Simulator: the reference numbers entities
                                  from counters that live as long as the Simulator (entities.py:213-216), so an entity's `index` is
                                  its index in the level + RELOADS * (entities of its type in the level); death balls stop repelling
                                  each other once it is non-zero (entity_death_ball.py:153-166) */
#define NPPSIM_NENV 8

#define NPPSIM_EV_EXIT_SWITCH 1  /* entity_exit_switch.py:128-129 -> gym_env.invalidate_switch_cache() */
#define NPPSIM_EV_LOCKED_DOOR 2  /* entity_door_locked.py:66-67 */
#define NPPSIM_EV_WIN         4
#define NPPSIM_EV_DEATH       8

/* dynamic slot layout: dyn_f[(slot*4 + k) * n_envs + i], dyn_i[(slot*3 + k) * n_envs + i] */
#define NPPSIM_DYN_F 4          /* xpos, ypos, a0, a1 */
#define NPPSIM_DYN_I 3          /* cell (cx*25+cy, -1 = not in grid_entity), stamp, aux */

typedef struct {
    int32_t n_envs;             /* N, also the SoA stride */
    int32_t ent_cap;            /* rows of ent_flag  (>= max n_ent over the loaded maps) */
    int32_t dyn_cap;            /* dynamic slots per env (>= max n_dyn over the loaded maps) */
    int32_t n_maps;
    double*  nin_f;             /* [NPPSIM_NF][N] */
    int32_t* nin_i;             /* [NPPSIM_NI][N] */
    int32_t* env_i;             /* [NPPSIM_NENV][N] */
    uint8_t* ent_flag;          /* [ent_cap][N]  per-entity byte: active / state / closed bits */
    double*  dyn_f;             /* [dyn_cap*4][N] */
    int32_t* dyn_i;             /* [dyn_cap*3][N] */
    const int32_t*  env_map;    /* [N] index into map_off */
    const uint8_t*  map_blob;   /* all compiled maps, concatenated (device) */
    const int64_t*  map_off;    /* [n_maps] byte offset of each map inside map_blob (device) */
    uint32_t kind_mask;         /* OR of NppMapHeader.kind_mask over the loaded maps (entity-kind census); 0 = unknown.
                                   Selects the step-kernel variant: level sets without moving / pushing entities run
                                   on a kernel compiled without those phases. */
    uint32_t stride;            /* SoA stride in elements; 0 = n_envs.  A sub-batch (envs [b, b + n) of a batch of S envs, b % 32 == 0) is the same
                                   struct with n_envs = n, stride = S and every per-env pointer advanced by b (warp_order / warp_cost by b / 32):
                                   groups of one batch can then be stepped independently on different streams. */
    int32_t*  warp_order;       /* optional [ceil(N/32)]: launch slot -> group of 32 consecutive envs (a permutation);
                                   NULL = identity.  Rewritten by nppsim_step* after each launch when warp_cost is set. */
    uint32_t* warp_cost;        /* optional [ceil(N/32)]: device-clock cost of each group in the last launch (written by
                                   the step kernel), the key of the longest-first order */
    const int32_t* ext_of;      /* optional [N]: state column i holds the caller's env ext_of[i] (a permutation).  The engine
                                   may keep envs in any order -- BatchedSimulator puts the envs of a level next to each other, so
                                   that the lanes of a warp walk the same level tables and run the same entity loops -- while
                                   EVERY per-env array that crosses this interface (actions, hor / jump rows, tick_limit,
                                   frames_executed, masks, stream offsets, row_off, switch_xy / door_xy, status, observation
                                   outputs, nppsim_step_host's block) stays in the caller's env order.  NULL = identity. */
    int32_t*  handoff;          /* optional [NPPSIM_HANDOFF_WORDS]: queue through which envs whose ninja is wedged in a corner (all
                                   32 depenetration iterations of every substep, ninja.py:303-379) leave their home lane and run
                                   their ticks alone on "solo" warps -- on SMs reserved for them from one full wave of blocks
                                   up, else in blocks at the front of the grid -- from the start of a launch when their previous
                                   tick was wedged, else from the end of their first wedged substep.  The caller
                                   zero-initialises it once and sets word [6] to a non-zero tag unique among the queues used
                                   on the same env arrays (sub-batches: one queue each); the library maintains the rest.
                                   NULL: every env stays in its home lane.  Results are identical either way. */
} nppsim_state_t;
#define NPPSIM_HANDOFF_CAP   4096
#define NPPSIM_HANDOFF_HDR   16     /* [0] entries reserved, [1] tickets taken, [2] regular blocks finished (per launch); [3] error
                                       bits, [4] mid-launch handoffs so far (both sticky); [5] length of the next launch's list;
                                       [6] tag (caller); [7] initial entries of the current launch; [8] regular block slots claimed,
                                       [9] initial entries claimed, [10] SMs reserved for solo blocks, [11] blocks that went solo on
                                       such an SM (all four per launch; "solo SMs", csrc/sim_variant.inc); [12] launches after which more envs
                                       asked for a solo start than the queue admits, [13] launches (both sticky: when [12] / [13] is
                                       large the wedged envs are the bulk of the work, not its tail, and the caller should step
                                       without a queue); [14..15] reserved */
#define NPPSIM_HANDOFF_WORDS (NPPSIM_HANDOFF_HDR + 2 * NPPSIM_HANDOFF_CAP)

/* Per-env input streams for nppsim_step_stream: env i owns data[offset[i] .. offset[i+1]) -- ragged, no padding. */
#define NPPSIM_STREAM_ACTIONS      0   /* Discrete(6) actions, ACTION_TO_INPUTS (action_replayer.py:30-37); > 5 -> NOOP (:236-240) */
#define NPPSIM_STREAM_REPLAY_BYTES 1   /* .replay input bytes: bit0 jump, bit1 right, bit2 left, both -> 0 (replay/replay_executor.py:61-86) */
typedef struct {
    const uint8_t* data;        /* device: all envs' entries, concatenated */
    const int64_t* offset;      /* device [N+1] */
    int32_t encoding;           /* NPPSIM_STREAM_* */
    int32_t repeat;             /* ticks each entry is held for (ActionReplayer.frame_skip, action_replayer.py:104-107); >= 1 */
    int64_t start_tick;         /* position (in ticks) inside every env's stream of the launch's first tick: chunked execution */
} nppsim_input_stream_t;

int         nppsim_version(void);
const char* nppsim_last_error(void);

/* ---- host side: Simulator.load's static part (nsim.py:51-56, map_loader.py:18-145) ---------
 * map_data is the map_loader format as float64 (generators emit non-integer coordinates):
 * [184:1150] 42x23 tile ids, [1156] exit-door count, [1200] death-ball count,
 * [1231],[1232] spawn, [1235:] entity records.  Writes one relocatable blob
 * (nclone_b200/csrc/map_blob.h) to `out`; upload it anywhere 16-byte aligned. */
int64_t nppsim_map_blob_bound(int32_t map_len);
int     nppsim_compile_map(const double* map_data, int32_t map_len, uint8_t* out, int64_t out_cap,
                           int64_t* out_used, int32_t* n_entities, int32_t* n_dyn_slots);

/* nppsim_compile_maps_device: the same compile for a whole batch of levels in ONE launch on the device
 * (replaces per-level Simulator.load work: map_loader.py:18-145,
 * utils/tile_segment_factory.py:169-344, utils/spatial_segment_index.py:69-110, nsim.py:194-219).
 *   map_data    device float64 [n_maps][row_stride], map_loader format;
 *   map_len     optional device int32 [n_maps] valid values per row (NULL: default_len for every row);
 *   blobs       device, 16-byte aligned, n_maps * blob_stride bytes; level m's blob starts at m * blob_stride and is
 *               byte-identical to nppsim_compile_map's output for the same row (blob_stride >=
 *               nppsim_map_blob_bound(longest row), multiple of 16);
 *   map_off     optional device int64 [n_maps], receives m * blob_stride (ready for nppsim_state_t.map_off);
 *   status      device int32 [n_maps]: 0 or the NPPSIM_E_* code nppsim_compile_map would return for that row (a failed
 *               level's blob has magic 0); levels with more than NPPSIM_DEVICE_COMPILE_MAX_ENT entities -> NPPSIM_E_CAPACITY;
 *   summary     device int32 [4]: {max n_entities, max n_dyn_slots, OR of kind_mask, number of failed levels}. */
#define NPPSIM_DEVICE_COMPILE_MAX_ENT 1024
int nppsim_compile_maps_device(const double* map_data, int64_t row_stride, const int32_t* map_len, int32_t default_len,
                               int32_t n_maps, uint8_t* blobs, int64_t blob_stride, int64_t* map_off, int32_t* status,
                               int32_t* summary, void* stream);

/* ---- device side ------------------------------------------------------------------------
 * nppsim_reset, for every env whose mask byte is non-zero (mask == NULL: all envs; mask is a device pointer):
 *   mode 0  Simulator.reset() (nsim.py:62-76) on the Simulator that already holds the level -- the call NPlayHeadless.reset
 *           (nplay_headless.py:272-276) and ActionReplayer.replay_to_checkpoint (action_replayer.py:99) make.  The reference
 *           never clears sim.entity_counts, so the re-created entities' `index` keeps counting (NPPSIM_ENV_RELOADS), and
 *           with it death balls stop repelling each other from the second episode on.
 *   mode 1  fast_reset() (nsim.py:78-140).
 *   mode 2  Simulator() + load() (nsim.py:51-56): the dynamic part of a load on a fresh Simulator, indices from 0.  Use it
 *           once after the state arrays are allocated (zero-filled state + mode 0 would leave RELOADS at 1). */
int nppsim_reset(const nppsim_state_t* st, const uint8_t* mask, int32_t mode, void* stream);

/* nppsim_reset_terminal: the env wrapper's episode turnover for a whole batch -- reset (mode as above)
 * exactly those envs whose ninja has won or died (state 6/7/8) or whose sim.frame >= max_frames
 * (max_frames <= 0: no truncation).  stats, optional device uint64[4], is incremented atomically by
 * {episodes ended, wins, deaths, truncations}. */
int nppsim_reset_terminal(const nppsim_state_t* st, int32_t max_frames, int32_t mode,
                          unsigned long long* stats, void* stream);

/* nppsim_set_exit_positions: the curriculum's mutation from outside tick -- IntermediateGoalManager.apply_to_simulator
 * (gym_environment/reward_calculation/intermediate_goal_manager.py:698-960) for every env whose mask byte is non-zero
 * (mask == NULL: all).  The last exit switch of entity_dic[3] moves to switch_xy and is re-armed while sim.frame <= 10;
 * its parent door moves to door_xy; cells and grid_entity order follow the reference (a re-celled switch is re-appended,
 * a re-celled door leaves the grid until the switch is taken).  A later reset / fast_reset restores the level's own positions.
 *   switch_xy, door_xy  device float64 [2][N] (row 0 x, row 1 y);
 *   status              optional device int32 [N]: 0, or -1 for a masked env whose level has no exit switch (the reference
 *                       raises RuntimeError); unmasked envs are left untouched.
 * Not modelled: re-arming a switch that was already taken in the first 10 frames and taking it again appends the door to
 * its cell list a second time in the reference; here the door stays listed once (the physics is identical -- the door's
 * collision is idempotent -- only later list positions in that cell differ). */
int nppsim_set_exit_positions(const nppsim_state_t* st, const uint8_t* mask, const double* switch_xy, const double* door_xy,
                              int32_t* status, void* stream);

/* nppsim_step: Simulator.tick (nsim.py:221-292) n_ticks times for every env in one launch.
 *   hor, jump   device int8, hor in {-1,0,1}, jump in {0,1}; input_ticks == 1: one value per
 *               env held for all ticks ([N]); input_ticks == n_ticks: per-tick rows ([n_ticks][N]).
 *   tick_limit  optional device int32 [N]: env i runs min(n_ticks, tick_limit[i]) ticks
 *               (ragged replay batches); NULL = no limit.
 *   stop_on_terminal  non-zero: an env stops after the tick that leaves it won or dead
 *               (gym_environment/base_environment.py:605-609 frame-skip early exit).
 *   frames_executed   optional device int32 [N]: ticks actually run. */
int nppsim_step(const nppsim_state_t* st, const int8_t* hor, const int8_t* jump, int32_t input_ticks,
                int32_t n_ticks, const int32_t* tick_limit, int32_t stop_on_terminal,
                int32_t* frames_executed, void* stream);

/* nppsim_step_actions: same, inputs given as Discrete(6) actions [N] (uint8) held for n_ticks,
 * decoded with ACTION_TO_INPUTS (action_replayer.py:30-37 == base_environment.py:384-402). */
int nppsim_step_actions(const nppsim_state_t* st, const uint8_t* actions, int32_t n_ticks,
                        int32_t stop_on_terminal, int32_t* frames_executed, void* stream);

/* nppsim_step_stream: up to n_ticks ticks with every env reading its own input stream -- the hot loop of
 * ActionReplayer.replay_to_checkpoint / replay_partial (action_replayer.py:79-163: reset, then frame_skip ticks per
 * stored action) and of ReplayExecutor.execute_replay (replay/replay_executor.py:431-436) for a whole ragged batch in
 * one launch.  Env i runs min(n_ticks, len_i * repeat - start_tick) ticks (0 once its stream is exhausted). */
int nppsim_step_stream(const nppsim_state_t* st, const nppsim_input_stream_t* in, int32_t n_ticks, int32_t stop_on_terminal,
                       int32_t* frames_executed, void* stream);

/* nppsim_replay_rows: one dataset row per env whose replay is still running -- the simulator-derived part of
 * ReplayExecutor._get_raw_observation (replay/replay_executor.py:470-600) written straight into ragged per-replay
 * tables.  Call after the step that executed stream ticks [.., obs_tick]; env i (replay of
 * in->offset[i+1]-in->offset[i] input bytes, in->repeat = 1) writes row row_off[i] + k, k = obs_tick / frame_skip - 1,
 * iff obs_tick <= its length:
 *   game_state [rows][41] f32   get_ninja_state()[:40] + time_remaining = max(0, (len - sim.frame) / len)   (:530-545)
 *   pos        [rows][6]  f64   player x,y | exit switch x,y | exit door x,y   (nplay_headless.py:415, 577-616)
 *   flags      [rows][4]  u8    switch_activated, player_won, player_dead, action = map_input_to_action(byte of the
 *                               row's last tick) (:42-58, :448)
 *   frame      [rows]     i32   frame_idx = obs_tick - 1 (:453)
 * Any output pointer may be NULL. */
int nppsim_replay_rows(const nppsim_state_t* st, const nppsim_input_stream_t* in, const int64_t* row_off, int32_t obs_tick,
                       int32_t frame_skip, float* game_state, double* pos, uint8_t* flags, int32_t* frame, void* stream);

/* nppsim_step_host: the call a host-resident caller makes (one NppEnvironment.step worth of physics -- and, with
 * with_obs != 0, of the observations the wrapper pulls every step, base_environment.py:2822, 3107 -- for the batch):
 * runs nppsim_step_actions on `actions_host` [N] and delivers the results in the CALLER's env order in `out_host`, then
 * synchronises the stream.  How the bytes travel depends on the host buffers:
 *   - pinned `actions_host` is read in place by the step kernel (mapped host memory, no copy operation; `actions_dev` is then
 *     unused); pageable actions are copied to `actions_dev` first;
 *   - when `out_host` is pinned (page-locked, 8-byte aligned) and the state is not permuted (ext_of == NULL), the step kernel
 *     stores position / velocity / state into it directly as each env finishes (no pack, no copy for that part), and only the
 *     observations (with_obs) are packed into `staging_dev` and copied;
 *   - otherwise a pack kernel assembles the whole block in `staging_dev` and ONE device-to-host copy moves it.
 * All arguments are checked before anything is queued.
 * Block layout (nppsim_host_block_bytes(N, with_obs) bytes; staging_dev 8-byte aligned, out_host pinned for speed):
 *   [0, 32N)      f64 [4][N]   ninja xpos, ypos, xspeed, yspeed
 *   [32N, 36N)    i32 [N]      ninja state
 *   [36N, 196N)   f32 [N][40]  NPlayHeadless.get_ninja_state (nplay_headless.py:735-917)        -- with_obs only
 *   [196N, 202N)  u8  [N][6]   Ninja.get_valid_action_mask (ninja.py:628-839), no path hints      -- with_obs only */
int64_t nppsim_host_block_bytes(int32_t n_envs, int32_t with_obs);
int nppsim_step_host(const nppsim_state_t* st, const uint8_t* actions_host, uint8_t* actions_dev,
                     int32_t n_ticks, int32_t stop_on_terminal, uint8_t* staging_dev, uint8_t* out_host,
                     int32_t with_obs, void* stream);
/* nppsim_step_host_reset: nppsim_step_host followed by nppsim_reset_terminal(max_frames, reset_mode, stats) in the same call --
 * a vectorised wrapper's whole step: the block receives the step's results (the terminal observation of an env that just won
 * or died included), the episode turnover is queued behind it on the stream, and ONE synchronisation ends the call, so the
 * device does not idle while the host comes back for the reset (base_environment.py: step(), then reset() of the done envs). */
int nppsim_step_host_reset(const nppsim_state_t* st, const uint8_t* actions_host, uint8_t* actions_dev,
                           int32_t n_ticks, int32_t stop_on_terminal, uint8_t* staging_dev, uint8_t* out_host,
                           int32_t with_obs, int32_t max_frames, int32_t reset_mode, unsigned long long* stats, void* stream);

/* nppsim_step_host_static: nppsim_step_host for a caller that has no way to allocate device memory.  (No counterpart in the
 * reference -- it has no device memory to manage; this exists for the hosts the library gets embedded in.)
 * Who that is: a ctypes / cffi stub that loads this library but links no CUDA runtime of its own to call cudaMalloc with; an
 * actor process next to a trainer or inference server that has claimed the card's memory up front for its own pool, where a
 * late cudaMalloc fails or is against the deployment's rules; a step loop that must not allocate because cudaMalloc
 * synchronises the whole device under the other streams' feet.  Such a caller still owns the simulation state it passes in
 * `st` and the two host buffers; what it cannot supply is nppsim_step_host's pair of scratch buffers (`actions_dev`,
 * `staging_dev`).  Here they are a fixed array inside the library image instead: NPPSIM_STATIC_MAX_ENVS * 203 bytes
 * (3.2 MiB) of every GPU that loads the library, held until the process exits whether or not this function is ever called.
 *   - Same arguments, same block layout, same results bit for bit as nppsim_step_host without `actions_dev` / `staging_dev`;
 *     host buffers may be pinned or pageable (both are copied; nothing is read or written in place).  st->handoff is ignored:
 *     every env runs in its home lane (results do not depend on it).
 *   - n_envs > NPPSIM_STATIC_MAX_ENVS: returns NPPSIM_E_CAPACITY before anything is launched; split the batch into sub-batch
 *     views (nppsim_state_t.stride) and call once per view, one after the other.
 *   - The function calls no cudaMalloc / cudaFree and nothing that allocates on the device.  Its FIRST call on each device
 *     looks up the array's device address and sets the step kernels' launch attributes (under a lock, with that device
 *     current); make that first call outside any stream capture.
 *   - All calls on one device share the one array.  It needs no clearing (a call writes every byte it copies out) and it keeps
 *     the previous call's data.  The function returns only after its stream has drained, so calls made one after the other are
 *     safe; calls that overlap on the GPU -- which takes two host threads stepping on the same device at once -- overwrite each
 *     other's actions and results.  Serialise them, or give each thread buffers of its own through nppsim_step_host. */
#define NPPSIM_STATIC_MAX_ENVS 16384
int nppsim_step_host_static(const nppsim_state_t* st, const uint8_t* actions_host, int32_t n_ticks, int32_t stop_on_terminal,
                            uint8_t* out_host, int32_t with_obs, void* stream);

/* nppsim_observe: the two per-step observation primitives the env wrapper calls.
 *   ninja_state_out  optional device float32 [N][40]: NPlayHeadless.get_ninja_state (nplay_headless.py:735-917);
 *   mask_out         optional device uint8 [N][6]: Ninja.get_valid_action_mask (ninja.py:628-839), masking mode
 *                    "hard"; path_dir optional device uint8 [N] = sim._path_straightness_direction
 *                    (0 None, 1 "left", 2 "right", 3 "down"; NULL = None). */
int nppsim_observe(const nppsim_state_t* st, const uint8_t* path_dir, uint8_t* mask_out, float* ninja_state_out, void* stream);

/* Test hooks.  nppsim_debug_sq: out[i] = the device's SQ(in[i]) -- the reference's `x ** 2`, i.e. glibc
 * pow(x, 2.0) (ninja.py:412 and ~45 other sites) -- for cross-checking against the host libm.
 * nppsim_sq_mode: 1 = exact pow emulation (default build), 0 = x*x. */
int nppsim_debug_sq(const double* in_dev, double* out_dev, int32_t n, void* stream);
int nppsim_sq_mode(void);
/* Test hook: counts (into *mismatches, device uint64, not cleared) the i for which the kernels' shared-reciprocal double division
 * div2(a1[i], a2[i], b[i]) differs in any bit from the device's own a1[i] / b[i], a2[i] / b[i]. */
int nppsim_debug_div2(const double* a1_dev, const double* a2_dev, const double* b_dev, unsigned long long* mismatches_dev, int32_t n, void* stream);
/* (To run a basic level set on the full kernel -- the variant cross-check of the tests -- pass kind_mask = 0, "census unknown".) */

/* Kernel launch count since library load (bench.py's gpu_launches claim). */
int64_t nppsim_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif
