"""CUDA path (through the C ABI of include/nppsim.h) against the reference's own outputs
(tests/golden, produced by the unmodified reference) and against the oracle.
Bar: bit-exact -- every double compared by bit pattern, every int by value."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from oracle import oracle, schema  # noqa: E402  (checker only)


def _run_batch_vs_golden(G, idx):
    """One env per golden trajectory, ragged lengths; one launch per frame, full canonical record
    of every still-running env hashed and compared against the reference's digest."""
    import torch
    from nclone_b200 import BatchedSimulator
    trajs = [G.traj(i) for i in idx]
    N = len(trajs)
    sim = BatchedSimulator(N)
    sim.load_maps([t["map_data"] for t in trajs], np.arange(N))
    lens = np.array([len(t["hor"]) for t in trajs])
    T = int(lens.max())
    hor = np.zeros((T, N), dtype=np.int8); jump = np.zeros((T, N), dtype=np.int8)
    for k, t in enumerate(trajs):
        hor[:lens[k], k] = t["hor"]; jump[:lens[k], k] = t["jump"]
    hor_d = torch.from_numpy(hor).cuda(); jump_d = torch.from_numpy(jump).cuda()
    bad = []
    frames = 0

    def apply_moves(f):          # curriculum fixtures: apply_to_simulator before tick f for the envs that have one
        hit = [(k, m) for k, t in enumerate(trajs) for m in t["moves"] if int(m[0]) == f]
        if hit:
            sxy = np.zeros((2, N)); dxy = np.zeros((2, N)); mask = np.zeros(N, dtype=np.uint8)
            for k, m in hit:
                sxy[:, k] = m[1:3]; dxy[:, k] = m[3:5]; mask[k] = 1
            sim.set_exit_positions(sxy, dxy, mask)

    apply_moves(0)
    recs = sim.canonical_state()
    for k in range(N):
        if schema.state_hash(*recs[k]) != trajs[k]["hash"][0]:
            bad.append((trajs[k]["name"], 0))
    alive = np.ones(N, dtype=bool)
    for f in range(T):
        limit = (lens > f).astype(np.int32)
        if f > 0:
            apply_moves(f)
        sim.tick(hor_d[f], jump_d[f], n_ticks=1, tick_limit=limit)
        ids = np.nonzero((lens > f) & alive)[0]
        recs = sim.canonical_state(ids)
        for k, rec in zip(ids, recs):
            frames += 1
            if schema.state_hash(*rec) != trajs[k]["hash"][f + 1]:
                bad.append((trajs[k]["name"], f + 1))
                alive[k] = False           # report the first differing frame only
            else:
                assert rec[0][0] == trajs[k]["pv"][f + 1, 0]
    return frames, bad, sim, trajs


def test_bc_replays_bit_exact(golden_replays):
    """BASELINE.json config 5: every bc_replays trajectory, every frame, full state."""
    frames, bad, sim, trajs = _run_batch_vs_golden(golden_replays, range(len(golden_replays)))
    assert not bad, f"{len(bad)} replays differ; first: {bad[:5]}"
    assert frames == 31461
    won = (sim.ninja_state().cpu().numpy() == 8).sum()
    assert won == 128


def test_generated_maps_random_actions_bit_exact(golden_generated):
    frames, bad, _, _ = _run_batch_vs_golden(golden_generated, range(len(golden_generated)))
    assert not bad, f"{len(bad)} episodes differ; first: {bad[:5]}"
    assert frames == 40800


def test_doors_and_scenario_maps_bit_exact(golden_doors):
    """SURVEY e5 (regular door, entity_door_regular.py:38-63) + the thin rows e7 / e16 + the reference's own scenario maps:
    163 episodes from the unmodified reference (oracle/gen_golden_doors.py) -- type-5 doors injected on the ninja's path
    (all eight orientations) and across drone / thwump paths, so open / re-close and the grid-edge toggling
    (entity_door_base.py:70-108 -> is_empty_row / is_empty_column) are both exercised; trap-door closures; shove-thwump
    activations; nclone/test_maps (17), nclone/maps/test-maps (5), bc_replays_tmp (8).  Every frame, full record."""
    g = golden_doors
    frames, bad, sim, trajs = _run_batch_vs_golden(g, range(len(g)))
    assert not bad, f"{len(bad)} episodes differ; first: {bad[:5]}"
    assert frames == 52396
    st = dict(zip((str(n) for n in g.d["stats_names"]), (int(v) for v in g.d["stats"])))
    assert st["door_reclosed"] >= 20 and st["trap_closed"] >= 10 and st["shwump_activated"] >= 10
    assert sim.kind_mask & (1 << 5)                      # the regular-door code of the full kernel ran


def test_curriculum_repositioning_bit_exact(golden_curriculum, golden_replays):
    """SURVEY 8(b) "mutation from outside tick": nppsim_set_exit_positions against the reference's own
    IntermediateGoalManager.apply_to_simulator (fixtures from the unmodified reference), every frame, full state."""
    import torch
    from nclone_b200 import BatchedSimulator
    frames, bad, sim, trajs = _run_batch_vs_golden(golden_curriculum, range(len(golden_curriculum)))
    assert not bad, f"{len(bad)} episodes differ; first: {bad[:5]}"
    assert frames == 16800
    # a level without an exit switch refuses, like the reference (RuntimeError)
    md = golden_replays.traj(0)["map_data"].copy(); md[1235:] = 0
    one = BatchedSimulator(1); one.load_maps([md])
    with pytest.raises(RuntimeError):
        one.set_exit_positions(torch.zeros(2, 1, dtype=torch.float64), torch.zeros(2, 1, dtype=torch.float64))
    # reset restores the level's own positions
    first = golden_curriculum.traj(0)
    two = BatchedSimulator(1); two.load_maps([first["map_data"]])
    before = two.canonical_state()
    m = first["moves"][0]
    two.set_exit_positions(np.array([[m[1]], [m[2]]]), np.array([[m[3]], [m[4]]]))
    assert schema.state_hash(*two.canonical_state()[0]) == first["hash"][0]
    two.reset()
    after = two.canonical_state()[0]
    ref = oracle.OracleSim(); ref.load(first["map_data"]); ref.reset()       # Simulator.reset() re-creates the entities: `index` counts on
    assert schema.state_hash(*after) == schema.state_hash(*ref.state())
    (nf0, ni0, ef0, ei0), (nf1, ni1, ef1, ei1) = before[0], after                 # and everything but `index` is as after the load
    assert np.array_equal(nf0, nf1) and np.array_equal(ni0, ni1) and np.array_equal(ef0, ef1)
    assert np.array_equal(np.delete(ei0, 1, axis=1), np.delete(ei1, 1, axis=1)) and not np.array_equal(ei0[:, 1], ei1[:, 1])


def test_first_replay_known_answer(golden_replays):
    """spawn (828,540); after 37 inputs (812.0036311681246, 542.0), state 8."""
    import torch
    from nclone_b200 import BatchedSimulator
    t = golden_replays.traj(0)
    sim = BatchedSimulator(1)
    sim.load_maps([t["map_data"]])
    x, y = sim.ninja_position()
    assert (x.item(), y.item()) == (828.0, 540.0)
    sim.tick(torch.from_numpy(t["hor"][:, None].copy()), torch.from_numpy(t["jump"][:, None].copy()), n_ticks=37)
    x, y = sim.ninja_position()
    assert (x.item(), y.item()) == (812.0036311681246, 542.0)
    assert bool(sim.ninja_has_won().item()) and sim.frame().item() == 37
    assert bool(sim.exit_switch_activated().item())


def test_fused_frame_skip_vs_oracle(golden_replays, golden_generated):
    """Config-2/3 shape: many envs, Discrete(6) actions held for 4 fused ticks with the env's
    early exit on win/death; ninja record and frames_executed compared with the oracle."""
    import torch
    from nclone_b200 import BatchedSimulator, ACTION_TO_INPUTS
    maps = [golden_generated.traj(i)["map_data"] for i in range(0, len(golden_generated), 2)]
    maps += [golden_replays.traj(i)["map_data"] for i in range(0, 130, 5)]
    M = len(maps)
    reps = 8
    N = M * reps
    rng = np.random.default_rng(5)
    steps = 60
    actions = rng.integers(0, 6, size=(steps, N)).astype(np.uint8)
    sim = BatchedSimulator(N)
    sim.load_maps(maps, np.arange(N) % M)
    osims = []
    for k in range(N):
        s = oracle.OracleSim(); s.load(maps[k % M]); osims.append(s)
    done = np.zeros(N, dtype=bool)
    for st in range(steps):
        fe = sim.step(torch.from_numpy(actions[st]).cuda(), frame_skip=4, stop_on_terminal=True).cpu().numpy()
        recs = sim.canonical_state()
        for k in range(N):
            h, j = ACTION_TO_INPUTS[actions[st, k]]
            n = 0
            for _ in range(4):
                osims[k].tick(h, j); n += 1
                if osims[k].state()[1][0] in (6, 7, 8):
                    break
            assert fe[k] == n, (k, st)
            ref = osims[k].state()
            assert schema.state_hash(*ref) == schema.state_hash(*recs[k]), (k, st, maps and k % M)


def test_reset_and_fast_reset(golden_replays):
    import torch
    from nclone_b200 import BatchedSimulator
    idx = [0, 1, 5, 40, 77]
    trajs = [golden_replays.traj(i) for i in idx]
    sim = BatchedSimulator(len(idx))
    sim.load_maps([t["map_data"] for t in trajs])
    first = sim.canonical_state()
    rng = np.random.default_rng(0)
    for _ in range(50):
        sim.step(torch.from_numpy(rng.integers(0, 6, len(idx)).astype(np.uint8)).cuda(), 4, False)
    sim.reset()
    # Simulator.reset() on the Simulator that loaded the level: the initial state again, entity indices counting on
    # (entities.py:213-216) -- the oracle's reset() is pinned to the reference's on exactly this (tests/golden/reset_reuse.npz)
    osims = []
    for t in trajs:
        s = oracle.OracleSim(); s.load(t["map_data"]); s.reset(); osims.append(s)
    for a, b, s in zip(first, sim.canonical_state(), osims):
        assert schema.state_hash(*b) == schema.state_hash(*s.state())
        assert np.array_equal(a[0].view(np.int64), b[0].view(np.int64)) and np.array_equal(a[1], b[1])
        assert np.array_equal(a[2].view(np.int64), b[2].view(np.int64)) and np.array_equal(np.delete(a[3], 1, axis=1), np.delete(b[3], 1, axis=1))
    sim.reset(mode="load")                                # a fresh Simulator loading the level: indices from 0 again
    for a, b in zip(first, sim.canonical_state()):
        assert schema.state_hash(*a) == schema.state_hash(*b)
    sim.reset()
    # fast_reset against the oracle's fast_reset after the same history
    acts = rng.integers(0, 6, size=(40, len(idx))).astype(np.uint8)
    from nclone_b200 import ACTION_TO_INPUTS
    for st in range(40):
        sim.step(torch.from_numpy(acts[st]).cuda(), 4, False)
        for k, s in enumerate(osims):
            h, j = ACTION_TO_INPUTS[acts[st, k]]
            for _ in range(4):
                s.tick(h, j)
    sim.fast_reset()
    for s in osims:
        s.fast_reset()
    for st in range(40):
        sim.step(torch.from_numpy(acts[st]).cuda(), 4, False)
        recs = sim.canonical_state()
        for k, s in enumerate(osims):
            h, j = ACTION_TO_INPUTS[acts[st, k]]
            for _ in range(4):
                s.tick(h, j)
            assert schema.state_hash(*s.state()) == schema.state_hash(*recs[k]), (k, st)


def test_reset_on_a_reused_simulator_vs_reference(golden_replays):
    """nppsim_reset mode 0 is Simulator.reset() the way the reference's callers use it -- on the Simulator that already holds
    the level (NPlayHeadless.reset, ActionReplayer).  tests/golden/reset_reuse.npz (unmodified reference,
    oracle/gen_golden_reset.py): load -> 150 ticks -> reset() -> 400 ticks -> reset() -> 400 ticks on every bc_replays level
    with two or more death balls; entity indices keep counting (entities.py:213-216) and the balls stop repelling each other
    after the first episode (entity_death_ball.py:153-166) -- 10 of the 19 levels, 41 and 86 among them, then move differently
    from a fresh Simulator.  Full canonical record, `index` included, after the load, every tick and each reset."""
    import torch
    from nclone_b200 import BatchedSimulator
    d = np.load(os.path.join(ROOT, "tests", "golden", "reset_reuse.npz"))
    levels = [int(v) for v in d["level"]]
    starts = list(np.cumsum(d["segments"])[:-1])
    assert int((d["differs_from_fresh"] >= 0).sum()) >= 8 and {41, 86} <= set(levels)
    N = len(levels)
    sim = BatchedSimulator(N)
    sim.load_maps([golden_replays.traj(i)["map_data"] for i in levels], np.arange(N))
    assert sim.kernel_variant() == "sim_full"
    hor_d = torch.from_numpy(np.ascontiguousarray(d["hor"].T)).cuda(); jump_d = torch.from_numpy(np.ascontiguousarray(d["jump"].T)).cuda()
    bad = [(levels[k], "load") for k, r in enumerate(sim.canonical_state()) if schema.state_hash(*r) != d["hash_init"][k]]
    alive = np.ones(N, dtype=bool)
    for t in range(hor_d.shape[0]):
        if t in starts:
            sim.reset()
            for k, r in enumerate(sim.canonical_state()):
                if alive[k] and schema.state_hash(*r) != d["hash_reset"][k, starts.index(t)]:
                    bad.append((levels[k], f"reset before tick {t}")); alive[k] = False
        sim.tick(hor_d[t], jump_d[t], n_ticks=1)
        ids = np.nonzero(alive)[0]
        for k, r in zip(ids, sim.canonical_state(ids)):
            if schema.state_hash(*r) != d["hash_tick"][k, t]:
                bad.append((levels[k], t)); alive[k] = False       # first differing frame only
    assert not bad, bad
    # a batch that is reloaded instead (fresh Simulators) does NOT follow these goldens where balls meet: the distinction is real
    fresh = BatchedSimulator(N); fresh.load_maps([golden_replays.traj(i)["map_data"] for i in levels], np.arange(N))
    for t in range(hor_d.shape[0]):
        if t in starts:
            fresh.reset(mode="load")
        fresh.tick(hor_d[t], jump_d[t], n_ticks=1)
    differ = (fresh.dyn_f.view(torch.int64) != sim.dyn_f.view(torch.int64)).any(0).cpu().numpy()
    assert differ.sum() >= 5, differ


def test_step_host_matches_device_path(golden_replays, golden_generated):
    """nppsim_step_host: host actions in, one packed block out (position / velocity / state and, with_obs, the 40-float
    ninja state + action mask), everything in the CALLER's env order although the engine keeps the envs of a level in
    consecutive state columns (nppsim_state_t.ext_of: 3 levels interleaved over 96 envs here)."""
    import torch
    from nclone_b200 import BatchedSimulator
    maps = [golden_replays.traj(3)["map_data"], golden_generated.traj(0)["map_data"], golden_generated.traj(40)["map_data"]]
    N = 96
    e2m = np.arange(N) % 3
    a = np.random.default_rng(1).integers(0, 6, size=(20, N)).astype(np.uint8)
    s1 = BatchedSimulator(N); s1.load_maps(maps, e2m)
    s2 = BatchedSimulator(N); s2.load_maps(maps, e2m)
    s3 = BatchedSimulator(N, placement="caller"); s3.load_maps(maps, e2m)
    assert s1.slot_env is not None and s3.slot_env is None and np.array_equal(s1.env_map_host, e2m)
    block, v = s2.host_block(with_obs=True)
    block3, v3 = s3.host_block(with_obs=False)
    # pinned blocks are written by the step kernel itself (zero-copy stores as each env finishes); a pageable block takes the
    # staged path (pack kernel + one copy): both must deliver the same bytes
    s4 = BatchedSimulator(N); s4.load_maps(maps, e2m)
    block4 = torch.zeros(block.numel(), dtype=torch.uint8)
    assert block.is_pinned() and not block4.is_pinned()
    for k in range(20):
        s1.step(torch.from_numpy(a[k]).cuda(), 4, True)
        s2.step_host(a[k], block, 4, True, with_obs=True)
        s3.step_host(a[k], block3, 4, True)
        s4.step_host(a[k], block4, 4, True, with_obs=True)
        assert np.array_equal(block4.numpy(), block.numpy()), k
    assert np.array_equal(s1.nin_f[:4].cpu().numpy().view(np.int64), v["pos_vel"].view(np.int64))
    assert np.array_equal(s1.nin_i[0].cpu().numpy(), v["state"])
    assert np.array_equal(v3["pos_vel"].view(np.int64), v["pos_vel"].view(np.int64)) and np.array_equal(v3["state"], v["state"])
    assert np.array_equal(s1.get_ninja_state().cpu().numpy().view(np.int32), v["ninja_state"].view(np.int32))
    assert np.array_equal(s1.get_action_mask().cpu().numpy().astype(np.uint8), v["action_mask"])
    assert len(np.unique(v["pos_vel"][0])) > 3                     # the three levels really behave differently
    with pytest.raises(ValueError):
        s2.step_host(a[0].astype(np.int64), block)                 # dtype is checked, not reinterpreted
    # nppsim_step_host_reset: the same call queues the episode turnover behind the results -- the block shows the step's
    # (terminal) states, the device state is the one after reset_terminal
    s5 = BatchedSimulator(N); s5.load_maps(maps, e2m)
    s6 = BatchedSimulator(N, placement="caller"); s6.load_maps(maps, e2m)
    block5, v5 = s5.host_block(with_obs=True)
    block6, v6 = s6.host_block(with_obs=True)
    terminal = 0
    a_pinned = torch.from_numpy(a).pin_memory().numpy()           # pinned actions are read in place by the step kernel, pageable ones copied first
    for k in range(20):
        s5.step_host(a[k], block5, 4, True, with_obs=True)
        s6.step_host(a_pinned[k], block6, 4, True, with_obs=True, reset_max_frames=24)
        assert np.array_equal(block5.numpy(), block6.numpy()), k
        terminal += int(np.isin(v6["state"], (6, 7, 8)).sum())
        s5.reset_terminal(24)
        for x, y in ((s5.nin_f.view(torch.int64), s6.nin_f.view(torch.int64)), (s5.nin_i, s6.nin_i), (s5.ent_flag, s6.ent_flag)):
            assert torch.equal(x, y), k
    assert int(s6.stats[0].item()) == int(s5.stats[0].item()) >= N     # every env was truncated at 24 frames at least once


def _oracle_rollout(maps, env_to_map, actions, frame_skip, max_frames, check_every):
    """Oracle side of test_large_random_rollout: list of (step, nin_f[N,31], nin_i[N,27]) checkpoints.
    Same policy as bench.py: early exit on win/death inside the frame-skip, Simulator.reset() on
    terminal / max_frames after the step."""
    from oracle import oracle as orc
    N = len(env_to_map)
    sims = []
    for k in range(N):
        s = orc.OracleSim(); s.load(maps[env_to_map[k]]); sims.append(s)
    L = orc.lib()
    out = []
    nf = np.zeros((N, 31)); ni = np.zeros((N, 27), dtype=np.int64)
    import ctypes as C
    for st in range(actions.shape[0]):
        for k, s in enumerate(sims):
            h, j = ((0, 0), (-1, 0), (1, 0), (0, 1), (-1, 1), (1, 1))[actions[st, k]]
            for _ in range(frame_skip):
                L.osim_tick(s.h, h, j)
                L.osim_get_state(s.h, nf[k].ctypes.data_as(C.POINTER(C.c_double)), ni[k].ctypes.data_as(C.POINTER(C.c_int64)), None, None)
                if ni[k, 0] in (6, 7, 8):
                    break
        if (st + 1) % check_every == 0:
            out.append((st, nf.copy(), ni.copy()))
        for k, s in enumerate(sims):
            if ni[k, 0] in (6, 7, 8) or ni[k, 26] >= max_frames:
                s.reset()                             # Simulator.reset() on the same Simulator, as nppsim_reset mode 0
    return out


def test_large_random_rollout_vs_oracle(golden_replays):
    """Guard-band / pooling safety net: 2,048 envs over the benchmark's generated levels and 512 over
    entity-rich replay levels, 240 agent steps (frame-skip 4, auto-reset): the ninja record of every env
    is compared bit-for-bit with the oracle every 20 steps, and full entity state at the end."""
    import torch
    from nclone_b200 import BatchedSimulator
    import bench
    arr, lens = bench.load_maps("simplest_with_mines")
    maps = [arr[k, :lens[k]] for k in range(0, 2048)]
    arr2, lens2 = bench.load_maps("locked_doors")
    maps += [arr2[k, :lens2[k]] for k in range(0, 256)]
    maps += [golden_replays.traj(i)["map_data"] for i in range(0, 130)]
    M = len(maps)
    N = 2048 + 256 + 130 * 2
    env_to_map = np.concatenate([np.arange(2048), 2048 + np.arange(256), 2304 + np.arange(260) % 130])
    rng = np.random.default_rng(11)
    steps, every = 240, 20
    actions = rng.integers(0, 6, size=(steps, N)).astype(np.uint8)
    # long holds make the ninja run / wall-slide / fall far: repeat each action for a few agent steps
    hold = rng.integers(1, 6, size=(steps, N))
    for st in range(1, steps):
        keep = (st % hold[st]) != 0
        actions[st, keep] = actions[st - 1, keep]
    sim = BatchedSimulator(N, placement="level")              # state columns permuted against the caller's env order (ext_of)
    sim.load_maps(maps, env_to_map)
    assert sim.slot_env is not None
    ref = _oracle_rollout(maps, env_to_map, actions, 4, 2000, every)
    ci = 0
    for st in range(steps):
        sim.step(torch.from_numpy(actions[st]).cuda(), 4, True)
        if (st + 1) % every == 0:
            _, nf, ni = ref[ci]; ci += 1
            gf = sim.nin_f.cpu().numpy().T; gi = sim.nin_i.cpu().numpy().T.astype(np.int64)
            badf = np.nonzero((gf.view(np.int64) != nf.view(np.int64)).any(1))[0]
            badi = np.nonzero((gi != ni).any(1))[0]
            assert badf.size == 0 and badi.size == 0, (st, badf[:5], badi[:5], env_to_map[badf[:5]])
        sim.reset_terminal(2000)


def test_soak_slice_random_rollout_vs_oracle():
    """A bounded slice of tools/soak_parity.py inside the suite (about 30 s): 12,288 envs on random levels of both bench level
    sets (config 3's mines, config 4's locked doors), 160 agent steps of held random actions (7.4 M ticks): every env's
    ninja record compared bit-for-bit with the oracle every 20 steps.  The oracle side is the threaded C rollout
    (oracle/baseline_runner.c: osim_rollout_records), itself pinned to the per-tick loop in test_oracle_golden.py.
    The seed is fixed so that a failure reproduces; tools/soak_parity.py takes any seed and size."""
    import torch
    from nclone_b200 import BatchedSimulator
    from oracle import oracle as orc
    import bench
    arr, lens = bench.load_maps("simplest_with_mines")
    arr2, lens2 = bench.load_maps("locked_doors")
    stride = max(arr.shape[1], arr2.shape[1])
    maps = np.zeros((len(arr) + len(arr2), stride)); maps[:len(arr), :arr.shape[1]] = arr; maps[len(arr):, :arr2.shape[1]] = arr2
    mlens = np.concatenate([lens, lens2]).astype(np.int32)
    N, steps, every = 12288, 160, 20
    rng = np.random.default_rng(2026)
    env_to_map = rng.integers(0, len(maps), size=N)
    actions = rng.integers(0, 6, size=(steps, N)).astype(np.uint8)
    hold = rng.integers(1, 8, size=(steps, N))
    for st in range(1, steps):
        keep = (st % hold[st]) != 0
        actions[st, keep] = actions[st - 1, keep]
    used = np.unique(env_to_map)                                   # compile only the levels in use
    remap = np.zeros(len(maps), dtype=np.int64); remap[used] = np.arange(len(used))
    ticks, of, oi = orc.rollout_records(maps, mlens, env_to_map, actions, 4, 2000, every, os.cpu_count() or 1)
    sim = BatchedSimulator(N)
    sim.load_maps([maps[k, :mlens[k]] for k in used], remap[env_to_map])
    for st in range(steps):
        sim.step(torch.from_numpy(actions[st]).cuda(), 4, True)
        if (st + 1) % every == 0:
            c = st // every
            gf = sim.nin_f.cpu().numpy().T; gi = sim.nin_i.cpu().numpy().T.astype(np.int64)
            badf = np.nonzero((gf.view(np.int64) != of[c].view(np.int64)).any(1))[0]
            badi = np.nonzero((gi != oi[c]).any(1))[0]
            assert badf.size == 0 and badi.size == 0, (st, badf[:5], badi[:5], env_to_map[badf[:5]])
        sim.reset_terminal(2000)
    assert sim.total_ticks() == ticks and ticks > 6_000_000
    assert sim.handoff_status() == 0


def test_wedged_ninja_cooperative_loop_vs_oracle():
    """The kernel's critical path: a ninja wedged into a corner runs all 32 depenetration iterations of every substep
    (ninja.py:303-379) through the warp-cooperative loop with winner prediction.  bench_data/wedges.npz holds three
    levels + action prefixes that end in such a state (one against an arc, two against line segments); each is
    continued with every Discrete(6) action held for 30 agent steps.  Full ninja record against the oracle after
    every agent step, and the loop must actually have been taken."""
    import torch
    from nclone_b200 import BatchedSimulator
    d = np.load(os.path.join(ROOT, "bench_data", "wedges.npz"))
    K = len(d["lens"])
    maps = [d["maps"][k, :d["lens"][k]] for k in range(K)]
    pre = [d["acts"][d["act_off"][k]:d["act_off"][k + 1]] for k in range(K)]
    env_to_map = np.repeat(np.arange(K), 6)
    N = len(env_to_map)
    steps = max(len(p) for p in pre) + 30
    actions = np.zeros((steps, N), dtype=np.uint8)
    for e in range(N):
        p = pre[env_to_map[e]]
        actions[:len(p), e] = p
        actions[len(p):, e] = e % 6
    sim = BatchedSimulator(N)
    sim.load_maps(maps, env_to_map)
    ref = _oracle_rollout(maps, env_to_map, actions, 4, 2000, 1)
    most = 0
    for st in range(steps):
        sim.step(torch.from_numpy(actions[st]).cuda(), 4, True)
        most = max(most, int(sim.warp_cost[0].item()) - 1)
        _, nf, ni = ref[st]
        gf = sim.nin_f.cpu().numpy().T; gi = sim.nin_i.cpu().numpy().T.astype(np.int64)
        badf = np.nonzero((gf.view(np.int64) != nf.view(np.int64)).any(1))[0]
        badi = np.nonzero((gi != ni).any(1))[0]
        assert badf.size == 0 and badi.size == 0, (st, badf[:5], badi[:5])
        sim.reset_terminal(2000)
    assert most >= 1000, most      # several envs of the warp wedged in the same agent step (>= 354 iterations each)


@pytest.mark.parametrize("solo_sm", ["1", "2"])
def test_wedged_ninja_handoff_to_solo_warps_vs_oracle(solo_sm, monkeypatch):
    """In-launch handoff (nppsim_state_t.handoff, csrc/sim_variant.inc): a wedged env leaves its home lane at the end of its
    first wedged tick and finishes the agent step on a solo warp; envs wedged when the previous launch ended start there.
    Same scenario as above with the queue switched on (it is on by default from 1,024 envs): full ninja record and
    frames_executed against the oracle after every agent step, handoffs must have happened, no solo warp may have timed out.
    Second half: per-tick input rows, per-env action streams and one held hor / jump value per env (the three other input
    forms) through the handoff -- an env that changes hands at tick >= 1 must keep the inputs it was given.
    solo_sm = "2": the same with the solo blocks dealt by SM (whole SMs reserved for solo warps, regular blocks taking their envs
    from a counter) -- what launches of a full wave and more do by themselves -- forced onto this small grid."""
    import torch
    from nclone_b200 import BatchedSimulator, ACTION_TO_INPUTS
    monkeypatch.setenv("NPPSIM_SOLO_SM", solo_sm)
    d = np.load(os.path.join(ROOT, "bench_data", "wedges.npz"))
    K = len(d["lens"])
    maps = [d["maps"][k, :d["lens"][k]] for k in range(K)]
    pre = [d["acts"][d["act_off"][k]:d["act_off"][k + 1]] for k in range(K)]
    env_to_map = np.repeat(np.arange(K), 12)
    N = len(env_to_map)
    steps = max(len(p) for p in pre) + 30
    actions = np.zeros((steps, N), dtype=np.uint8)
    for e in range(N):
        p = pre[env_to_map[e]]
        actions[:len(p), e] = p
        actions[len(p):, e] = e % 6
    ref = _oracle_rollout(maps, env_to_map, actions, 4, 2000, 1)
    for mode in ("actions", "rows", "stream", "held"):
        sim = BatchedSimulator(N, handoff=True)
        sim.load_maps(maps, env_to_map)
        hj = np.array(ACTION_TO_INPUTS, dtype=np.int8)
        sp = sim.make_stream([actions[:, e] for e in range(N)], "actions", 4) if mode == "stream" else None
        plain = None
        seen_queue = 0
        if mode == "actions":                          # frames_executed against the same batch without a queue
            plain = BatchedSimulator(N, handoff=False); plain.load_maps(maps, env_to_map)
        for st in range(steps):
            if mode == "actions":
                fe = sim.step(torch.from_numpy(actions[st]).cuda(), 4, True)
                ticks = plain.step(torch.from_numpy(actions[st]).cuda(), 4, True).cpu().numpy(); plain.reset_terminal(2000)
            elif mode == "rows":
                hor = np.repeat(hj[actions[st], 0][None, :], 4, 0); jump = np.repeat(hj[actions[st], 1][None, :], 4, 0)
                fe = sim.tick(hor, jump, n_ticks=4, stop_on_terminal=True)
            elif mode == "held":
                fe = sim.tick(hj[actions[st], 0].copy(), hj[actions[st], 1].copy(), n_ticks=4, stop_on_terminal=True)
            else:
                fe = sim.step_stream(sp, 4, start_tick=4 * st, stop_on_terminal=True)
            _, nf, ni = ref[st]
            gf = sim.nin_f.cpu().numpy().T; gi = sim.nin_i.cpu().numpy().T.astype(np.int64)
            badf = np.nonzero((gf.view(np.int64) != nf.view(np.int64)).any(1))[0]
            badi = np.nonzero((gi != ni).any(1))[0]
            assert badf.size == 0 and badi.size == 0, (mode, st, badf[:5], badi[:5])
            if mode == "actions":
                assert np.array_equal(fe.cpu().numpy(), ticks), (mode, st)
            seen_queue = max(seen_queue, sim.solo_queue_length())
            sim.reset_terminal(2000)
        assert sim.handoff_status() == 0
        assert sim.handoff_count() >= 18, sim.handoff_count()      # mid-launch handoffs: one per wedge episode of 36 envs
        assert seen_queue > 0                                       # and envs started launches on solo warps
        assert (sim.solo_sms() >= 1) == (solo_sm == "2")


@pytest.mark.parametrize("solo_sm", ["1", "2"])
def test_handoff_with_tick_limits_and_long_launches(solo_sm, monkeypatch):
    """The handoff under per-env tick limits (nppsim_step's tick_limit: 0 .. n_ticks ticks per env and launch, so that envs
    are handed over with any number of ticks left, including none) and under launches of 8 ticks (two agent steps fused):
    state and frames_executed must equal those of a batch stepped without a queue, launch by launch."""
    import torch
    from nclone_b200 import BatchedSimulator, ACTION_TO_INPUTS
    monkeypatch.setenv("NPPSIM_SOLO_SM", solo_sm)
    d = np.load(os.path.join(ROOT, "bench_data", "wedges.npz"))
    K = len(d["lens"])
    maps = [d["maps"][k, :d["lens"][k]] for k in range(K)]
    pre = [d["acts"][d["act_off"][k]:d["act_off"][k + 1]] for k in range(K)]
    env_to_map = np.repeat(np.arange(K), 16)
    N = len(env_to_map)
    rng = np.random.default_rng(9)
    hj = np.array(ACTION_TO_INPUTS, dtype=np.int8)
    for n_ticks in (4, 8):
        launches = (max(len(p) for p in pre) * 4) // n_ticks + 40
        sim = BatchedSimulator(N, handoff=True); sim.load_maps(maps, env_to_map)
        plain = BatchedSimulator(N, handoff=False); plain.load_maps(maps, env_to_map)
        budget = np.zeros(N, dtype=np.int64)             # ticks of the prefix already played, per env
        for st in range(launches):
            lim = rng.integers(0, n_ticks + 1, size=N).astype(np.int32)
            hor = np.zeros((n_ticks, N), dtype=np.int8); jump = np.zeros((n_ticks, N), dtype=np.int8)
            for e in range(N):
                p = pre[env_to_map[e]]
                for t in range(n_ticks):                 # the prefix action of the tick this env would be at, then its last one held
                    a = p[min((budget[e] + t) // 4, len(p) - 1)]
                    hor[t, e], jump[t, e] = hj[a]
            fe = sim.tick(hor, jump, n_ticks=n_ticks, tick_limit=lim, stop_on_terminal=True)
            fp = plain.tick(hor, jump, n_ticks=n_ticks, tick_limit=lim, stop_on_terminal=True)
            torch.cuda.synchronize()
            assert torch.equal(fe, fp), (n_ticks, st)
            budget += fe.cpu().numpy()
            for x, y in ((sim.nin_f.view(torch.int64), plain.nin_f.view(torch.int64)), (sim.nin_i, plain.nin_i), (sim.ent_flag, plain.ent_flag)):
                assert torch.equal(x, y), (n_ticks, st)
        assert sim.handoff_status() == 0
        assert sim.handoff_count() >= 8, (n_ticks, sim.handoff_count())


def test_automatic_handoff_policy_drops_a_saturated_queue():
    """BatchedSimulator's default policy (handoff=None): the queue is on from 1,024 envs and is dropped once most launches
    overflow it -- wedged ninjas are then the bulk of the work, not its tail (sched_kernel's counters, words 12 / 13 of the
    queue; BatchedSimulator._adapt_handoff).  1,200 envs on the three wedge levels, every replica driven into its corner and
    held there: the policy must switch kernels mid-run, and the state must stay bit-identical to a batch stepped without a
    queue from the start (which the tests above pin to the oracle)."""
    import torch
    from nclone_b200 import BatchedSimulator
    d = np.load(os.path.join(ROOT, "bench_data", "wedges.npz"))
    K = len(d["lens"])
    maps = [d["maps"][k, :d["lens"][k]] for k in range(K)]
    pre = [d["acts"][d["act_off"][k]:d["act_off"][k + 1]] for k in range(K)]
    env_to_map = np.repeat(np.arange(K), 400)
    N = len(env_to_map)
    steps = max(len(p) for p in pre) + 72
    actions = np.zeros((steps, N), dtype=np.uint8)
    for e in range(N):
        p = pre[env_to_map[e]]
        actions[:len(p), e] = p
        actions[len(p):, e] = p[-1]                      # keep pushing into the corner
    auto = BatchedSimulator(N); auto.load_maps(maps, env_to_map)
    plain = BatchedSimulator(N, handoff=False); plain.load_maps(maps, env_to_map)
    assert auto.handoff is not None and auto.kernel_variant() == "sim_basic_solo"
    for st in range(steps):
        a = torch.from_numpy(actions[st]).cuda()
        fa = auto.step(a, 4, True); fp = plain.step(a, 4, True)
        if st % 8 == 7 or st == steps - 1:
            torch.cuda.synchronize()
            assert torch.equal(fa, fp), st
            for x, y in ((auto.nin_f.view(torch.int64), plain.nin_f.view(torch.int64)), (auto.nin_i, plain.nin_i), (auto.ent_flag, plain.ent_flag)):
                assert torch.equal(x, y), st
        auto.reset_terminal(2000); plain.reset_terminal(2000)
    clipped, launches = auto.handoff_saturation()
    assert launches >= 32 and clipped >= 16, (clipped, launches)
    assert auto.handoff is None and auto.kernel_variant() == "sim_basic", auto.handoff_saturation()
    assert auto.handoff_status() == 0


def test_basic_variant_vs_oracle_and_full_kernel():
    """Level sets without moving / pushing entities run on the specialised `sim_basic` kernel
    (nppsim_state_t.kind_mask).  (1) 1,280 envs of the benchmark levels against the oracle;
    (2) 16,384 envs x 150 agent steps: basic and full kernels must leave identical state everywhere."""
    import torch
    from nclone_b200 import fields as FD
    from nclone_b200 import BatchedSimulator
    from nclone_b200._lib import lib
    import bench
    BASIC = (1 << 1) | (1 << 2) | (1 << 3) | (1 << 4) | (1 << 6) | (1 << 7) | 1
    arr, lens = bench.load_maps("simplest_with_mines")
    arr2, lens2 = bench.load_maps("locked_doors")
    maps = [arr[k, :lens[k]] for k in range(2048, 3072)] + [arr2[k, :lens2[k]] for k in range(256, 512)]
    N = len(maps)
    rng = np.random.default_rng(5)
    steps, every = 160, 20
    actions = rng.integers(0, 6, size=(steps, N)).astype(np.uint8)
    hold = rng.integers(1, 6, size=(steps, N))
    for st in range(1, steps):
        keep = (st % hold[st]) != 0
        actions[st, keep] = actions[st - 1, keep]
    sim = BatchedSimulator(N)
    sim.load_maps(maps, np.arange(N))
    assert sim.kind_mask != 0 and (sim.kind_mask & ~BASIC) == 0
    ref = _oracle_rollout(maps, np.arange(N), actions, 4, 2000, every)
    ci = 0
    for st in range(steps):
        sim.step(torch.from_numpy(actions[st]).cuda(), 4, True)
        if (st + 1) % every == 0:
            _, nf, ni = ref[ci]; ci += 1
            gf = sim.nin_f.cpu().numpy().T; gi = sim.nin_i.cpu().numpy().T.astype(np.int64)
            badf = np.nonzero((gf.view(np.int64) != nf.view(np.int64)).any(1))[0]
            badi = np.nonzero((gi != ni).any(1))[0]
            assert badf.size == 0 and badi.size == 0, (st, badf[:5], badi[:5])
        sim.reset_terminal(2000)
    # (2) kernel against kernel
    N2 = 16384
    allmaps = [arr[k, :lens[k]] for k in range(len(arr))] + [arr2[k, :lens2[k]] for k in range(len(arr2))]
    e2m = np.arange(N2) % len(allmaps)
    gen = torch.Generator(device="cuda"); gen.manual_seed(7)
    acts = torch.randint(0, 6, (150, N2), generator=gen, device="cuda", dtype=torch.uint8)
    states = []
    for force in (False, True):
        s2 = BatchedSimulator(N2)
        s2.load_maps(allmaps, e2m)
        s2.force_full_kernel(force)            # kind_mask = 0 ("census unknown") selects the full kernel
        for st in range(150):
            s2.step(acts[st], 4, True)
            s2.reset_terminal(2000)
        envi = s2.env_i.clone()
        envi[FD.ENV_WEDGE] = 0                 # handoff marks are scheduling metadata of the basic_solo kernel (the full kernel has no solo warps)
        states.append([t.clone() for t in (s2.nin_f, s2.nin_i, envi, s2.ent_flag, s2.dyn_f, s2.dyn_i)])
    for a, b in zip(*states):
        assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))


def test_observation_primitives_vs_reference():
    """SURVEY 8(f) f1 on the device: action mask bit-exact, the 40-float ninja state within one float32 ulp
    of the reference's float64 values (the kernel stores float32 like the env wrapper does; `x ** 0.5` is
    sqrt and atan2 is CUDA's, both within 2 double ulp of glibc's).  Tolerance: |d| <= 1.2e-7 * max(1, |x|)."""
    import os
    import torch
    from conftest import Golden, GOLDEN
    from nclone_b200 import BatchedSimulator
    d = dict(np.load(os.path.join(GOLDEN, "obs.npz")))          # NpzFile decompresses on every access
    G = {"bc_replays": Golden(os.path.join(GOLDEN, "bc_replays.npz")), "generated": Golden(os.path.join(GOLDEN, "generated.npz"))}
    trajs = [G[str(s)].traj(int(i)) for s, i in zip(d["src"], d["idx"])]
    N = len(trajs)
    lens = np.array([len(t["hor"]) for t in trajs]); T = int(lens.max())
    hor = np.zeros((T, N), dtype=np.int8); jump = np.zeros((T, N), dtype=np.int8)
    for k, t in enumerate(trajs):
        hor[:lens[k], k] = t["hor"]; jump[:lens[k], k] = t["jump"]
    sim = BatchedSimulator(N)
    sim.load_maps([t["map_data"] for t in trajs], np.arange(N))
    hor_d = torch.from_numpy(hor).cuda(); jump_d = torch.from_numpy(jump).cuda()
    off = d["frame_off"]
    worst = 0.0
    for f in range(T + 1):
        if f > 0:
            sim.tick(hor_d[f - 1], jump_d[f - 1], n_ticks=1, tick_limit=(lens > f - 1).astype(np.int32))
        pd = torch.full((N,), f % 4, dtype=torch.uint8, device="cuda")
        mask = sim.get_action_mask(pd).cpu().numpy().astype(np.uint8)
        st = sim.get_ninja_state().cpu().numpy().astype(np.float64)
        for k in np.nonzero(lens >= f)[0]:
            want = d["ninja_state"][off[k] + f]
            assert np.array_equal(mask[k], d["mask"][off[k] + f]), (k, f)
            err = np.abs(st[k] - want) / np.maximum(1.0, np.abs(want))
            worst = max(worst, float(err.max()))
            assert err.max() <= 1.2e-7, (k, f, int(err.argmax()), st[k][err.argmax()], want[err.argmax()])
    assert worst <= 1.2e-7


def _replay_rows_fixture():
    import os
    from conftest import Golden, GOLDEN
    return dict(np.load(os.path.join(GOLDEN, "replay_rows.npz"))), Golden(os.path.join(GOLDEN, "bc_replays.npz"))


def test_replay_dataset_rows_vs_reference():
    """execute_replays (per-env input-byte streams + nppsim_replay_rows, everything in HBM) against
    the reference's own ReplayExecutor loop (tests/golden/replay_rows.npz).  Positions, flags, actions, frames and
    time_remaining bit-exact; the 40 ninja-state floats within one float32 ulp (tolerance 1.2e-7 * max(1, |x|), as
    in test_observation_primitives_vs_reference: sqrt / atan2 differ from glibc's by <= 2 double ulp)."""
    from nclone_b200 import execute_replays
    d, G = _replay_rows_fixture()
    inp_all, inp_off = G.d["input_bytes"], G.d["inp_off"]
    replays = [(G.traj(int(i))["map_data"], inp_all[inp_off[i]:inp_off[i + 1]]) for i in d["idx"]]
    for fs in (1, 4):
        out = execute_replays(replays, frame_skip=fs)
        assert np.array_equal(out["row_off"].cpu().numpy(), d[f"row_off_fs{fs}"])
        assert np.array_equal(out["pos"].cpu().numpy().view(np.int64), d[f"pos_fs{fs}"].view(np.int64))
        fl = d[f"flags_fs{fs}"]
        assert np.array_equal(out["switch_activated"].cpu().numpy(), fl[:, 0].astype(bool))
        assert np.array_equal(out["player_won"].cpu().numpy(), fl[:, 1].astype(bool))
        assert np.array_equal(out["player_dead"].cpu().numpy(), fl[:, 2].astype(bool))
        assert np.array_equal(out["action"].cpu().numpy(), fl[:, 3])
        assert np.array_equal(out["frame"].cpu().numpy(), d[f"frame_fs{fs}"])
        gs = out["game_state"].cpu().numpy(); want = d[f"game_state_fs{fs}"]
        assert np.array_equal(gs[:, 40].view(np.int32), want[:, 40].view(np.int32))
        err = np.abs(gs.astype(np.float64) - want) / np.maximum(1.0, np.abs(want))
        assert err.max() <= 1.2e-7, (fs, np.unravel_index(err.argmax(), err.shape))
        assert int(out["player_won"].sum()) > 0


def test_checkpoint_replay_vs_reference():
    """replay_to_checkpoint (reset + ragged per-env action streams, one launch) against the
    reference's ActionReplayer(frame_skip=4).replay_to_checkpoint.  Every column bit-exact, entity `index` included: it
    keeps counting across the reset() on the reference's re-used Simulator (entities.py:213-216), and so does the batch's."""
    import torch
    from nclone_b200 import BatchedSimulator
    d, G = _replay_rows_fixture()
    trajs = [G.traj(int(i)) for i in d["idx"]]
    N = len(trajs)
    sim = BatchedSimulator(N)
    sim.load_maps([t["map_data"] for t in trajs], np.arange(N))
    rng = np.random.default_rng(3)
    for _ in range(7):                                   # the envs have been used before the restore
        sim.step(torch.from_numpy(rng.integers(0, 6, N).astype(np.uint8)).cuda(), 4, False)
    seqs = [d["ck_actions"][d["ck_off"][k]:d["ck_off"][k + 1]] for k in range(N)]
    x, y = sim.replay_to_checkpoint(seqs, frame_skip=4)
    assert np.array_equal(np.stack([x.cpu().numpy(), y.cpu().numpy()], 1).view(np.int64), d["ck_pos"].view(np.int64))
    assert np.array_equal(sim.frames_executed.cpu().numpy(), np.array([4 * len(q) for q in seqs]))
    for k, (nf, ni, ef, ei) in enumerate(sim.canonical_state()):
        e0, e1 = d["ck_ent_off"][k], d["ck_ent_off"][k + 1]
        assert np.array_equal(nf.view(np.int64), d["ck_nin_f"][k].view(np.int64)) and np.array_equal(ni, d["ck_nin_i"][k]), k
        assert np.array_equal(ef.view(np.int64), d["ck_ent_f"][e0:e1].view(np.int64)), k
        assert np.array_equal(ei, d["ck_ent_i"][e0:e1]), k
    # masked restore: only env 0 is touched
    before = sim.canonical_state()
    mask = np.zeros(N, dtype=np.uint8); mask[0] = 1
    sim.replay_to_checkpoint(seqs, frame_skip=4, mask=mask)
    after = sim.canonical_state()
    for k in range(1, N):
        assert schema.state_hash(*before[k]) == schema.state_hash(*after[k]), k
    # env 0 went through one more reset(): the same state again, its entities' indices one level's worth further on
    assert all(np.array_equal(a, b) for a, b in zip(before[0][:3], after[0][:3]))
    assert np.array_equal(np.delete(before[0][3], 1, axis=1), np.delete(after[0][3], 1, axis=1)) and (after[0][3][:, 1] > before[0][3][:, 1]).all()


def test_env_view_mirrors_single_env_interface(golden_replays):
    """EnvView = env i of a batch behind NPlayHeadless's scalar interface.  Env 3 of a batch is driven
    through view.tick() with a replay's inputs and read through the facade getters; the other envs must not move."""
    from nclone_b200 import BatchedSimulator, EnvView
    idx = [0, 7, 14, 40, 77]
    trajs = [golden_replays.traj(i) for i in idx]
    sim = BatchedSimulator(len(idx))
    sim.load_maps([t["map_data"] for t in trajs])
    k = 3
    t = trajs[k]
    others = [schema.state_hash(*r) for r in sim.canonical_state()]
    v = EnvView(sim, k, map_data=t["map_data"])
    assert v.ninja_position() == (t["pv"][0, 0], t["pv"][0, 1]) and v.sim.frame == 0
    assert not v.exit_switch_activated()
    o = oracle.OracleSim(); o.load(t["map_data"])
    for f in range(len(t["hor"])):
        v.tick(int(t["hor"][f]), int(t["jump"][f])); o.tick(int(t["hor"][f]), int(t["jump"][f]))
        assert v.ninja_position() == (t["pv"][f + 1, 0], t["pv"][f + 1, 1]), f
        assert v.ninja_velocity() == (t["pv"][f + 1, 2], t["pv"][f + 1, 3]), f
        assert v.sim.ninja.state == t["state"][f + 1] and v.sim.frame == f + 1
        if f % 50 == 0:
            assert v.get_action_mask() == [bool(b) for b in o.action_mask(0)]
            assert np.abs(v.get_ninja_state() - o.ninja_state()).max() <= 1.2e-7
    assert v.ninja_has_won() and v.exit_switch_activated() and not v.ninja_has_died()
    nf, ni, ef, ei = o.state()
    flat = [e for lst in v.sim.entity_dic.values() for e in lst]
    assert len(flat) == len(ei)
    for e, rf, ri in zip(flat, ef, ei):
        assert (e.type, e.active, e.xpos, e.ypos, e.cell) == (ri[0], bool(ri[2]), rf[0], rf[1], (ri[3], ri[4]))
    sw = [r for r in range(len(ei)) if ei[r, 0] == 4]
    assert v.exit_switch_position() == (ef[sw[-1], 0], ef[sw[-1], 1])
    assert len(v.sim.tile_dic) == 1100 and v.sim.tile_dic[(0, 0)] == 1
    for j, r in enumerate(sim.canonical_state()):
        if j != k:
            assert schema.state_hash(*r) == others[j]
    v.reset()
    assert v.ninja_position() == (t["pv"][0, 0], t["pv"][0, 1]) and v.sim.frame == 0


def test_groups_views_step_independently(golden_replays, golden_generated):
    """BatchedSimulator.groups(G): views over one allocation (nppsim_state_t.stride) stepped on their own streams, in a
    different interleaving per group, must leave exactly the state the whole batch reaches when stepped as one."""
    import torch
    from nclone_b200 import BatchedSimulator
    maps = [golden_generated.traj(i)["map_data"] for i in range(0, len(golden_generated), 3)]
    maps += [golden_replays.traj(i)["map_data"] for i in range(0, 130, 9)]
    N = 5 * 32 + 7
    env_to_map = np.arange(N) % len(maps)
    rng = np.random.default_rng(21)
    steps = 40
    acts = rng.integers(0, 6, size=(steps, N)).astype(np.uint8)
    whole = BatchedSimulator(N); whole.load_maps(maps, env_to_map)
    for s in range(steps):
        whole.step(torch.from_numpy(acts[s]).cuda(), 4, True); whole.reset_terminal(60)
    want = [schema.state_hash(*r) for r in whole.canonical_state()]
    sim = BatchedSimulator(N, placement="level"); sim.load_maps(maps, env_to_map)
    groups = sim.groups(3)
    assert [g.n_envs for g in groups] == [64, 64, 39]
    streams = [torch.cuda.Stream() for _ in groups]
    torch.cuda.synchronize()
    # a group is a range of state columns: the envs of the parent it holds (the engine keeps the envs of a level together)
    assert sim.slot_env is not None and sorted(np.concatenate([g.env_ids for g in groups]).tolist()) == list(range(N))
    blocks = [g.host_block() for g in groups]
    for k, (g, st) in enumerate(zip(groups, streams)):       # group k runs all its steps before group k+1 is even issued
        with torch.cuda.stream(st):
            for s in range(steps):
                a = acts[s, g.env_ids]
                if s % 2:
                    g.step_host(np.ascontiguousarray(a), blocks[k][0], 4, True)      # host-buffer entry point on a view
                else:
                    g.step(torch.from_numpy(np.ascontiguousarray(a)).cuda(), 4, True)
                g.reset_terminal(60)
    torch.cuda.synchronize()
    got = [schema.state_hash(*r) for r in sim.canonical_state()]
    assert got == want
    assert int(sim.stats[0]) == int(whole.stats[0]) > 0
    # getters of a view are slices of the parent's
    x, _ = groups[1].ninja_position()
    assert torch.equal(x, sim.ninja_position()[0][torch.from_numpy(groups[1].env_ids).cuda()])
    assert groups[2].get_action_mask().shape == (39, 6) and groups[2].get_ninja_state().shape == (39, 40)


def test_full_size_replica_consistency_and_oracle_sample():
    """BASELINE.json configs[2] at full size: 65,536 envs = 4,096 generated levels x 16 replicas.  Size-independent
    properties: (1) the 16 replicas of a level get the same action stream and sit in different warps / lanes / launch
    slots (longest-first scheduling reorders the groups every step), so all 16 must hold bit-identical state at every
    check -- any dependence of the pooled / cooperative paths on the neighbours in the warp would break it;
    (2) the envs whose warp did the most cooperative depenetration iterations (wedged ninjas: the cooperative loop with
    winner prediction) and a random sample are replayed on the oracle and compared bit-for-bit."""
    import torch
    from nclone_b200 import BatchedSimulator
    import bench
    arr, lens = bench.load_maps("simplest_with_mines")
    M = len(arr)
    maps = [arr[k, :lens[k]] for k in range(M)]
    R = 16
    N = M * R
    # replica r of level m lives at env (m * 7 + r * 4099) % N-ish: a permutation that spreads replicas over warps and lanes
    perm = (np.arange(N, dtype=np.int64) * 40503) % N          # 40503 is odd, N a power of two: a bijection
    env_to_map = np.empty(N, dtype=np.int64); env_to_map[perm] = np.arange(N) % M
    # the permuted layout at full size (the default here would be the caller's order: 16 replicas per level); the queue pinned
    # on: all 16 replicas of a level wedge together here, which the automatic policy would (rightly) answer by dropping it
    sim = BatchedSimulator(N, placement="level", handoff=True)
    sim.load_maps(maps, env_to_map)
    rng = np.random.default_rng(3)
    steps = 60
    acts_map = rng.integers(0, 6, size=(steps, M)).astype(np.uint8)
    hold = rng.integers(1, 6, size=(steps, M))
    for st in range(1, steps):
        keep = (st % hold[st]) != 0
        acts_map[st, keep] = acts_map[st - 1, keep]
    e2m_d = torch.from_numpy(env_to_map).cuda()
    order = torch.argsort(e2m_d, stable=True).view(M, R)       # row m: the 16 envs on level m
    heavy = np.zeros(N // 32, dtype=np.int64)
    from nclone_b200 import fields as FD
    solo_envs = set()
    for st in range(steps):
        a = torch.from_numpy(acts_map[st]).cuda()[e2m_d]
        sim.step(a, 4, True)
        heavy = np.maximum(heavy, sim.warp_cost.cpu().numpy().astype(np.int64) - 1)
        solo_envs.update(torch.nonzero(sim.env_i[FD.ENV_WEDGE]).flatten().tolist())     # envs that start the next launch on a solo warp
        if (st + 1) % 10 == 0:
            for t in (sim.nin_f, sim.nin_i):
                v = t.view(torch.int64) if t.dtype == torch.float64 else t
                rows = v[:, order]                              # [fields, M, R]
                assert bool((rows == rows[:, :, :1]).all()), st
            fl = sim.ent_flag[:, order]
            assert bool((fl == fl[:, :, :1]).all()), st
        if st + 1 < steps:
            sim.reset_terminal(2000)
    # wedged ninjas occurred, and they went through the solo-warp handoff (on by default at this size), both ways
    assert sim.handoff_status() == 0
    assert len(solo_envs) >= 8 and sim.handoff_count() >= 8, (len(solo_envs), sim.handoff_count())
    # (2) oracle replay of the envs that ran on solo warps, of those in the heaviest warps and of a random sample
    top = np.argsort(-heavy)[:4]
    cols = (top[:, None] * 32 + np.arange(32)[None, :]).ravel()   # warp groups are ranges of state columns
    heavy_envs = sim.slot_env[cols] if sim.slot_env is not None else cols
    pick = np.unique(np.concatenate([np.array(sorted(solo_envs), dtype=np.int64)[:160], heavy_envs, rng.integers(0, N, size=64)]))
    sub_maps = env_to_map[pick]
    acts = acts_map[:, sub_maps]
    ref = _oracle_rollout(maps, sub_maps, acts, 4, 2000, steps)
    _, nf, ni = ref[-1]
    # the oracle's checkpoint is taken before the post-step reset, as the big batch was left above
    tp = torch.from_numpy(pick).cuda()
    gf = sim.nin_f[:, tp].cpu().numpy().T.copy(); gi = sim.nin_i[:, tp].cpu().numpy().T.astype(np.int64)
    assert np.array_equal(gf.view(np.int64), nf.view(np.int64)) and np.array_equal(gi, ni)
    # and the same envs as a small batch of their own (no handoff queue below 1,024 envs)
    sim2 = BatchedSimulator(len(pick)); sim2.load_maps(maps, sub_maps)
    assert sim2.handoff is None
    for st in range(steps):
        sim2.step(torch.from_numpy(np.ascontiguousarray(acts[st])).cuda(), 4, True)
        if st + 1 < steps:
            sim2.reset_terminal(2000)
    gf = sim2.nin_f.cpu().numpy().T; gi = sim2.nin_i.cpu().numpy().T.astype(np.int64)
    assert np.array_equal(gf.view(np.int64), nf.view(np.int64)) and np.array_equal(gi, ni)
    # and the small batch agrees with the same envs inside the big one (state before the last reset is gone there; compare
    # the envs that did not terminate in the last step)
    alive = ~np.isin(gi[:, 0], (6, 7, 8)) & (gi[:, 26] < 2000)
    big_f = sim.nin_f[:, torch.from_numpy(pick).cuda()].cpu().numpy().T
    assert np.array_equal(big_f[alive].view(np.int64), gf[alive].view(np.int64))
