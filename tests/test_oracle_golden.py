"""The C restatement (oracle/nsim_oracle.c) against the reference's own outputs.

The fixtures were produced by oracle/gen_golden.py from the unmodified reference;
every frame of every trajectory must match bit for bit (64-bit digest of the full
canonical record: ninja's 31 doubles + 27 ints and every entity row)."""
import numpy as np
import pytest

from oracle import oracle, schema


def _check(traj):
    rec = oracle.run_trajectory(traj["map_data"], traj["hor"], traj["jump"], moves=traj["moves"])
    bad = np.nonzero(rec["hash"] != traj["hash"])[0]
    assert bad.size == 0, f"{traj['name']}: first differing frame {bad[0]} of {len(traj['hor'])}"
    assert np.array_equal(rec["nin_f"][:, :4].view(np.int64), traj["pv"].view(np.int64))
    assert np.array_equal(rec["nin_i"][:, 0], traj["state"])
    assert np.array_equal(rec["nin_f"][-1].view(np.int64), traj["final_nin_f"].view(np.int64))
    assert np.array_equal(rec["nin_i"][-1], traj["final_nin_i"])
    assert np.array_equal(rec["ent_f"][-1].view(np.int64), traj["final_ent_f"].view(np.int64))
    assert np.array_equal(rec["ent_i"][-1], traj["final_ent_i"])
    return len(traj["hor"])


def test_known_answers(golden_replays):
    """scripts/validate_replays.py:106-140 requires a win on the last frame."""
    g = golden_replays
    assert len(g) == 130
    finals = g.d["final_nin_i"][:, 0]
    assert int((finals == 8).sum()) == 128
    no_exit = {"20251022_151333_train_simple_100001", "20251023_095301_train_complex_100001"}
    assert {g.names[i] for i in np.nonzero(finals != 8)[0]} == no_exit
    t = g.traj(0)
    assert t["name"] == "20251022_142001_train_very_simple_100001"
    assert len(t["hor"]) == 37
    assert t["pv"][0, 0] == 828.0 and t["pv"][0, 1] == 540.0
    assert t["pv"][-1, 0] == 812.0036311681246 and t["pv"][-1, 1] == 542.0 and t["state"][-1] == 8
    # the win must happen exactly on the last frame, not earlier
    for i in np.nonzero(finals == 8)[0]:
        tr = g.traj(int(i))
        assert tr["state"][-1] == 8 and tr["state"][-2] != 8, tr["name"]


def test_oracle_bc_replays_bit_exact(golden_replays):
    frames = sum(_check(golden_replays.traj(i)) for i in range(len(golden_replays)))
    assert frames == 31461


def test_oracle_generated_bit_exact(golden_generated):
    frames = sum(_check(golden_generated.traj(i)) for i in range(len(golden_generated)))
    assert frames == 40800


def test_oracle_curriculum_repositioning_bit_exact(golden_curriculum):
    """the reference's IntermediateGoalManager.apply_to_simulator (run unmodified by
    oracle/gen_golden.py) moves the exit switch / door at load and mid-episode; the restatement must follow."""
    g = golden_curriculum
    frames = sum(_check(g.traj(i)) for i in range(len(g)))
    assert frames == 16800 and len(g) == 56
    assert sum(len(g.traj(i)["moves"]) > 1 for i in range(len(g))) >= 10      # mid-episode moves are covered
    s = oracle.OracleSim(); s.load(g.traj(0)["map_data"])
    md = g.traj(0)["map_data"].copy(); md[1235:] = 0                          # a level without exit switch refuses
    s2 = oracle.OracleSim(); s2.load(md)
    with pytest.raises(RuntimeError):
        s2.apply_curriculum((100.0, 100.0), (200.0, 200.0))


def test_oracle_doors_and_scenario_maps_bit_exact(golden_doors):
    """Rows no reference replay exercises: regular doors (entity_door_regular.py:38-63, type-5 records injected into real
    levels, on the ninja's path and across drone / thwump paths), trap doors, shove thwumps; plus the reference's 17
    test_maps, 5 official maps and 8 bc_replays_tmp replays.  Fixtures: the unmodified reference (oracle/gen_golden_doors.py)."""
    g = golden_doors
    frames = sum(_check(g.traj(i)) for i in range(len(g)))
    assert frames == 52396 and len(g) == 163
    st = dict(zip((str(n) for n in g.d["stats_names"]), (int(v) for v in g.d["stats"])))
    assert st["door_reclosed"] >= 20 and st["trap_closed"] >= 10 and st["shwump_activated"] >= 10 and st["door_on_mover_path"] >= 20
    groups = {n.split("/")[0] for n in g.names}
    assert groups == {"door5", "trap8", "shwump28", "testmap", "official", "tmp"}


def test_pow_is_libm_not_multiply():
    """H1: x**2 is libm pow; the oracle must not have folded it to x*x."""
    import ctypes as C
    import math
    L = oracle.lib()
    L.osim_sq_probe.argtypes = [C.c_double]; L.osim_sq_probe.restype = C.c_double
    rng = np.random.default_rng(7)
    xs = rng.uniform(-1100, 1100, 200000)
    diff = 0
    for x in xs:
        x = float(x)
        assert L.osim_sq_probe(x) == x ** 2
        diff += (x ** 2) != (x * x)
    assert diff > 0, "expected pow(x,2) != x*x for some inputs on this libm"
    assert math.isfinite(diff)


def test_reset_and_fast_reset_replay_first_golden(golden_replays):
    """reset()/fast_reset() bring the env back to the post-load state (nsim.py:62-140)."""
    t = golden_replays.traj(0)
    s = oracle.OracleSim()
    s.load(t["map_data"])
    first = s.state()
    s.run(t["hor"], t["jump"])
    s.fast_reset()
    again = s.state()
    for a, b in zip(first, again):
        assert np.array_equal(np.asarray(a).view(np.int64), np.asarray(b).view(np.int64))
    nf, ni, _, _ = s.run(t["hor"], t["jump"])
    assert nf[-1][0] == 812.0036311681246 and ni[-1][0] == 8


def test_oracle_observation_primitives_match_reference():
    """get_ninja_state (40 floats) and get_valid_action_mask of the C restatement against
    the reference's own outputs (tests/golden/obs.npz from oracle/gen_golden_obs.py): bit-exact."""
    import os
    from conftest import Golden, GOLDEN
    d = dict(np.load(os.path.join(GOLDEN, "obs.npz")))          # NpzFile decompresses on every access
    G = {"bc_replays": Golden(os.path.join(GOLDEN, "bc_replays.npz")), "generated": Golden(os.path.join(GOLDEN, "generated.npz"))}
    frames = 0
    for k, (src, i) in enumerate(zip(d["src"], d["idx"])):
        t = G[str(src)].traj(int(i)); f0 = int(d["frame_off"][k])
        s = oracle.OracleSim(); s.load(t["map_data"])
        for f in range(len(t["hor"]) + 1):
            if f > 0:
                s.tick(t["hor"][f - 1], t["jump"][f - 1])
            assert np.array_equal(s.ninja_state().view(np.int64), d["ninja_state"][f0 + f].view(np.int64)), (src, i, f)
            assert np.array_equal(s.action_mask(f % 4), d["mask"][f0 + f]), (src, i, f)
            frames += 1
    assert frames == len(d["mask"]) == 11327


def _replay_rows_fixture():
    import os
    from conftest import Golden, GOLDEN
    return dict(np.load(os.path.join(GOLDEN, "replay_rows.npz"))), Golden(os.path.join(GOLDEN, "bc_replays.npz"))


def test_oracle_replay_rows_match_reference():
    """the simulator-derived columns of ReplayExecutor._get_raw_observation, recorded from the
    reference's own loop (oracle/gen_golden_replay_rows.py), against the restatement -- bit-exact (float32 rows
    compared after the same float64 -> float32 rounding)."""
    from nclone_b200.replay import decode_input_byte, INPUT_TO_ACTION
    d, G = _replay_rows_fixture()
    inp_all, inp_off = G.d["input_bytes"], G.d["inp_off"]
    for fs in (1, 4):
        off = d[f"row_off_fs{fs}"]
        for k, i in enumerate(d["idx"]):
            t = G.traj(int(i)); inp = inp_all[inp_off[i]:inp_off[i + 1]]
            hor, jump = decode_input_byte(inp)
            assert np.array_equal(hor, t["hor"]) and np.array_equal(jump, t["jump"])
            s = oracle.OracleSim(); s.load(t["map_data"])
            row = int(off[k])
            for f, b in enumerate(inp):
                s.tick(hor[f], jump[f])
                if (f + 1) % fs:
                    continue
                nf, ni, ef, ei = s.state()
                gs = np.append(s.ninja_state().astype(np.float32), max(0.0, (len(inp) - ni[schema.NINJA_I.index("frame")]) / len(inp))).astype(np.float32)
                assert np.array_equal(gs.view(np.int32), d[f"game_state_fs{fs}"][row].view(np.int32)), (fs, i, f)
                sw = [e for e in range(len(ei)) if ei[e, 0] == 4]
                pos = [nf[0], nf[1], 0, 0, 0, 0]
                act = 1
                if sw:
                    door = [e for e in range(len(ei)) if ei[e, 0] == 3][len(sw) - 1]     # parent of the last switch
                    pos[2:] = [ef[sw[-1], 0], ef[sw[-1], 1], ef[door, 0], ef[door, 1]]
                    act = int(not ei[sw[-1], 2])
                assert np.array_equal(np.array(pos, dtype=np.float64).view(np.int64), d[f"pos_fs{fs}"][row].view(np.int64)), (fs, i, f)
                assert list(d[f"flags_fs{fs}"][row]) == [act, int(ni[0] == 8), int(ni[0] in (6, 7)), INPUT_TO_ACTION[b & 7]], (fs, i, f)
                assert d[f"frame_fs{fs}"][row] == f
                row += 1
            assert row == off[k + 1]


def test_oracle_checkpoint_replay_matches_reference():
    """ActionReplayer.replay_to_checkpoint (reset + frame_skip ticks per stored action) run by the
    reference on re-used Simulators; the oracle's reset() keeps entity_counts the same way, so the whole record matches."""
    from nclone_b200 import ACTION_TO_INPUTS
    d, G = _replay_rows_fixture()
    for k, i in enumerate(d["idx"]):
        t = G.traj(int(i))
        s = oracle.OracleSim(); s.load(t["map_data"])
        for _ in range(5):
            s.tick(1, 1)
        s.reset()
        for a in d["ck_actions"][d["ck_off"][k]:d["ck_off"][k + 1]]:
            for _ in range(4):
                s.tick(*ACTION_TO_INPUTS[a])
        nf, ni, ef, ei = s.state()
        e0, e1 = d["ck_ent_off"][k], d["ck_ent_off"][k + 1]
        assert (nf[0], nf[1]) == tuple(d["ck_pos"][k])
        assert np.array_equal(nf.view(np.int64), d["ck_nin_f"][k].view(np.int64)) and np.array_equal(ni, d["ck_nin_i"][k])
        assert np.array_equal(ef.view(np.int64), d["ck_ent_f"][e0:e1].view(np.int64)) and np.array_equal(ei, d["ck_ent_i"][e0:e1])


def test_threaded_rollout_records_match_the_per_tick_loop():
    """oracle/baseline_runner.c: osim_rollout_records (the soak slice's checker) == the per-env, per-tick OracleSim loop."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from oracle import oracle as orc
    arr, lens = bench.load_maps("locked_doors")
    N, S, every, M = 96, 60, 10, 24
    rng = np.random.default_rng(5)
    em = rng.integers(0, M, size=N)
    acts = rng.integers(0, 6, size=(S, N)).astype(np.uint8)
    ticks, of, oi = orc.rollout_records(arr[:M], lens[:M], em, acts, 4, 300, every, 4)
    L = orc.lib()
    import ctypes as C
    total = 0
    nf = np.zeros(31); ni = np.zeros(27, dtype=np.int64)
    for e in range(N):
        s = orc.OracleSim(); md = arr[em[e], :lens[em[e]]]; s.load(md)
        for st in range(S):
            h, j = ((0, 0), (-1, 0), (1, 0), (0, 1), (-1, 1), (1, 1))[acts[st, e]]
            for _ in range(4):
                L.osim_tick(s.h, h, j); total += 1
                L.osim_get_state(s.h, nf.ctypes.data_as(C.POINTER(C.c_double)), ni.ctypes.data_as(C.POINTER(C.c_int64)), None, None)
                if ni[0] in (6, 7, 8):
                    break
            if (st + 1) % every == 0:
                c = st // every
                assert (nf.view(np.int64) == of[c, e].view(np.int64)).all() and (ni == oi[c, e]).all(), (e, st)
            if ni[0] in (6, 7, 8) or ni[26] >= 300:
                s.load(md)
    assert total == ticks


def test_reset_on_a_reused_simulator_vs_reference(golden_replays):
    """tests/golden/reset_reuse.npz (oracle/gen_golden_reset.py, unmodified reference): load -> 150 ticks -> reset() -> 400 ticks
    -> reset() -> 400 ticks on ONE Simulator, every bc_replays level with two or more death balls.  entity_counts outlives
    reset() (entities.py:213-216), so entity indices keep counting and death balls stop repelling each other after the first
    episode (entity_death_ball.py:153-166): 10 of the 19 levels then move differently from a fresh Simulator.  The oracle's
    reset() must follow the reference's, every frame, full record (entity `index` included)."""
    import os
    from conftest import GOLDEN
    d = np.load(os.path.join(GOLDEN, "reset_reuse.npz"))
    starts = list(np.cumsum(d["segments"])[:-1])
    assert int((d["differs_from_fresh"] >= 0).sum()) >= 8 and {41, 86} <= set(int(v) for v in d["level"])
    for k, lvl in enumerate(d["level"]):
        s = oracle.OracleSim(); s.load(golden_replays.traj(int(lvl))["map_data"])
        assert schema.state_hash(*s.state()) == d["hash_init"][k], lvl
        for t in range(d["hor"].shape[1]):
            if t in starts:
                s.reset()
                assert schema.state_hash(*s.state()) == d["hash_reset"][k, starts.index(t)], (lvl, t)
            s.tick(int(d["hor"][k, t]), int(d["jump"][k, t]))
            assert schema.state_hash(*s.state()) == d["hash_tick"][k, t], (lvl, t)

