"""Golden vectors for the replay -> dataset path and for checkpoint replay (f2), from
the UNMODIFIED reference (imported read-only from /root/reference).  Build-container only.

    python -m oracle.gen_golden_replay_rows   ->  tests/golden/replay_rows.npz

replay rows: the loop of ReplayExecutor.execute_replay (replay/replay_executor.py:431-460) is run here over a subset
of the bc_replays with the reference's own pieces -- Simulator, the NPlayHeadless getters (called unbound on a
stand-in that only carries `.sim`, because the real constructor wants pygame), and map_input_to_action /
decode_input_to_controls lifted from replay_executor.py's source by AST (the module itself imports gymnasium) -- and
the simulator-derived columns of _get_raw_observation (:470-600) are recorded at frame_skip 1 and 4:
  game_state [rows, 41] float32, pos [rows, 6] float64, flags [rows, 4] uint8, frame [rows] int32
  (layout of nppsim_replay_rows, include/nppsim.h).
checkpoints: ActionReplayer(frame_skip=4).replay_to_checkpoint (action_replayer.py:79-134, imported as is) on the
same levels with random Discrete(6) action sequences of ragged lengths; final ninja position + full canonical record.
"""
import ast
import os
import random
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import refdump, schema  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")


def lift_functions(path, names):
    """Compile the named top-level functions of a reference source file without importing the module."""
    tree = ast.parse(open(path).read())
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    ns = {}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return [ns[n] for n in names]


class _Facade:
    """`.sim` plus every NPlayHeadless method, unbound -> bound to this stand-in."""

    def __init__(self, sim):
        self.sim = sim

    def __getattr__(self, name):
        from nclone.nplay_headless import NPlayHeadless
        fn = getattr(NPlayHeadless, name)
        return lambda *a, **k: fn(self, *a, **k)


def replay_rows(map_data, inputs, frame_skip, decode, to_action):
    sim = refdump.new_sim(list(map_data))
    fac = _Facade(sim)
    gs, pos, flags, frames = [], [], [], []
    for frame_idx, b in enumerate(inputs):
        h, j = decode(int(b))
        sim.tick(h, j)                                          # NPlayHeadless.tick -> sim.tick (nplay_headless.py:322-327)
        if (frame_idx + 1) % frame_skip == 0:
            ninja_state = np.array(fac.get_ninja_state(), dtype=np.float32)
            est = len(inputs) or 2000
            time_remaining = max(0.0, (est - sim.frame) / est)
            gs.append(np.append(ninja_state[:40], time_remaining).astype(np.float32))
            px, py = fac.ninja_position(); sx, sy = fac.exit_switch_position(); dx, dy = fac.exit_door_position()
            pos.append((px, py, sx, sy, dx, dy))
            flags.append((int(fac.exit_switch_activated()), int(fac.ninja_has_won()), int(sim.ninja.has_died()), to_action(int(b))))
            frames.append(frame_idx)
    return (np.array(gs, dtype=np.float32).reshape(-1, 41), np.array(pos, dtype=np.float64).reshape(-1, 6),
            np.array(flags, dtype=np.uint8).reshape(-1, 4), np.array(frames, dtype=np.int32))


def main():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import Golden
    refdump.import_reference()
    to_action, decode = lift_functions(os.path.join(refdump.REF_ROOT, "nclone", "replay", "replay_executor.py"),
                                       ["map_input_to_action", "decode_input_to_controls"])
    G = Golden(os.path.join(OUT, "bc_replays.npz"))
    inp_all = G.d["input_bytes"]; inp_off = G.d["inp_off"]
    picks = list(range(0, len(G), 7))
    d = {"idx": np.array(picks, dtype=np.int32)}
    for fs in (1, 4):
        cols = [[], [], [], []]; off = [0]
        for i in picks:
            t = G.traj(i)
            r = replay_rows([int(v) for v in t["map_data"]], inp_all[inp_off[i]:inp_off[i + 1]], fs, decode, to_action)
            for c, v in zip(cols, r):
                c.append(v)
            off.append(off[-1] + len(r[3]))
        d[f"row_off_fs{fs}"] = np.array(off, dtype=np.int64)
        for name, c in zip(("game_state", "pos", "flags", "frame"), cols):
            d[f"{name}_fs{fs}"] = np.concatenate(c)
    # ---- checkpoints ------------------------------------------------------------------------------------
    from nclone.action_replayer import ActionReplayer
    from nclone.state_checkpoint import StateCheckpoint
    import inspect
    rng = random.Random(77)
    seqs, ck_pos, ck_rec = [], [], []
    rep = ActionReplayer(validate_position=False, frame_skip=4)
    for i in picks:
        t = G.traj(i)
        sim = refdump.new_sim([int(v) for v in t["map_data"]])
        for _ in range(rng.randrange(0, 30)):                 # the env has been used before the restore
            sim.tick(rng.choice((-1, 0, 1)), rng.randrange(2))
        seq = [rng.randrange(6) for _ in range(rng.choice((0, 1, 7, 30, 90)))]
        kw = {}
        for name, p in inspect.signature(StateCheckpoint).parameters.items():     # fill whatever the dataclass requires
            if name == "action_sequence":
                kw[name] = seq
            elif name == "ninja_position":
                kw[name] = (0.0, 0.0)
            elif p.default is inspect.Parameter.empty:
                kw[name] = 0
        res = rep.replay_to_checkpoint(sim, StateCheckpoint(**kw), skip_validation=True)
        seqs.append(np.array(seq, dtype=np.uint8))
        ck_pos.append(res.actual_position)
        ck_rec.append(refdump.canonical(sim))
    d["ck_off"] = np.cumsum([0] + [len(q) for q in seqs]).astype(np.int64)
    d["ck_actions"] = np.concatenate(seqs) if seqs else np.zeros(0, np.uint8)
    d["ck_pos"] = np.array(ck_pos, dtype=np.float64)
    # full final records (entity `index` keeps counting across reset() on a re-used Simulator, entities.py:213-216 --
    # a quirk the batched reset does not model -- so consumers compare every column but that one)
    d["ck_nin_f"] = np.stack([r[0] for r in ck_rec]); d["ck_nin_i"] = np.stack([r[1] for r in ck_rec])
    d["ck_ent_off"] = np.cumsum([0] + [len(r[3]) for r in ck_rec]).astype(np.int64)
    d["ck_ent_f"] = np.concatenate([r[2] for r in ck_rec]); d["ck_ent_i"] = np.concatenate([r[3] for r in ck_rec])
    path = os.path.join(OUT, "replay_rows.npz")
    np.savez_compressed(path, **d)
    print(len(picks), "replays; rows fs1", int(d["row_off_fs1"][-1]), "fs4", int(d["row_off_fs4"][-1]),
          "; checkpoint actions", int(d["ck_off"][-1]), ";", os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
