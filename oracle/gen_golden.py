"""Generate the committed golden fixtures under tests/golden/ from the UNMODIFIED
reference (imported read-only from /root/reference).  Build-container only.

    python -m oracle.gen_golden

Outputs (all inputs *and* expected outputs, so the GPU box needs nothing else):

tests/golden/bc_replays.npz
    every /root/reference/bc_replays/*.replay: the map bytes and input bytes as
    stored in the file (CompactReplay container, gameplay_recorder.py:60-127),
    and per frame (frame 0 = after load, then one per input byte):
    ninja (xpos, ypos, xspeed, yspeed) as float64, ninja state byte, and the
    64-bit digest of the full canonical record (oracle/schema.py) -- ninja's 31
    doubles + 27 ints and every entity's row -- plus the full final record.
tests/golden/generated.npz
    maps from the reference's generators (float map_data) and random-action
    episodes on them and on entity-rich replay maps, same per-frame records.

Known answers these reproduce: 128/130 replays end in state 8 on
their last frame; the first replay ends at (812.0036311681246, 542.0).
"""
import glob
import os
import random
import sys
import zlib

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import schema, refdump  # noqa: E402
from oracle.fuzz_vs_ref import generated_map, random_inputs, curriculum_moves  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def pack(trajs):
    """trajs: list of dict(name, map_data(np array), hor, jump, rec) -> flat npz dict."""
    d = {}
    d["names"] = np.array([t["name"] for t in trajs])
    d["map_off"] = np.cumsum([0] + [len(t["map_data"]) for t in trajs]).astype(np.int64)
    d["map_data"] = np.concatenate([np.asarray(t["map_data"]) for t in trajs])
    d["frame_off"] = np.cumsum([0] + [len(t["hor"]) + 1 for t in trajs]).astype(np.int64)
    d["inp_off"] = np.cumsum([0] + [len(t["hor"]) for t in trajs]).astype(np.int64)
    d["hor"] = np.concatenate([np.asarray(t["hor"], dtype=np.int8) for t in trajs])
    d["jump"] = np.concatenate([np.asarray(t["jump"], dtype=np.int8) for t in trajs])
    d["pv"] = np.concatenate([t["rec"]["nin_f"][:, :4] for t in trajs])
    d["state"] = np.concatenate([t["rec"]["nin_i"][:, 0] for t in trajs]).astype(np.uint8)
    d["hash"] = np.concatenate([t["rec"]["hash"] for t in trajs])
    d["final_nin_f"] = np.stack([t["rec"]["nin_f"][-1] for t in trajs])
    d["final_nin_i"] = np.stack([t["rec"]["nin_i"][-1] for t in trajs])
    d["ent_off"] = np.cumsum([0] + [t["rec"]["ent_i"].shape[1] for t in trajs]).astype(np.int64)
    d["final_ent_f"] = np.concatenate([t["rec"]["ent_f"][-1] for t in trajs])
    d["final_ent_i"] = np.concatenate([t["rec"]["ent_i"][-1] for t in trajs])
    return d


def curriculum():
    """tests/golden/curriculum.npz: random-action episodes in which the curriculum's
    IntermediateGoalManager.apply_to_simulator (the reference's own code, see refdump.apply_curriculum) moves the
    exit switch and door onto the ninja's path -- at load and, in half the episodes, again mid-episode.
    Extra arrays: mv_off [n+1], mv [*, 5] = (tick before which the move is applied, sx, sy, dx, dy)."""
    refdump.import_reference()
    from nclone.map_generation.generator_configs import CATEGORIES  # noqa
    files = sorted(glob.glob(os.path.join(refdump.REF_ROOT, "bc_replays", "*.replay")))
    sources = [(f"gen/{cat}/{41000 + s}", generated_map(cat, 41000 + s)) for cat in CATEGORIES for s in range(2)]
    sources += [(f"rand/{os.path.basename(f)[:-7]}", list(schema.parse_replay(open(f, "rb").read())[0])) for f in files[::13]]
    trajs, touched = [], 0
    for name, md in sources:
        for ep in range(2):
            rng = random.Random(zlib.crc32(f"{name}:{ep}".encode()))
            hor, jump = random_inputs(rng, 300)
            base = refdump.run_trajectory(md, hor, jump)
            moves = curriculum_moves(rng, base["nin_f"][:, :2], 300)
            try:
                rec = refdump.run_trajectory(md, hor, jump, full=True, moves=moves)
            except RuntimeError:
                continue                      # level without an exit switch
            touched += int(not np.array_equal(rec["hash"][1:], base["hash"][1:]) and (rec["ent_i"][-1][:, 2] == 0).any())
            trajs.append(dict(name=f"{name}/{ep}", map_data=np.asarray(md, dtype=np.float64), hor=hor, jump=jump, rec=rec,
                              moves=np.asarray(moves, dtype=np.float64)))
    d = pack(trajs)
    d["mv_off"] = np.cumsum([0] + [len(t["moves"]) for t in trajs]).astype(np.int64)
    d["mv"] = np.concatenate([t["moves"] for t in trajs])
    np.savez_compressed(os.path.join(OUT, "curriculum.npz"), **d)
    won = int((d["final_nin_i"][:, 0] == 8).sum())
    print(f"curriculum: {len(trajs)} episodes, {len(d['hor'])} frames, {touched} with an entity taken after a move, {won} won;",
          os.path.getsize(os.path.join(OUT, "curriculum.npz")) // 1024, "KiB")


def main():
    os.makedirs(OUT, exist_ok=True)
    if "--curriculum-only" in sys.argv:
        return curriculum()
    refdump.import_reference()
    # ---- bc_replays ---------------------------------------------------------
    files = sorted(glob.glob(os.path.join(refdump.REF_ROOT, "bc_replays", "*.replay")))
    trajs = []
    raw_inputs = []
    for f in files:
        md, inp, ok = schema.parse_replay(open(f, "rb").read())
        hj = [schema.decode_input_byte(b) for b in inp]
        hor = [h for h, _ in hj]; jump = [j for _, j in hj]
        rec = refdump.run_trajectory(list(md), hor, jump, full=True)
        trajs.append(dict(name=os.path.basename(f)[:-7], map_data=np.frombuffer(md, dtype=np.uint8),
                          hor=hor, jump=jump, rec=rec))
        raw_inputs.append(np.frombuffer(inp, dtype=np.uint8))
    d = pack(trajs)
    d["input_bytes"] = np.concatenate(raw_inputs)
    np.savez_compressed(os.path.join(OUT, "bc_replays.npz"), **d)
    won = int((d["final_nin_i"][:, 0] == 8).sum())
    print(f"bc_replays: {len(trajs)} replays, {len(d['hor'])} frames, {won} end in state 8")
    assert trajs[0]["rec"]["nin_f"][-1][0] == 812.0036311681246 and trajs[0]["rec"]["nin_f"][-1][1] == 542.0

    # ---- generated maps + random actions -------------------------------------
    from nclone.map_generation.generator_configs import CATEGORIES  # noqa
    trajs = []
    for cat in CATEGORIES:
        for s in range(4):
            seed = 31000 + s
            md = generated_map(cat, seed)
            for ep in range(2):
                hor, jump = random_inputs(random.Random(1000 * seed + ep), 300)
                rec = refdump.run_trajectory(md, hor, jump, full=True)
                trajs.append(dict(name=f"gen/{cat}/{seed}/{ep}", map_data=np.asarray(md, dtype=np.float64),
                                  hor=hor, jump=jump, rec=rec))
    # entity-rich replay maps under random actions: pick the maps with most entity kinds
    scored = []
    for f in files:
        md = schema.parse_replay(open(f, "rb").read())[0]
        kinds = set(md[i] for i in range(1235, len(md) - 4, 5))
        scored.append((len(kinds & {5, 6, 8, 10, 11, 14, 17, 20, 24, 25, 26, 28}), f))
    scored.sort(reverse=True)
    for _, f in scored[:24]:
        md = list(schema.parse_replay(open(f, "rb").read())[0])
        for ep in range(2):
            hor, jump = random_inputs(random.Random(zlib.crc32(os.path.basename(f).encode()) % 100000 + ep), 400)
            rec = refdump.run_trajectory(md, hor, jump, full=True)
            trajs.append(dict(name=f"rand/{os.path.basename(f)[:-7]}/{ep}", map_data=np.asarray(md, dtype=np.float64),
                              hor=hor, jump=jump, rec=rec))
    d = pack(trajs)
    np.savez_compressed(os.path.join(OUT, "generated.npz"), **d)
    print(f"generated: {len(trajs)} episodes, {len(d['hor'])} frames")
    for n in ("bc_replays.npz", "generated.npz"):
        print(n, os.path.getsize(os.path.join(OUT, n)) // 1024, "KiB")
    curriculum()


if __name__ == "__main__":
    main()
