"""Differential fuzz: nsim_oracle.c vs the unmodified reference, random actions.

TEST INFRASTRUCTURE, build-container only (needs /root/reference).  Usage:

    python -m oracle.fuzz_vs_ref --maps-per-category 8 --episodes 4 --frames 400

Map sources: the reference's own generators (all categories of
map_generation/generator_configs.py:695-825, through GeneratorRegistry,
gym_environment/generator_registry.py:31-58) and the maps embedded in
bc_replays (which carry every entity type the reference instantiates).
Every frame's full canonical state (oracle/schema.py) is compared by hash.
"""
import argparse
import glob
import multiprocessing as mp
import os
import random
import sys
import zlib

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import schema, refdump, oracle  # noqa: E402


def random_inputs(rng, T):
    """Discrete(6) actions held for a random 1..12 frames (frame-skip 4 on average)."""
    hor = np.zeros(T, dtype=np.int8); jump = np.zeros(T, dtype=np.int8)
    t = 0
    while t < T:
        a = rng.randrange(6)
        hold = rng.choice((1, 2, 4, 4, 4, 8, 12))
        h, j = schema.ACTION_TO_INPUTS[a]
        hor[t:t + hold] = h; jump[t:t + hold] = j
        t += hold
    return hor, jump


def generated_map(category, seed):
    refdump.import_reference()
    # gym_environment/generator_registry.py:31-58 restated (that package imports gymnasium,
    # which is absent here): pick (generator, preset) with the registry's rng, build, generate.
    from nclone.map_generation.generator_factory import GeneratorFactory  # noqa
    from nclone.map_generation.generator_configs import CATEGORIES  # noqa
    gen_type, preset = random.Random(seed).choice(CATEGORIES[category].generators)
    map_gen = GeneratorFactory.create_from_preset(gen_type, preset, seed)
    map_gen.generate(seed=seed)
    return list(map_gen.map_data())


def curriculum_moves(rng, path_xy, T):
    """Curriculum repositioning events for a trajectory whose un-moved ninja path is path_xy [T+1, 2]: the switch
    (and then the door) are put ON the path the ninja is about to walk, so the moved entities are actually
    touched; some events come mid-episode (the reference warns but still moves, intermediate_goal_manager.py:843-856)."""
    def on_path(t_lo):
        t = rng.randrange(min(t_lo, T), T + 1)
        return float(path_xy[t, 0]) + rng.uniform(-8, 8), float(path_xy[t, 1]) + rng.uniform(-8, 8), t
    sx, sy, ts = on_path(5)
    dx, dy, _ = on_path(ts)
    moves = [(0, sx, sy, dx, dy)]
    if rng.random() < 0.5:
        t = rng.choice((1, 3, 9, 10, 11, 40, 120))
        if t < T:
            sx, sy, ts = on_path(t)
            dx, dy, _ = on_path(rng.choice((t, ts)))
            moves.append((t, sx, sy, dx, dy))
    return moves


def job(args):
    kind, a, b, episodes, frames = args
    curriculum = kind.endswith("+c")
    kind = kind[:-2] if curriculum else kind
    try:
        if kind == "gen":
            md = generated_map(a, b)
            name = f"{a}/{b}"
        else:
            md = list(schema.parse_replay(open(a, "rb").read())[0])
            name = os.path.basename(a)
        bad = []
        nframes = 0
        for ep in range(episodes):
            rng = random.Random(zlib.crc32(f"{name}:{ep}".encode()))
            hor, jump = random_inputs(rng, frames)
            moves = ()
            if curriculum:
                base = oracle.run_trajectory(md, hor, jump)
                moves = curriculum_moves(rng, base["nin_f"][:, :2], frames)
            try:
                ref = refdump.run_trajectory(md, hor, jump, moves=moves)
            except RuntimeError:
                if not curriculum:
                    raise
                try:                      # level without an exit switch: both sides must refuse
                    oracle.run_trajectory(md, hor, jump, moves=moves)
                    bad.append((name, ep, -1))
                except RuntimeError:
                    pass
                continue
            orc = oracle.run_trajectory(md, hor, jump, moves=moves)
            nframes += frames
            if not np.array_equal(ref["hash"], orc["hash"]):
                t = int(np.argmax(ref["hash"] != orc["hash"]))
                bad.append((name, ep, t))
        return name, nframes, bad, None
    except Exception as exc:  # generator failures are reported, not hidden
        return str(args[:3]), 0, [], repr(exc)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--maps-per-category", type=int, default=4)
    ap.add_argument("--episodes", type=int, default=3)
    ap.add_argument("--frames", type=int, default=400)
    ap.add_argument("--replay-maps", type=int, default=130)
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    ap.add_argument("--seed-base", type=int, default=9000)
    ap.add_argument("--curriculum", action="store_true", help="with curriculum exit switch / door repositioning events")
    a = ap.parse_args()
    refdump.import_reference()
    from nclone.map_generation.generator_configs import CATEGORIES  # noqa
    jobs = []
    for cat in CATEGORIES:
        for s in range(a.maps_per_category):
            jobs.append(("gen" + ("+c" if a.curriculum else ""), cat, a.seed_base + s, a.episodes, a.frames))
    files = sorted(glob.glob(os.path.join(refdump.REF_ROOT, "bc_replays", "*.replay")))[:a.replay_maps]
    for f in files:
        jobs.append(("replay" + ("+c" if a.curriculum else ""), f, 0, a.episodes, a.frames))
    oracle.build()
    with mp.Pool(a.procs) as pool:
        res = pool.map(job, jobs, chunksize=1)
    frames = sum(r[1] for r in res)
    bad = [b for r in res for b in r[2]]
    errs = [(r[0], r[3]) for r in res if r[3]]
    print(f"jobs {len(jobs)} frames {frames} mismatching-episodes {len(bad)} errors {len(errs)}")
    for b in bad[:20]:
        print("  MISMATCH", b)
    for e in errs[:10]:
        print("  ERROR", e)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
