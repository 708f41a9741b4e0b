"""tests/golden/reset_reuse.npz -- Simulator.reset() on a RE-USED Simulator, from the unmodified reference.  Build container only.

    python -m oracle.gen_golden_reset

The reference's callers (NPlayHeadless.reset, nplay_headless.py:272-276; ActionReplayer.replay_to_checkpoint,
action_replayer.py:99) keep one Simulator for its whole life and call reset() on it.  sim.entity_counts is never cleared
(entities.py:213-216), so every entity created after the first episode carries an `index` that goes on counting, and death
balls -- which repel only balls with a larger index while `index + 1 < map_data[1200]` (entity_death_ball.py:153-166) --
stop repelling each other from the second episode on.  This fixture pins that behaviour:

  every bc_replays level with at least two death balls: Simulator() -> load -> 150 ticks -> reset() -> 400 ticks -> reset() ->
  400 ticks under seeded random inputs (each held 4 ticks), the 64-bit digest of the full canonical record (oracle/schema.py,
  entity `index` included) after the load, after every tick and after each reset().

`differs_from_fresh[k]` is the first tick (0-based, over all 950) at which the same inputs on a Simulator that is re-created
for every episode give different death-ball positions, -1 if never: how many levels actually show the effect.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import schema, refdump  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "reset_reuse.npz")
SEGMENTS = (150, 400, 400)


def balls(sim):
    return [(e.xpos, e.ypos) for e in sim.entity_dic.get(25, [])]


def main():
    g = np.load(os.path.join(ROOT, "tests", "golden", "bc_replays.npz"))
    n = len(g["names"])
    levels = []
    for i in range(n):
        md = g["map_data"][g["map_off"][i]:g["map_off"][i + 1]]
        if int(md[1200]) >= 2:
            levels.append(i)
    T = sum(SEGMENTS)
    starts = np.cumsum((0,) + SEGMENTS[:-1])[1:]                  # ticks before which reset() is called
    rng = np.random.default_rng(20260311)
    hor = np.repeat(rng.integers(-1, 2, size=(len(levels), (T + 3) // 4)), 4, axis=1)[:, :T].astype(np.int8)
    jump = np.repeat(rng.integers(0, 2, size=(len(levels), (T + 3) // 4)), 4, axis=1)[:, :T].astype(np.int8)
    h_init = np.zeros(len(levels), dtype=np.uint64)
    h_tick = np.zeros((len(levels), T), dtype=np.uint64)
    h_reset = np.zeros((len(levels), len(starts)), dtype=np.uint64)
    differs = np.full(len(levels), -1, dtype=np.int64)
    n_balls = np.zeros(len(levels), dtype=np.int64)
    for k, i in enumerate(levels):
        md = [int(v) for v in g["map_data"][g["map_off"][i]:g["map_off"][i + 1]]]
        sim = refdump.new_sim(md)                                   # the long-lived Simulator
        fresh = refdump.new_sim(md)                                 # re-created per episode, for `differs_from_fresh`
        n_balls[k] = len(balls(sim))
        h_init[k] = schema.state_hash(*refdump.canonical(sim))
        for t in range(T):
            if t in starts:
                sim.reset()
                h_reset[k, list(starts).index(t)] = schema.state_hash(*refdump.canonical(sim))
                fresh = refdump.new_sim(md)
            sim.tick(int(hor[k, t]), int(jump[k, t]))
            fresh.tick(int(hor[k, t]), int(jump[k, t]))
            h_tick[k, t] = schema.state_hash(*refdump.canonical(sim))
            if differs[k] < 0 and balls(sim) != balls(fresh):
                differs[k] = t
        print(f"level {i:3d} ({str(g['names'][i])}): {n_balls[k]} death balls, differs from a fresh Simulator at tick {differs[k]}")
    np.savez_compressed(OUT, level=np.array(levels, dtype=np.int64), segments=np.array(SEGMENTS, dtype=np.int64), hor=hor, jump=jump,
                        hash_init=h_init, hash_tick=h_tick, hash_reset=h_reset, differs_from_fresh=differs, n_balls=n_balls)
    print(f"{OUT}: {len(levels)} levels, {int((differs >= 0).sum())} show the effect")


if __name__ == "__main__":
    main()
