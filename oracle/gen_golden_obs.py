"""Golden vectors for the observation primitives, from the UNMODIFIED reference
(imported read-only from /root/reference).  Build-container only.

    python -m oracle.gen_golden_obs   ->  tests/golden/obs.npz

For a subset of the trajectories already in tests/golden/{bc_replays,generated}.npz (same maps, same
inputs) it records, after load and after every tick:
  ninja_state  [frames, 40] float64   NPlayHeadless.get_ninja_state()          nplay_headless.py:735-917
  mask         [frames, 6]  uint8     Ninja.get_valid_action_mask()            ninja.py:628-839
with sim._path_straightness_direction cycling None/"left"/"right"/"down" by frame (path_dir = frame % 4)
and the default masking mode ("hard").
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import refdump  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")
PATH_DIRS = [None, "left", "right", "down"]


class _Facade:
    """Just enough of NPlayHeadless for the unbound get_ninja_state."""

    def __init__(self, sim):
        self.sim = sim


def record(map_data, hor, jump):
    from nclone.nplay_headless import NPlayHeadless
    sim = refdump.new_sim(map_data)
    fac = _Facade(sim)
    T = len(hor)
    st = np.zeros((T + 1, 40)); mk = np.zeros((T + 1, 6), dtype=np.uint8)
    for f in range(T + 1):
        if f > 0:
            sim.tick(int(hor[f - 1]), int(jump[f - 1]))
        st[f] = NPlayHeadless.get_ninja_state(fac)
        sim._path_straightness_direction = PATH_DIRS[f % 4]
        mk[f] = sim.ninja.get_valid_action_mask()
    return st, mk


def main():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import Golden
    refdump.import_reference()
    picks = []
    G = Golden(os.path.join(OUT, "bc_replays.npz"))
    picks += [("bc_replays", i) for i in range(0, len(G), 6)]
    H = Golden(os.path.join(OUT, "generated.npz"))
    picks += [("generated", i) for i in range(0, len(H), 9)]
    src, idx, off, states, masks = [], [], [0], [], []
    for which, i in picks:
        t = (G if which == "bc_replays" else H).traj(i)
        md = [int(v) if float(v).is_integer() else float(v) for v in t["map_data"]]      # generators emit ints and a few floats
        st, mk = record(md, t["hor"], t["jump"])
        src.append(which); idx.append(i); off.append(off[-1] + len(st)); states.append(st); masks.append(mk)
    path = os.path.join(OUT, "obs.npz")
    np.savez_compressed(path, src=np.array(src), idx=np.array(idx, dtype=np.int32), frame_off=np.array(off, dtype=np.int64),
                        ninja_state=np.concatenate(states), mask=np.concatenate(masks))
    print(len(picks), "trajectories,", off[-1], "frames,", os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
