import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


class Golden:
    """Flat npz written by oracle/gen_golden.py, sliced per trajectory."""

    def __init__(self, path):
        self.d = np.load(path)
        self.names = [str(n) for n in self.d["names"]]

    def __len__(self):
        return len(self.names)

    def traj(self, i):
        d = self.d
        m0, m1 = d["map_off"][i], d["map_off"][i + 1]
        f0, f1 = d["frame_off"][i], d["frame_off"][i + 1]
        i0, i1 = d["inp_off"][i], d["inp_off"][i + 1]
        e0, e1 = d["ent_off"][i], d["ent_off"][i + 1]
        moves = ()
        if "mv_off" in d:
            moves = [tuple(r) for r in d["mv"][d["mv_off"][i]:d["mv_off"][i + 1]]]
        return dict(name=self.names[i], moves=moves, map_data=d["map_data"][m0:m1].astype(np.float64),
                    hor=d["hor"][i0:i1], jump=d["jump"][i0:i1],
                    pv=d["pv"][f0:f1], state=d["state"][f0:f1], hash=d["hash"][f0:f1],
                    final_nin_f=d["final_nin_f"][i], final_nin_i=d["final_nin_i"][i],
                    final_ent_f=d["final_ent_f"][e0:e1], final_ent_i=d["final_ent_i"][e0:e1])


@pytest.fixture(scope="session")
def golden_replays():
    return Golden(os.path.join(GOLDEN, "bc_replays.npz"))


@pytest.fixture(scope="session")
def golden_generated():
    return Golden(os.path.join(GOLDEN, "generated.npz"))


@pytest.fixture(scope="session")
def golden_curriculum():
    return Golden(os.path.join(GOLDEN, "curriculum.npz"))


@pytest.fixture(scope="session")
def golden_doors():
    """Synthesised regular-door / trap-door / shove-thwump levels and the reference's own scenario maps
    (oracle/gen_golden_doors.py, outputs of the unmodified reference)."""
    return Golden(os.path.join(GOLDEN, "doors.npz"))
