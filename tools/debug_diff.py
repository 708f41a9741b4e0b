"""Debug helper (GPU box): run golden trajectories on the CUDA path and the oracle in lockstep,
print the first differing fields.  usage: python tools/debug_diff.py bc|gen name_substring [...]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from conftest import Golden, GOLDEN
from oracle import oracle, schema
from nclone_b200 import BatchedSimulator

G = Golden(os.path.join(GOLDEN, "bc_replays.npz" if sys.argv[1] == "bc" else "generated.npz"))
for pat in sys.argv[2:]:
    i = [k for k, n in enumerate(G.names) if pat in n][0]
    t = G.traj(i)
    sim = BatchedSimulator(1); sim.load_maps([t["map_data"]])
    o = oracle.OracleSim(); o.load(t["map_data"])
    print("==", t["name"], "frames", len(t["hor"]), "entities", o.num_entities)
    for f in range(len(t["hor"])):
        sim.tick(torch.tensor([t["hor"][f]], dtype=torch.int8), torch.tensor([t["jump"][f]], dtype=torch.int8))
        o.tick(t["hor"][f], t["jump"][f])
        a = sim.canonical_state()[0]; b = o.state()
        if schema.state_hash(*a) != schema.state_hash(*b):
            print(" first diff at frame", f + 1)
            for k in np.nonzero(a[0].view(np.int64) != b[0].view(np.int64))[0]:
                print("   nin_f", schema.NINJA_F[k], "cuda", repr(a[0][k]), "oracle", repr(b[0][k]))
            for k in np.nonzero(a[1] != b[1])[0]:
                print("   nin_i", schema.NINJA_I[k], "cuda", a[1][k], "oracle", b[1][k])
            de = np.nonzero((a[3] != b[3]).any(1) | (a[2].view(np.int64) != b[2].view(np.int64)).any(1))[0]
            for k in de[:8]:
                print("   ent", k, "cuda", a[3][k], a[2][k], "| oracle", b[3][k], b[2][k])
            break
    else:
        print(" no diff")
