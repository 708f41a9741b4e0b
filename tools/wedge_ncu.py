"""One wedged agent step of bench_data/wedges.npz alone on the GPU, on a solo warp, bracketed by cudaProfilerStart/Stop:
    ncu --profile-from-start off --set full --import-source on --warp-sampling-interval 0 -k regex:step_kernel -c 1 -o rep python tools/wedge_ncu.py K
then tools/ncu_lines.py rep.ncu-rep nclone_b200/libnppsim.so attributes the stall samples of that lone chain to source lines."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nclone_b200 import BatchedSimulator  # noqa: E402

k = int(sys.argv[1]) if len(sys.argv) > 1 else 0
handoff = None if "--home" in sys.argv else True
d = np.load(os.path.join(ROOT, "bench_data", "wedges.npz"))
md = d["maps"][k, :d["lens"][k]]
acts = d["acts"][d["act_off"][k]:d["act_off"][k + 1]]
sim = BatchedSimulator(1, handoff=handoff)
sim.load_maps([md])
sim.replay_to_checkpoint([acts[:-1]], frame_skip=4)
snap = [t.clone() for t in (sim.nin_f, sim.nin_i, sim.env_i, sim.ent_flag, sim.dyn_f, sim.dyn_i)]
a = torch.tensor([int(acts[-1])], dtype=torch.uint8, device="cuda")
for r in range(3):
    for t, s in zip((sim.nin_f, sim.nin_i, sim.env_i, sim.ent_flag, sim.dyn_f, sim.dyn_i), snap):
        t.copy_(s)
    torch.cuda.synchronize()
    if r == 2:
        torch.cuda.profiler.start()
    sim.step(a, 4, False)
    torch.cuda.synchronize()
    if r == 2:
        torch.cuda.profiler.stop()
print("queue", sim.solo_queue_length(), "status", sim.handoff_status())
