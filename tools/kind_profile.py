"""Cycles per roaming-entity update by kind and phase (-DNPP_PROFILE build; GPU box).
usage: python tools/kind_profile.py [envs] [--lib=variants/profile.so]   -- the 130 bc_replays levels, random actions"""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
lib = next((a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--lib=")), os.path.join(ROOT, "variants", "profile.so"))
os.environ["NPPSIM_LIB"] = lib
import torch
import bench
from nclone_b200 import BatchedSimulator
from nclone_b200._lib import lib as L
n_envs = next((int(a) for a in sys.argv[1:] if a.isdigit()), 65536)
arr, lens = bench.load_maps("bc_replays")
sim = BatchedSimulator(n_envs)
sim.load_maps([arr[k, :lens[k]] for k in range(len(arr))], np.arange(n_envs) % len(arr))
gen = torch.Generator(device="cuda"); gen.manual_seed(1)
acts = torch.randint(0, 6, (64, n_envs), generator=gen, device="cuda", dtype=torch.uint8)
for s in range(60):
    sim.step(acts[s % 64], 4, True); sim.reset_terminal(2000)
lib_ = L()
lib_.nppsim_debug_kind_cycles.argtypes = [C.c_void_p, C.c_int]
out = (C.c_ulonglong * 128)()
lib_.nppsim_debug_kind_cycles(None, 1)
K = 20
for s in range(K):
    sim.step(acts[s % 64], 4, True); sim.reset_terminal(2000)
lib_.nppsim_debug_kind_cycles(out, 0)
v = np.array(list(out), dtype=np.float64).reshape(2, 32, 2)
names = {5: "regular door", 10: "drone (zap)", 11: "bounce block", 12: "thwump", 14: "death ball", 15: "mini drone", 16: "shove thwump"}
print(f"{n_envs} envs on the 130 bc_replays levels, {K} launches; per warp-level call (lane 0's clock; all 32 lanes run the same level)")
for ph, pn in enumerate(("move", "think")):
    for k in range(32):
        if v[ph, k, 1] > 0:
            print(f"  {pn:5s} {names.get(k, 'kind %d (no-op in this phase)' % k):32s} calls {int(v[ph, k, 1]):10d}  cycles per call {v[ph, k, 0] / v[ph, k, 1]:9.0f}  share of the phase {100 * v[ph, k, 0] / v[ph, :, 0].sum():5.1f} %")
