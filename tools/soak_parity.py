"""Soak parity run (GPU box): a larger random rollout than the test suite's against the C oracle.
usage: python tools/soak_parity.py [n_envs] [agent_steps] [seed]     (checker only: imports oracle/ like the tests do)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
from nclone_b200 import BatchedSimulator
from oracle import oracle as orc

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
seed = int(sys.argv[3]) if len(sys.argv) > 3 else 1
every = 20
arr, lens = bench.load_maps("simplest_with_mines")
arr2, lens2 = bench.load_maps("locked_doors")
stride = max(arr.shape[1], arr2.shape[1])
marr = np.zeros((len(arr) + len(arr2), stride)); marr[:len(arr), :arr.shape[1]] = arr; marr[len(arr):, :arr2.shape[1]] = arr2
mlens = np.concatenate([lens, lens2]).astype(np.int32)
maps = [marr[k, :mlens[k]] for k in range(len(marr))]
rng = np.random.default_rng(seed)
env_to_map = rng.integers(0, len(maps), size=N)
actions = rng.integers(0, 6, size=(steps, N)).astype(np.uint8)
hold = rng.integers(1, 8, size=(steps, N))
for st in range(1, steps):
    keep = (st % hold[st]) != 0
    actions[st, keep] = actions[st - 1, keep]
t0 = time.time()
ticks, of, oi = orc.rollout_records(marr, mlens, env_to_map, actions, 4, 2000, every, os.cpu_count() or 1)      # threaded C rollout
print(f"oracle: {N} envs x {steps} agent steps = {ticks:,} ticks in {time.time() - t0:.1f} s", flush=True)
sim = BatchedSimulator(N); sim.load_maps(maps, env_to_map)
ci = 0; coop = 0
for st in range(steps):
    sim.step(torch.from_numpy(actions[st]).cuda(), 4, True)
    coop += int(sim.warp_cost.sum().item()) - sim.warp_cost.numel()
    if (st + 1) % every == 0:
        nf, ni = of[ci], oi[ci]; ci += 1
        gf = sim.nin_f.cpu().numpy().T; gi = sim.nin_i.cpu().numpy().T.astype(np.int64)
        badf = np.nonzero((gf.view(np.int64) != nf.view(np.int64)).any(1))[0]
        badi = np.nonzero((gi != ni).any(1))[0]
        if badf.size or badi.size:
            print("MISMATCH at step", st, badf[:8], badi[:8], env_to_map[badf[:8]]); sys.exit(1)
    sim.reset_terminal(2000)
print(f"bit-exact: {N} envs x {steps} agent steps ({N * steps * 4:,} ticks), {coop:,} cooperative iterations on the device, seed {seed}")
