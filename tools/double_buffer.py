"""Experiment: the same env batch stepped as G independent groups on G streams (double buffering, as an async RL
loop does: the policy acts on one group while the other steps).  A group's long-tailed step kernel then overlaps
the other groups' kernels instead of idling the GPU.  usage: python tools/double_buffer.py [envs] [steps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from nclone_b200 import BatchedSimulator  # noqa: E402
from nclone_b200.sharding import shard_env_to_map  # noqa: E402


def run(n_envs, groups, steps, arr, lens):
    M = len(arr)
    per = n_envs // groups
    full = shard_env_to_map(n_envs, 0, M)
    sims, streams, acts = [], [], []
    for g in range(groups):
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            sim = BatchedSimulator(per)
            sim.load_maps([arr[k, :lens[k]] for k in range(M)], full[g * per:(g + 1) * per])
            gen = torch.Generator(device="cuda"); gen.manual_seed(1234 + g)
            a = torch.randint(0, 6, (64, per), generator=gen, device="cuda", dtype=torch.uint8)
            for k in range(320):
                sim.step(a[k % 64], 4, True); sim.reset_terminal(2000)
        sims.append(sim); streams.append(s); acts.append(a)
    torch.cuda.synchronize()
    t0 = [s.total_ticks() for s in sims]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in streams:
        s.wait_event(e0)
    for k in range(steps):
        for g in range(groups):
            with torch.cuda.stream(streams[g]):
                sims[g].step(acts[g][k % 64], 4, True); sims[g].reset_terminal(2000)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    ticks = sum(s.total_ticks() - t for s, t in zip(sims, t0))
    ms = e0.elapsed_time(e1)
    return ticks / ms * 1e3, ms / steps


def main():
    n_envs = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    arr, lens = bench.load_maps("simplest_with_mines")
    for groups in (1, 2, 4):
        v, ms = run(n_envs, groups, steps, arr, lens)
        print(f"envs {n_envs} groups {groups}: {v / 1e6:.1f} M env-steps/s, {ms:.3f} ms per agent step of the whole batch")


if __name__ == "__main__":
    main()
