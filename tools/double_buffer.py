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


def run(n_envs, n_groups, steps, arr, lens):
    M = len(arr)
    sim = BatchedSimulator(n_envs)
    sim.load_maps([arr[k, :lens[k]] for k in range(M)], shard_env_to_map(n_envs, 0, M))
    groups = sim.groups(n_groups)
    streams = [torch.cuda.Stream() for _ in groups]
    gen = torch.Generator(device="cuda"); gen.manual_seed(1234)
    acts = torch.randint(0, 6, (64, n_envs), generator=gen, device="cuda", dtype=torch.uint8)
    bounds = np.cumsum([0] + [g.n_envs for g in groups])
    ga = [acts[:, bounds[k]:bounds[k + 1]].contiguous() for k in range(len(groups))]
    for k in range(320):
        sim.step(acts[k % 64], 4, True); sim.reset_terminal(2000)
    torch.cuda.synchronize()
    t0 = sim.total_ticks()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in streams:
        s.wait_event(e0)
    for k in range(steps):
        for g in range(len(groups)):
            with torch.cuda.stream(streams[g]):
                groups[g].step(ga[g][k % 64], 4, True); groups[g].reset_terminal(2000)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    return (sim.total_ticks() - t0) / ms * 1e3, ms / steps


def main():
    n_envs = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    arr, lens = bench.load_maps("simplest_with_mines")
    for groups in (1, 2, 4):
        v, ms = run(n_envs, groups, steps, arr, lens)
        print(f"envs {n_envs} groups {groups}: {v / 1e6:.1f} M env-steps/s, {ms:.3f} ms per agent step of the whole batch")


if __name__ == "__main__":
    main()
