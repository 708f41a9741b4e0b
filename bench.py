#!/usr/bin/env python
"""bench.py -- env-steps/s of the batched Simulator.tick path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c4]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...     (one rank per GPU)

One "step" = one agent step of the whole per-GPU batch: a fused launch of frame_skip=4
Simulator.tick calls for every env (with the env wrapper's early exit on win/death) followed by
the episode-turnover launch (reset envs that won, died or reached 2,000 frames).  env-steps are
physics ticks actually executed, counted on the device.

Default workload = BASELINE.json configs[2], the one the metric is quoted on: 65,536 envs/GPU over
heterogeneous generated maps (4,096 distinct `simplest_with_mines` levels from the reference's
generator, bench_data/, x16 envs each with independent action streams), uniform Discrete(6) actions.
`value` is measured with actions resident in HBM; `e2e` goes through the host-buffer C-ABI entry
(nppsim_step_host: pinned host actions -> H2D -> kernels -> D2H of pos/vel/state) every step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAME_SKIP = 4
MAX_FRAMES = 2000
WORKLOADS = {
    # name: (fixture, envs per GPU, description)
    "c2": ("first_replay", 4096, "configs[1]: 4,096 envs on one level (exit + switch only), random Discrete(6)"),
    "c3": ("simplest_with_mines", 65536, "configs[2]: 65,536 envs/GPU over 4,096 distinct generated simplest_with_mines levels (x16), random Discrete(6)"),
    "c4": ("locked_doors", 131072, "configs[3]: 131,072 envs/GPU over 2,048 distinct generated levels with a locked door (x64), random Discrete(6)"),
    "c5": ("bc_replays", 65536, "configs[4] throughput form (SURVEY 8(d) config 5): the 130 bc_replays levels (all 16 entity kinds, full kernel) replicated to 65,536 envs/GPU, random Discrete(6)"),
}


def load_maps(fixture):
    if fixture == "first_replay":
        g = np.load(os.path.join(ROOT, "tests", "golden", "bc_replays.npz"))
        md = g["map_data"][g["map_off"][0]:g["map_off"][1]].astype(np.float64)
        return md[None, :].copy(), np.array([len(md)], dtype=np.int32)
    if fixture == "bc_replays":
        g = np.load(os.path.join(ROOT, "tests", "golden", "bc_replays.npz"))
        off = g["map_off"]
        lens = np.diff(off).astype(np.int32)
        arr = np.zeros((len(lens), int(lens.max())), dtype=np.float64)
        for k in range(len(lens)):
            arr[k, :lens[k]] = g["map_data"][off[k]:off[k + 1]]
        return arr, lens
    d = np.load(os.path.join(ROOT, "bench_data", fixture + ".npz"))
    arr = d["ints"].astype(np.float64)
    arr[d["frac_row"], d["frac_col"]] = d["frac_val"]
    return arr, d["lens"].astype(np.int32)


def entity_stats(arr, lens):
    """Per map: (#entities that never move, sum over moving ones of (17 + 2*state_bytes)) for the
    SURVEY.md 8(d) algorithmic-bytes formula."""
    dyn_bytes = {14: 42, 26: 42, 17: 32, 25: 32, 20: 9, 28: 35}
    e_static = np.zeros(len(arr)); dyn = np.zeros(len(arr))
    for k in range(len(arr)):
        m = arr[k, :lens[k]]
        i = 1235
        while i + 4 < len(m):
            t = int(m[i])
            if t in dyn_bytes:
                dyn[k] += 17 + 2 * dyn_bytes[t]
            elif t in (1, 2, 3, 5, 6, 8, 10, 11, 21, 24):
                e_static[k] += 2 if t == 3 else 1
            i += (10 if (i + 9 < len(m) and m[i + 7] != 0 and m[i + 8] == 0 and m[i + 9] == 0) else 9) if t in (6, 8) else 5
    return e_static, dyn


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.proc = index, [], False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
                if self.stop_flag:
                    break
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
                for n, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def cpu_baseline(arr, lens, threads, target_seconds, seed=99):
    """Oracle port (oracle/nsim_oracle.c) on the host cores: same level set, same action
    distribution, same frame-skip / early-exit / turnover policy.  Bounded sample."""
    from oracle import oracle
    rng = np.random.default_rng(seed)
    probe_maps = min(len(arr), max(threads, 8))
    acts = rng.integers(0, 6, size=(probe_maps, 25)).astype(np.uint8)
    ticks, secs = oracle.baseline_run(arr[:probe_maps], lens[:probe_maps], acts, FRAME_SKIP, MAX_FRAMES, threads)
    rate = ticks / max(secs, 1e-6)
    n_steps = 500                                   # agent steps per env: several episodes each
    n_maps = int(min(len(arr), max(threads * 2, target_seconds * rate / (n_steps * FRAME_SKIP))))
    n_maps -= n_maps % threads if n_maps > threads else 0
    # the level set is finite: once every level is in the sample, lengthen the rollouts instead (about target_seconds of work)
    n_steps = int(min(20000, max(n_steps, target_seconds * rate / (n_maps * FRAME_SKIP))))
    acts = rng.integers(0, 6, size=(n_maps, n_steps)).astype(np.uint8)
    ticks, secs = oracle.baseline_run(arr[:n_maps], lens[:n_maps], acts, FRAME_SKIP, MAX_FRAMES, threads)
    return ticks / secs, f"{n_maps} envs (first {n_maps} levels of the workload) x {n_steps} agent steps = {ticks} ticks in {secs:.2f} s on {threads} threads"


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path.  The reference is pure
    Python and cannot travel to the GPU box, so this times its C restatement (oracle/, pinned
    bit-exact to the reference) on all host threads: cpu_baseline.kind = "port"."""
    if rank != 0:
        return
    fixture, n_envs, desc = WORKLOADS[args.workload]
    arr, lens = load_maps(fixture)
    threads = os.cpu_count() or 1
    per_step = max(2.0, min(20.0, 150.0 / max(1, args.steps + args.warmup)))
    vals, sample = [], ""
    t0 = time.perf_counter()
    for s in range(args.warmup + args.steps):
        v, sample = cpu_baseline(arr, lens, threads, per_step, seed=s)
        if s >= args.warmup:
            vals.append(v)
    wall = time.perf_counter() - t0
    value = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * wall / max(1, args.steps + args.warmup),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "frame_skip": FRAME_SKIP, "max_frames": MAX_FRAMES, "host_threads": threads},
        "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": threads, "kind": "port", "sample": "per step: " + sample},
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="override envs per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between timed steps")
    ap.add_argument("--async-groups", type=int, default=4, help="also time the batch stepped as this many independent groups on separate streams (0: skip)")
    ap.add_argument("--placement", default="scattered", choices=["scattered", "adjacent"], help="caller's env -> level assignment: replicas of a level n_levels envs apart, or consecutive")
    ap.add_argument("--spread", type=int, default=300, help="untimed agent steps run first so envs sit at mixed episode phases")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    from nclone_b200 import BatchedSimulator

    fixture, n_envs, desc = WORKLOADS[args.workload]
    if args.envs:
        n_envs = args.envs
    arr, lens = load_maps(fixture)
    M = len(arr)
    # shard: every rank owns its own env batch (weak scaling); ranks use different level rotations and seeds
    from nclone_b200.sharding import shard_env_to_map, reduce_rollout
    env_to_map = shard_env_to_map(n_envs, rank, M, adjacent=(args.placement == "adjacent"))
    sim = BatchedSimulator(n_envs, device=dev)
    sim.load_maps([arr[k, :lens[k]] for k in range(M)], env_to_map)

    gen = torch.Generator(device=dev); gen.manual_seed(1234 + rank)
    POOL = 64
    actions = torch.randint(0, 6, (POOL, n_envs), generator=gen, device=dev, dtype=torch.uint8)
    actions_host = actions.cpu().pin_memory()
    flush = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def one_step(s):
        sim.step(actions[s % POOL], FRAME_SKIP, True)
        sim.reset_terminal(MAX_FRAMES)

    # spread the envs over episode phases before timing: 300 agent steps (about 1.5 mean episodes)
    for s in range(args.spread):
        one_step(s)
    for s in range(args.warmup):
        one_step(s)
    torch.cuda.synchronize(dev)

    sampler = ClockSampler(torch.cuda.current_device() if "CUDA_VISIBLE_DEVICES" not in os.environ else local_rank)
    if rank == 0:
        sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ticks0 = sim.total_ticks()
    launches0 = sim.launch_count()
    handoffs0 = sim.handoff_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    for s in range(args.steps):
        if flush is not None:
            flush.zero_()                       # untimed: evict the state arrays and map blobs from L2
        ev[s][0].record()
        sim.step(actions[s % POOL], FRAME_SKIP, True)
        ev[s][1].record()
        sim.reset_terminal(MAX_FRAMES)
        ev[s][2].record()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    launches = sim.launch_count() - launches0
    handoff_info = {"enabled": sim.handoff is not None, "mid_launch_per_step": (sim.handoff_count() - handoffs0) / args.steps,
                    "solo_start_queue": sim.solo_queue_length(), "status": sim.handoff_status()}
    step_ms = np.array([e[0].elapsed_time(e[1]) for e in ev]); reset_ms = np.array([e[1].elapsed_time(e[2]) for e in ev])
    elapsed_ms = float(step_ms.sum() + reset_ms.sum())
    ticks = sim.total_ticks() - ticks0

    # ---- e2e: host buffers through the C ABI, copies inside the timed region -------------------------
    pv = np.zeros((4, n_envs)); st = np.zeros(n_envs, dtype=np.int32)
    pv_t = torch.from_numpy(pv).pin_memory(); st_t = torch.from_numpy(st).pin_memory()
    pv, st = pv_t.numpy(), st_t.numpy()
    ah = actions_host.numpy()
    for s in range(3):
        sim.step_host(ah[s % POOL], pv, st, FRAME_SKIP, True); sim.reset_terminal(MAX_FRAMES)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    t1 = sim.total_ticks()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(args.steps):
        sim.step_host(ah[s % POOL], pv, st, FRAME_SKIP, True)
        sim.reset_terminal(MAX_FRAMES)
    e1.record()
    torch.cuda.synchronize(dev)
    e2e_ms = e0.elapsed_time(e1)
    e2e_ticks = sim.total_ticks() - t1

    # ---- the same batch as independent groups on separate streams (BatchedSimulator.groups): what an asynchronous RL loop
    # does -- every group's launches in stream order, no per-step synchronisation of the whole batch, so one group's
    # long-tailed step kernel overlaps the other groups' kernels.  Reported beside `value`, never instead of it.
    grp_ms, grp_ticks = 0.0, 0
    if args.async_groups > 1:
        groups = sim.groups(args.async_groups)
        gstreams = [torch.cuda.Stream(device=dev) for _ in groups]
        bounds = np.cumsum([0] + [g.n_envs for g in groups])
        ga = [actions[:, bounds[k]:bounds[k + 1]].contiguous() for k in range(len(groups))]
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        t2 = sim.total_ticks()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for gs in gstreams:
            gs.wait_event(g0)
        for s in range(args.steps):
            for k, g in enumerate(groups):
                with torch.cuda.stream(gstreams[k]):
                    g.step(ga[k][s % POOL], FRAME_SKIP, True)
                    g.reset_terminal(MAX_FRAMES)
        for gs in gstreams:
            torch.cuda.current_stream(dev).wait_stream(gs)
        g1.record()
        torch.cuda.synchronize(dev)
        grp_ms = g0.elapsed_time(g1); grp_ticks = sim.total_ticks() - t2
    clocks = sampler.finish() if rank == 0 else None

    # ---- reduce over ranks: max time, summed work ------------------------------------------------------
    # end-of-rollout reduction: the only collective (NCCL), outside the timed regions
    (elapsed_ms, e2e_ms, grp_ms), (ticks_all, e2e_ticks_all, grp_ticks_all), stats = reduce_rollout(
        [elapsed_ms, e2e_ms, grp_ms], [ticks, e2e_ticks, grp_ticks], sim.stats, dist if world > 1 else None)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    value = ticks_all / (elapsed_ms / 1000.0)
    e2e_value = e2e_ticks_all / (e2e_ms / 1000.0)

    # ---- roofline of the dominant kernel (step_kernel): SURVEY.md 8(d) algorithmic bytes ---------------
    e_static, dyn = entity_stats(arr, lens)
    bytes_per_tick = float(np.mean((892 + 19 * e_static + dyn)[env_to_map]))
    ticks_per_launch = ticks / args.steps
    step_s = float(step_ms.mean()) / 1000.0
    achieved = bytes_per_tick * ticks_per_launch / step_s / 1e9
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback"
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "step_kernel_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if tj.get("workload") == args.workload and tj.get("envs") == n_envs:
            traffic = tj.get("dram_bytes_per_launch")
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "kernel": "step_kernel", "kernel_ms": float(step_ms.mean()), "algorithmic_bytes_per_env_step": bytes_per_tick,
                "env_steps_per_launch": ticks_per_launch, "peak_source": peak_src,
                "note": "branchy fp64 scalar code: bound by issue slots / fp64 latency, not HBM (DESIGN.md); frac is reported as required"}

    cpu = None
    if not args.no_cpu_baseline and world == 1:          # rank 0 at N=1 only
        threads = os.cpu_count() or 1
        v, sample = cpu_baseline(arr, lens, threads, 25.0)
        cpu = {"value": v, "unit": "env-steps/s", "cores": threads, "kind": "port", "sample": sample}

    line = {
        "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "envs_per_gpu": n_envs, "frame_skip": FRAME_SKIP, "max_frames": MAX_FRAMES,
                   "distinct_levels": int(M), "l2": "inputs (state + level blobs) exceed L2 only partly; L2 flushed between timed steps (256 MiB memset, untimed)" if flush is not None else "not flushed",
                   "sq_mode": "exact glibc pow(x,2) emulation"},
        "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": int(n_envs), "d2h_bytes_per_step": int(n_envs * (4 * 8 + 4)),
                "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": int(launches),
        "kernel_ms": {"step_kernel": float(step_ms.mean()), "reset_kernel": float(reset_ms.mean())},
        "agent_steps_per_s": value / FRAME_SKIP,
        "episode_stats": {"episodes": int(stats[0]), "wins": int(stats[1]), "deaths": int(stats[2]), "truncations": int(stats[3])},
        "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks, "handoff": handoff_info,
    }
    if grp_ms > 0:
        line["async_groups"] = {"groups": args.async_groups, "value": grp_ticks_all / (grp_ms / 1000.0), "unit": "env-steps/s",
                                "ms_per_step": grp_ms / args.steps,
                                "note": "same batch, same work, stepped as independent groups on separate streams (no L2 flush, no per-step sync)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
