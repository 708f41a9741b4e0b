#!/usr/bin/env python
"""bench.py -- env-steps/s of the batched Simulator.tick path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c4|c5]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...     (one rank per GPU)

One "step" = one agent step of the whole per-GPU batch: a fused launch of frame_skip=4 Simulator.tick calls for every
env (with the env wrapper's early exit on win/death) followed by the episode-turnover launch (reset envs that won, died
or reached 2,000 frames).  env-steps are physics ticks actually executed, counted on the device.

Default workload = BASELINE.json configs[2], the one the metric is quoted on: 65,536 envs/GPU over heterogeneous
generated maps (4,096 distinct `simplest_with_mines` levels from the reference's generator, bench_data/, x16 envs each
with independent action streams; the caller numbers the envs scattered, env i -> level i % 4096), uniform Discrete(6)
actions.  `value` is measured with actions resident in HBM; `e2e` goes through the host-buffer C-ABI entry
(nppsim_step_host: pinned host actions -> H2D -> kernels -> one D2H of pos/vel/state) every step, `e2e_obs` the same
with the 40-float ninja state and the action mask in the block (what NppEnvironment.step pulls per step).

The default line also carries, as secondary blocks measured with short step counts in the same process: config2 (4,096
envs on one level), config4 (131,072 envs/GPU, levels with a locked door: BASELINE.json's target config) and full_kernel
(config 5's throughput form: the 130 bc_replays levels, all 16 entity kinds, x 65,536 envs).
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
    """Per map, for the SURVEY.md 8(d) algorithmic-bytes formula: #entities that never move, sum over moving ones of
    (17 + 2*state_bytes), sum over moving ones of state_bytes."""
    dyn_bytes = {14: 42, 26: 42, 17: 32, 25: 32, 20: 9, 28: 35}
    e_static = np.zeros(len(arr)); dyn = np.zeros(len(arr)); dyn_state = np.zeros(len(arr)); n_dyn = np.zeros(len(arr))
    for k in range(len(arr)):
        m = arr[k, :lens[k]]
        i = 1235
        while i + 4 < len(m):
            t = int(m[i])
            if t in dyn_bytes:
                dyn[k] += 17 + 2 * dyn_bytes[t]; dyn_state[k] += dyn_bytes[t]; n_dyn[k] += 1
            elif t in (1, 2, 3, 5, 6, 8, 10, 11, 21, 24):
                e_static[k] += 2 if t == 3 else 1
            i += (10 if (i + 9 < len(m) and m[i + 7] != 0 and m[i + 8] == 0 and m[i + 9] == 0) else 9) if t in (6, 8) else 5
    return e_static, dyn, dyn_state, n_dyn


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


def config_dict(workload, n_envs, n_levels, placement, flush):
    """`config` of the JSON line: the same keys and values in both arms (ours / reference) for the same workload."""
    return {"workload": WORKLOADS[workload][2], "envs_per_gpu": int(n_envs), "frame_skip": FRAME_SKIP, "max_frames": MAX_FRAMES,
            "distinct_levels": int(n_levels), "caller_env_order": placement,
            "l2": ("GPU arm: L2 flushed between timed steps (256 MiB memset, untimed); state + level blobs exceed L2 only partly" if flush else "not flushed"),
            "arithmetic": "IEEE double, x**2 == glibc pow(x, 2) (GPU arm: exact emulation)"}


def headline_config(args, n_levels):
    """`config` of the headline workload from the parsed command line -- the ONE place both arms get it from."""
    return config_dict(args.workload, args.envs or WORKLOADS[args.workload][1], n_levels, args.placement, not args.no_flush)


# ---------------------------------------------------------------------------------------------------- CPU baselines
def cpu_baseline_port(arr, lens, threads, target_seconds, seed=99):
    """Oracle port (oracle/nsim_oracle.c) on the host cores: same level set, same action distribution, same frame-skip /
    early-exit / turnover policy.  Bounded sample."""
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


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def usable_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_baseline_reference(arr, lens, processes, seconds, seed=0):
    """The reference's own nsim.Simulator, headless, under multiprocessing on this box's host cores (north_star's baseline;
    oracle/ref_baseline.py runs the compiled reference of oracle/_ref).  -> dict, or None when the reference is absent."""
    from oracle import ref_baseline as rb
    if not rb.available():
        return None
    per = 8
    n = min(len(arr), processes * per)
    maps = [arr[k, :lens[k]] for k in range(n)]
    rate, ticks, wall, procs, envs = rb.run(maps, processes, seconds, FRAME_SKIP, MAX_FRAMES, seed=seed, envs_per_process=per)
    return {"value": rate, "unit": "env-steps/s", "cores": procs, "kind": "reference", "per_core": rate / procs, "cpu": cpu_model(),
            "sample": f"nsim.Simulator (unmodified reference, byte-compiled by oracle/make_ref.py) x {envs} envs on the first {n} levels of the workload, "
                      f"{procs} processes x {seconds:.0f} s = {ticks} ticks ({wall:.1f} s wall incl. process start and level loading)"}


def config1_reference(reps):
    """BASELINE.json configs[0]: first bc_replay through the reference, 1 env, 1 core, load / tick timed separately."""
    from oracle import ref_baseline as rb
    if not rb.available():
        return None
    g = np.load(os.path.join(ROOT, "tests", "golden", "bc_replays.npz"))
    md = g["map_data"][g["map_off"][0]:g["map_off"][1]]
    i0, i1 = int(g["inp_off"][0]), int(g["inp_off"][1])
    r = rb.config1(md, g["hor"][i0:i1], g["jump"][i0:i1], reps=reps)
    r["replay"] = str(g["names"][0])
    r["known_answer_ok"] = bool(r["final_pos"] == [812.0036311681246, 542.0] and r["final_state"] == 8)   # BASELINE.json configs[0]
    return r


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores, same metric / config.
    Times the reference itself (oracle/_ref, kind "reference") when present, else its C restatement (kind "port").
    A "step" is a bounded time slice of the workload (the whole run is sized to about 2.5 minutes)."""
    if rank != 0:
        return
    fixture, n_envs, desc = WORKLOADS[args.workload]
    if args.envs:
        n_envs = args.envs
    arr, lens = load_maps(fixture)
    threads = usable_cores()
    n_slices = max(1, args.steps + args.warmup)
    from oracle import ref_baseline as rb
    use_ref = rb.available() and not args.port_baseline
    t0 = time.perf_counter()
    if use_ref:
        per_step = max(0.25, min(20.0, 150.0 / n_slices))
        per = 8
        n = min(len(arr), threads * per)
        slices, _, procs, envs = rb.run_slices([arr[k, :lens[k]] for k in range(n)], threads, per_step, n_slices, FRAME_SKIP, MAX_FRAMES, envs_per_process=per)
        vals = [v for v, _ in slices[args.warmup:]]
        info = {"unit": "env-steps/s", "cores": procs, "kind": "reference", "cpu": cpu_model(),
                "sample": f"nsim.Simulator (unmodified reference, byte-compiled by oracle/make_ref.py) x {envs} envs on the first {n} levels of the workload, "
                          f"{procs} processes; per step: a {per_step:.2f} s slice, {int(np.mean([t for _, t in slices[args.warmup:]]))} ticks"}
    else:
        per_step = max(0.5, min(20.0, 150.0 / n_slices))
        vals = []
        for s in range(n_slices):
            v, sample = cpu_baseline_port(arr, lens, threads, per_step, seed=s)
            if s >= args.warmup:
                vals.append(v)
        info = {"unit": "env-steps/s", "cores": threads, "kind": "port", "cpu": cpu_model(), "sample": "per step: " + sample}
    wall = time.perf_counter() - t0
    value = float(np.mean(vals))
    info["value"] = value; info["per_core"] = value / info["cores"]
    line = {
        "impl": "reference", "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * wall / n_slices,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": headline_config(args, len(arr)),
        "cpu_baseline": info,
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "host_threads": threads,
    }
    if use_ref:
        line["config1"] = config1_reference(1000)
    print(json.dumps(line))


def replay_parity(dev):
    """The second half of BASELINE.json's metric ("bit-exact replay %"), computed on this box: every bc_replays recording
    through nppsim_step_stream (one env per replay, raw input bytes, one tick per launch), the ninja's position / velocity /
    state compared bit for bit with the unmodified reference's after EVERY frame and the full canonical record (ninja +
    every entity) after the last one (tests/golden/bc_replays.npz, written by oracle/gen_golden.py)."""
    import torch
    from nclone_b200 import BatchedSimulator
    g = np.load(os.path.join(ROOT, "tests", "golden", "bc_replays.npz"))
    n = len(g["names"])
    maps = [g["map_data"][g["map_off"][i]:g["map_off"][i + 1]].astype(np.float64) for i in range(n)]
    seqs = []
    for i in range(n):
        h = g["hor"][g["inp_off"][i]:g["inp_off"][i + 1]].astype(np.int64); j = g["jump"][g["inp_off"][i]:g["inp_off"][i + 1]].astype(np.int64)
        seqs.append((j | ((h == 1) << 1) | ((h == -1) << 2)).astype(np.uint8))            # replay_executor.py:61-86
    lens = np.array([len(q) for q in seqs])
    sim = BatchedSimulator(n, device=dev)
    sim.load_maps(maps, np.arange(n))
    sp = sim.make_stream(seqs, "replay", 1)
    T = int(lens.max())
    pv = torch.empty((T, 4, n), dtype=torch.float64, device=dev); st = torch.empty((T, n), dtype=torch.int32, device=dev)
    for t in range(T):
        sim.step_stream(sp, 1, start_tick=t)
        pv[t].copy_(sim.nin_f[:4]); st[t].copy_(sim.nin_i[0])
    pv = pv.cpu().numpy(); st = st.cpu().numpy()
    recs = sim.canonical_state()
    ok, frames = 0, 0
    for i in range(n):
        f0 = int(g["frame_off"][i]); L = int(lens[i])
        same = np.array_equal(pv[:L, :, i].view(np.int64), g["pv"][f0 + 1:f0 + 1 + L].view(np.int64)) and np.array_equal(st[:L, i], g["state"][f0 + 1:f0 + 1 + L])
        e0, e1 = int(g["ent_off"][i]), int(g["ent_off"][i + 1])
        nf, ni, ef, ei = recs[i]
        same = same and np.array_equal(nf.view(np.int64), g["final_nin_f"][i].view(np.int64)) and np.array_equal(ni, g["final_nin_i"][i])
        same = same and np.array_equal(ef.view(np.int64), g["final_ent_f"][e0:e1].view(np.int64)) and np.array_equal(ei, g["final_ent_i"][e0:e1])
        ok += int(same); frames += L
    return {"replays": n, "bit_exact": ok, "frames": int(frames), "pct": 100.0 * ok / n,
            "check": "pos/vel/state every frame + full record at the end vs the unmodified reference"}


# ---------------------------------------------------------------------------------------------------- GPU measurement
def measure(workload, n_envs, args, rank, world, dev, dist, steps, warmup, spread, flush, do_async=False, do_obs=True, sampler=None):
    """Time one workload on this rank's GPU -> dict of this rank's numbers (reduced over ranks by reduce_block)."""
    import torch
    from nclone_b200 import BatchedSimulator
    from nclone_b200.sharding import shard_env_to_map
    arr, lens = load_maps(WORKLOADS[workload][0])
    M = len(arr)
    # shard: every rank owns its own env batch (weak scaling); ranks use different level rotations and seeds
    env_to_map = shard_env_to_map(n_envs, rank, M, adjacent=(args.placement == "adjacent"))
    sim = BatchedSimulator(n_envs, device=dev, handoff=None if args.handoff == "auto" else (args.handoff == "on"), placement=args.engine_order)
    sim.load_maps([arr[k, :lens[k]] for k in range(M)], env_to_map)
    gen = torch.Generator(device=dev); gen.manual_seed(1234 + rank)
    POOL = 64
    actions = torch.randint(0, 6, (POOL, n_envs), generator=gen, device=dev, dtype=torch.uint8)
    actions_host = actions.cpu().pin_memory()

    def one_step(s):
        sim.step(actions[s % POOL], FRAME_SKIP, True)
        sim.reset_terminal(MAX_FRAMES)

    for s in range(spread + warmup):             # `spread` steps put the envs at mixed episode phases (about 1.5 mean episodes)
        one_step(s)
    torch.cuda.synchronize(dev)
    if sampler is not None:
        sampler.start()
    if world > 1:
        dist.barrier()
        for s in range(3):                       # the first collective's lazy set-up and the sampler's start stay outside the timed window
            one_step(spread + warmup + s)
    torch.cuda.synchronize(dev)
    ticks0, launches0, handoffs0 = sim.total_ticks(), sim.launch_count(), sim.handoff_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for s in range(steps):
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
    out = {"n_envs": n_envs, "M": M, "arr": arr, "lens": lens, "env_to_map": env_to_map, "sim": sim}
    out["launches"] = sim.launch_count() - launches0
    step_ms = np.array([e[0].elapsed_time(e[1]) for e in ev]); reset_ms = np.array([e[1].elapsed_time(e[2]) for e in ev])
    out["step_ms"], out["reset_ms"] = step_ms, reset_ms
    out["elapsed_ms"] = float(step_ms.sum() + reset_ms.sum())
    out["ticks"] = sim.total_ticks() - ticks0
    out["handoff"] = {"enabled": sim.handoff is not None, "mid_launch_per_step": (sim.handoff_count() - handoffs0) / steps,
                      "solo_start_queue": sim.solo_queue_length(), "solo_sms": sim.solo_sms(), "status": sim.handoff_status(),
                      "policy": args.handoff, "saturated_launches": list(sim.handoff_saturation())}
    out["engine_order"] = "envs of a level in consecutive state columns" if sim.slot_env is not None else "caller's order"

    # ---- e2e: host buffers through the C ABI, copies inside the timed region -------------------------
    ah = actions_host.numpy()

    def e2e_leg(with_obs):
        block, _ = sim.host_block(with_obs)
        for s in range(3):
            sim.step_host(ah[s % POOL], block, FRAME_SKIP, True, with_obs, reset_max_frames=MAX_FRAMES)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        t1 = sim.total_ticks()
        # timed per step like the device-resident loop, so that the same untimed L2 flush can sit between the steps
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for s in range(steps):
            if flush is not None:
                flush.zero_()
            evs[s][0].record()
            sim.step_host(ah[s % POOL], block, FRAME_SKIP, True, with_obs, reset_max_frames=MAX_FRAMES)     # nppsim_step_host_reset: step + episode turnover, one call
            evs[s][1].record()
        torch.cuda.synchronize(dev)
        return float(sum(a.elapsed_time(b) for a, b in evs)), sim.total_ticks() - t1, int(block.numel())

    out["e2e_ms"], out["e2e_ticks"], out["e2e_d2h"] = e2e_leg(False)
    if do_obs:
        out["e2e_obs_ms"], out["e2e_obs_ticks"], out["e2e_obs_d2h"] = e2e_leg(True)

    # ---- the same batch as independent groups on separate streams (BatchedSimulator.groups): what an asynchronous RL loop
    # does -- every group's launches in stream order, no per-step synchronisation of the whole batch, so one group's
    # long-tailed step kernel overlaps the other groups' kernels.  Reported beside `value`, never instead of it.
    out["grp_ms"], out["grp_ticks"] = 0.0, 0
    if do_async and args.async_groups > 1:
        groups = sim.groups(args.async_groups)
        gstreams = [torch.cuda.Stream(device=dev) for _ in groups]
        ga = [actions[:, torch.from_numpy(np.asarray(g.env_ids, dtype=np.int64)).to(dev)].contiguous() for g in groups]   # view env k = parent env env_ids[k]
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        t2 = sim.total_ticks()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for gs in gstreams:
            gs.wait_event(g0)
        for s in range(steps):
            for k, g in enumerate(groups):
                with torch.cuda.stream(gstreams[k]):
                    g.step(ga[k][s % POOL], FRAME_SKIP, True)
                    g.reset_terminal(MAX_FRAMES)
        for gs in gstreams:
            torch.cuda.current_stream(dev).wait_stream(gs)
        g1.record()
        torch.cuda.synchronize(dev)
        out["grp_ms"] = g0.elapsed_time(g1); out["grp_ticks"] = sim.total_ticks() - t2
    return out


def reduce_block(m, steps, dist, world, dev):
    """Max over ranks of the elapsed times, sum over ranks of the work: whole-job numbers of one workload.
    The only collectives of the bench (NCCL), outside the timed regions."""
    import torch
    from nclone_b200.sharding import reduce_rollout
    times = [m["elapsed_ms"], m["e2e_ms"], m.get("e2e_obs_ms", 0.0), m["grp_ms"]]
    work = [m["ticks"], m["e2e_ticks"], m.get("e2e_obs_ticks", 0), m["grp_ticks"]]
    (el, e2e, e2eo, grp), (tk, e2etk, e2eotk, grptk), stats = reduce_rollout(times, work, m["sim"].stats, dist if world > 1 else None)
    r = {"value": tk / (el / 1e3), "ms_per_step": el / steps, "e2e_value": e2etk / (e2e / 1e3), "e2e_ms_per_step": e2e / steps, "stats": stats}
    if e2eo > 0:
        r["e2e_obs_value"] = e2eotk / (e2eo / 1e3); r["e2e_obs_ms_per_step"] = e2eo / steps
    if grp > 0:
        r["grp_value"] = grptk / (grp / 1e3); r["grp_ms_per_step"] = grp / steps
    # attribution of the max-over-ranks time: every rank's step-kernel ms (min / median / max / mean) and wedge handoffs
    if world > 1:
        sm = m["step_ms"]
        mine = torch.tensor([float(sm.min()), float(np.median(sm)), float(sm.max()), float(sm.mean()), m["e2e_ms"] / steps,
                             m["handoff"]["mid_launch_per_step"], float(m["handoff"]["solo_start_queue"])], dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        keys = ("step_ms_min", "step_ms_median", "step_ms_max", "step_ms_mean", "e2e_ms_per_step", "wedge_handoffs_per_step", "wedged_envs_at_launch")
        r["per_rank"] = [{k: round(float(v), 4) for k, v in zip(keys, t.tolist())} for t in allr]
        slow = int(np.argmax([p["step_ms_mean"] for p in r["per_rank"]]))
        r["slowest_rank"] = dict(r["per_rank"][slow], rank=slow)
    return r


def secondary_block(tag, m, r, n_envs, steps):
    b = {"workload": WORKLOADS[tag][2], "envs_per_gpu": n_envs, "steps": steps, "value": r["value"], "unit": "env-steps/s", "ms_per_step": r["ms_per_step"],
         "kernel_ms": {"step_kernel": float(m["step_ms"].mean()), "reset_kernel": float(m["reset_ms"].mean())},
         "e2e": {"value": r["e2e_value"], "unit": "env-steps/s", "ms_per_step": r["e2e_ms_per_step"], "h2d_bytes_per_step": int(n_envs), "d2h_bytes_per_step": int(m["e2e_d2h"])},
         "handoff": m["handoff"], "kernel": m["sim"].kernel_variant(), "engine_order": m["engine_order"]}
    for k in ("per_rank", "slowest_rank"):
        if k in r:
            b[k] = r[k]
    return b


def build_parser():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="override envs per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--port-baseline", action="store_true", help="--impl reference: time the C restatement even when the compiled reference is present")
    ap.add_argument("--no-secondary", action="store_true", help="skip the config2 / config4 / full_kernel blocks")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between timed steps")
    ap.add_argument("--async-groups", type=int, default=4, help="also time the batch stepped as this many independent groups on separate streams (0: skip)")
    ap.add_argument("--placement", default="scattered", choices=["scattered", "adjacent"], help="caller's env -> level assignment: replicas of a level n_levels envs apart, or consecutive")
    ap.add_argument("--engine-order", default="auto", choices=["level", "caller", "auto"], help="BatchedSimulator placement: envs of a level in consecutive state columns, the caller's order, or (default) by replicas per level")
    ap.add_argument("--handoff", default="auto", choices=["auto", "on", "off"], help="solo-warp handoff of wedged envs")
    ap.add_argument("--spread", type=int, default=300, help="untimed agent steps run first so envs sit at mixed episode phases")
    return ap


def main():
    args = build_parser().parse_args()
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

    n_envs = args.envs or WORKLOADS[args.workload][1]
    flush = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
    sampler = ClockSampler(torch.cuda.current_device() if "CUDA_VISIBLE_DEVICES" not in os.environ else local_rank) if rank == 0 else None
    m = measure(args.workload, n_envs, args, rank, world, dev, dist, args.steps, args.warmup, args.spread, flush, do_async=True, sampler=sampler)
    clocks = sampler.finish() if sampler is not None else None
    r = reduce_block(m, args.steps, dist, world, dev)

    # ---- secondary blocks: the other BASELINE.json configs, short runs in the same process ---------------------------------
    secondary = {}
    if not args.no_secondary and args.workload == "c3" and not args.envs:
        s_steps = max(10, min(40, args.steps))
        for tag, key in (("c2", "config2"), ("c4", "config4"), ("c5", "full_kernel")):
            ne = WORKLOADS[tag][1]
            m2 = measure(tag, ne, args, rank, world, dev, dist, s_steps, 3, 150, flush, do_async=False, do_obs=False)
            r2 = reduce_block(m2, s_steps, dist, world, dev)
            secondary[key] = secondary_block(tag, m2, r2, ne, s_steps)
            del m2, r2
            torch.cuda.empty_cache()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    sim = m["sim"]
    value, e2e_value, stats = r["value"], r["e2e_value"], r["stats"]
    arr, lens, env_to_map, M = m["arr"], m["lens"], m["env_to_map"], m["M"]
    step_ms, reset_ms, ticks = m["step_ms"], m["reset_ms"], m["ticks"]

    # ---- roofline of the dominant kernel (step_kernel) ---------------------------------------------------------------------
    # Algorithmic bytes per env-step.  SURVEY.md 8(d)'s per-tick figure is 892 + 19 E + sum_dyn(17 + 2 state) for ONE tick per
    # launch.  The kernel fuses K = frame_skip ticks per launch, so the per-env state (281 B ninja + entity state, read + written)
    # moves once per launch, not once per tick: fused = (562 + 2 sum_state) / K' + 330 + 19 E + 17 n_dyn, with K' = ticks actually
    # executed per env per launch (< 4 with early exits).  Both are reported; `achieved` uses the fused form.
    e_static, dyn, dyn_state, n_dyn = entity_stats(arr, lens)
    ticks_per_launch = ticks / args.steps
    k_eff = max(1e-9, ticks_per_launch / n_envs)
    unfused = float(np.mean((892 + 19 * e_static + dyn)[env_to_map]))
    fused = float(np.mean(((562 + 2 * dyn_state) / k_eff + 330 + 19 * e_static + 17 * n_dyn)[env_to_map]))
    step_s = float(step_ms.mean()) / 1000.0
    achieved = fused * ticks_per_launch / step_s / 1e9
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback"
    traffic = warp_inst = None
    tpath = os.path.join(ROOT, "profiles", "step_kernel_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if tj.get("workload") == args.workload and tj.get("envs") == n_envs:
            traffic = tj.get("dram_bytes_per_launch"); warp_inst = tj.get("warp_inst_per_launch")
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "kernel": "step_kernel", "kernel_ms": float(step_ms.mean()), "algorithmic_bytes_per_env_step": fused,
                "algorithmic_bytes_per_env_step_unfused": unfused, "achieved_unfused": unfused * ticks_per_launch / step_s / 1e9,
                "env_steps_per_launch": ticks_per_launch, "peak_source": peak_src,
                "note": "branchy fp64 scalar code bound by per-warp latency and issue slots, not HBM (DESIGN.md 3): the HBM fraction is reported as the "
                        "contract requires, `secondary` is the ceiling that binds"}
    cpath = os.path.join(ROOT, "profiles", "compute_peaks.json")
    if os.path.exists(cpath) and warp_inst:
        cp = json.load(open(cpath))
        ips = warp_inst / step_s
        roofline["secondary"] = {
            "bound": "issue", "warp_inst_per_launch": warp_inst, "achieved": ips, "unit": "warp-inst/s",
            "peak": cp["issue_slots_per_s_nominal"], "frac_issue": ips / cp["issue_slots_per_s_nominal"],
            "peak_fp64_or_imad_pipe": cp["dfma_warp_inst_per_s"], "frac_of_fp64_pipe_rate": ips / cp["dfma_warp_inst_per_s"],
            "peak_source": "profiles/compute_peaks.json: 4 schedulers x 148 SMs x max SM clock (nominal issue slots); DFMA and IMAD each measured at 1.96 "
                           "warp-inst/cycle/SM on this pool's B200 by tools/microbench/fp64_issue.cu -- a kernel made only of those would top out at frac_issue 0.49",
            "dfma_tflops_measured": cp.get("dfma_tflops"),
            "warp_inst_source": "ncu smsp__inst_executed.sum of the same workload (profiles/step_kernel_traffic.json)"}

    cpu = cpu_port = None
    if not args.no_cpu_baseline and world == 1:          # rank 0 at N=1 only
        threads = usable_cores()
        cpu = cpu_baseline_reference(arr, lens, threads, 20.0)
        v, sample = cpu_baseline_port(arr, lens, threads, 8.0 if cpu else 25.0)
        cpu_port = {"value": v, "unit": "env-steps/s", "cores": threads, "kind": "port", "sample": sample}
        if cpu is None:
            cpu, cpu_port = cpu_port, None
        else:
            cpu["config1"] = config1_reference(100)

    line = {
        "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": headline_config(args, M),
        "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": int(n_envs), "d2h_bytes_per_step": int(m["e2e_d2h"]),
                "ms_per_step": r["e2e_ms_per_step"]},
        "gpu_launches": int(m["launches"]),
        "kernel_ms": {"step_kernel": float(step_ms.mean()), "reset_kernel": float(reset_ms.mean())},
        "agent_steps_per_s": value / FRAME_SKIP,
        "episode_stats": {"episodes": int(stats[0]), "wins": int(stats[1]), "deaths": int(stats[2]), "truncations": int(stats[3])},
        "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks, "handoff": m["handoff"],
        "engine_order": m["engine_order"], "kernel": sim.kernel_variant(),
    }
    if cpu_port is not None:
        line["cpu_baseline_port"] = cpu_port
    if "e2e_obs_value" in r:
        line["e2e_obs"] = {"value": r["e2e_obs_value"], "unit": "env-steps/s", "ms_per_step": r["e2e_obs_ms_per_step"], "h2d_bytes_per_step": int(n_envs),
                           "d2h_bytes_per_step": int(m["e2e_obs_d2h"]),
                           "note": "nppsim_step_host(with_obs=1): the same call also returns get_ninja_state (40 x f32) and the action mask (6 x u8) per env"}
    if "grp_value" in r:
        line["async_groups"] = {"groups": args.async_groups, "value": r["grp_value"], "unit": "env-steps/s", "ms_per_step": r["grp_ms_per_step"],
                                "note": "same batch, same work, stepped as independent groups on separate streams (no L2 flush, no per-step sync)"}
    line.update(secondary)
    for k in ("per_rank", "slowest_rank"):       # (last: the driver's record keeps the tail of the line)
        if k in r:
            line[k] = r[k]
    if not args.no_secondary:
        line["replay_parity"] = replay_parity(dev)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
