"""CPU baseline = the reference's own `nsim.Simulator`, headless, under multiprocessing (BASELINE.json north_star;
SURVEY.md 8(d) "CPU baseline beside it"; the reference's own multi-process runner is run_multiple_headless.py:128-137).

TEST / MEASUREMENT INFRASTRUCTURE: imported only by bench.py (`--impl reference`, `cpu_baseline`).  Runs the compiled
reference modules of oracle/_ref (oracle/make_ref.py) -- or the source tree when /root/reference is present -- through
the reference's public API: Simulator(SimConfig(enable_anim=False, log_data=False, debug=False)) (what NPlayHeadless builds,
nplay_headless.py:63-67), sim.load(map_data), sim.tick(hor, jump), sim.reset().

Policy = bench.py's: uniform Discrete(6) actions (ACTION_TO_INPUTS, action_replayer.py:30-37) held for frame_skip ticks
with the env wrapper's early exit on win / death (base_environment.py:605-609), Simulator.reset() when the episode ended
or reached max_frames.  Each worker process owns the Simulators of its share of the levels.
"""
import multiprocessing as mp
import os
import random
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ACTION_TO_INPUTS = ((0, 0), (-1, 0), (1, 0), (0, 1), (-1, 1), (1, 1))


def ref_path():
    """Where the reference's `nclone` package can be imported from, or None."""
    compiled = os.path.join(HERE, "_ref")
    if os.path.isfile(os.path.join(compiled, "nclone", "nsim.pyc")):
        return compiled
    src = os.environ.get("NCLONE_REF", "/root/reference")
    if os.path.isfile(os.path.join(src, "nclone", "nsim.py")):
        return src
    return None


def available():
    return ref_path() is not None


def _import():
    p = ref_path()
    if p is None:
        raise RuntimeError("reference not available: neither oracle/_ref nor /root/reference")
    scratch = os.environ.get("NCLONE_REF_HOME", "/tmp/nclone_ref_home")      # persistent_collision_cache.py:47-50 writes ~/.cache
    os.makedirs(scratch, exist_ok=True)
    os.environ["HOME"] = scratch
    sys.dont_write_bytecode = True
    if p not in sys.path:
        sys.path.insert(0, p)
    from nclone.nsim import Simulator
    from nclone.sim_config import SimConfig
    return Simulator, SimConfig


def _as_list(md):
    # map_data the way the reference's loaders hand it over: ints where integral (entity types index lists)
    return [int(v) if float(v).is_integer() else float(v) for v in md]


def _worker(args):
    maps, seed, frame_skip, max_frames, seconds = args
    Simulator, SimConfig = _import()
    sims = []
    for md in maps:
        s = Simulator(SimConfig(enable_anim=False, log_data=False, debug=False))
        s.load(_as_list(md))
        sims.append(s)
    rng = random.Random(seed)
    ticks = 0
    t0 = time.perf_counter()
    deadline = t0 + seconds
    while time.perf_counter() < deadline:
        for s in sims:
            h, j = ACTION_TO_INPUTS[rng.randrange(6)]
            for _ in range(frame_skip):
                s.tick(h, j)
                ticks += 1
                if s.ninja.state in (6, 7, 8):
                    break
            if s.ninja.state in (6, 7, 8) or s.frame >= max_frames:
                s.reset()
    return ticks, time.perf_counter() - t0


def run(maps, processes, seconds, frame_skip=4, max_frames=2000, seed=0, envs_per_process=8):
    """-> (ticks/s over all processes, ticks, wall seconds, processes used, envs used).  `maps`: list of map_data rows;
    every process gets `envs_per_process` of them (round-robin)."""
    processes = max(1, int(processes))
    shares = [[maps[(p * envs_per_process + k) % len(maps)] for k in range(envs_per_process)] for p in range(processes)]
    jobs = [(shares[p], seed * 1000 + p, frame_skip, max_frames, seconds) for p in range(processes)]
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(processes) as pool:
        res = pool.map(_worker, jobs)
    wall = time.perf_counter() - t0
    ticks = sum(r[0] for r in res)
    busy = max(r[1] for r in res)          # the measured region of the slowest worker (startup / map loading excluded)
    return ticks / busy, ticks, wall, processes, processes * envs_per_process


def config1(map_data, hor, jump, reps=1000):
    """BASELINE.json configs[0]: the first bc_replay through the reference, 1 env, 1 core; `load` and `tick` timed
    separately over `reps` repetitions.  -> dict."""
    Simulator, SimConfig = _import()
    md = _as_list(map_data)
    s = Simulator(SimConfig(enable_anim=False, log_data=False, debug=False))
    t_load = t_tick = 0.0
    for _ in range(reps):
        t0 = time.perf_counter()
        s.load(list(md))
        t1 = time.perf_counter()
        for h, j in zip(hor, jump):
            s.tick(int(h), int(j))
        t2 = time.perf_counter()
        t_load += t1 - t0; t_tick += t2 - t1
    return {"reps": reps, "frames": len(hor), "load_ms": 1e3 * t_load / reps, "ticks_per_s": reps * len(hor) / t_tick,
            "final_pos": [s.ninja.xpos, s.ninja.ypos], "final_state": int(s.ninja.state)}
