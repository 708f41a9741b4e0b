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
import importlib.abc
import importlib.util
import marshal
import multiprocessing as mp
import os
import random
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ACTION_TO_INPUTS = ((0, 0), (-1, 0), (1, 0), (0, 1), (-1, 1), (1, 1))


def _stale(compiled, src):
    """True when `src` holds the reference and some module listed in the compiled copy's MANIFEST is missing from it or
    newer there than its compiled image."""
    if not os.path.isfile(os.path.join(src, "nclone", "nsim.py")):
        return False
    try:
        with open(os.path.join(compiled, "MANIFEST.txt")) as f:
            mods = [ln.strip() for ln in f.read().splitlines()[1:] if ln.strip()]
    except OSError:
        return True
    for name in mods:
        rel = name.replace(".", os.sep)
        for cand in (rel + ".py", os.path.join(rel, "__init__.py")):
            if os.path.isfile(os.path.join(src, cand)):
                img = os.path.join(compiled, cand[:-3] + ".refc")
                if not os.path.isfile(img) or os.path.getmtime(img) < os.path.getmtime(os.path.join(src, cand)):
                    return True
                break
        else:
            return True
    return False


def ref_path():
    """Where the reference's `nclone` package can be imported from, or None."""
    compiled = os.path.join(HERE, "_ref")
    src = os.environ.get("NCLONE_REF", "/root/reference")
    if os.path.isfile(os.path.join(compiled, "nclone", "nsim.refc")):
        # where the reference tree is present too (the build container) a compiled copy older than any of its sources
        # would be timed silently in its place: refuse it and import the tree itself (python -m oracle.make_ref rebuilds)
        if not _stale(compiled, src):
            return compiled
    if os.path.isfile(os.path.join(src, "nclone", "nsim.py")):
        return src
    return None


def available():
    return ref_path() is not None


class _RefLoader(importlib.abc.Loader):
    def __init__(self, path):
        self.path = path

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        with open(self.path, "rb") as f:
            code = marshal.loads(f.read()[16:])          # .pyc image: 16-byte header, then the marshalled code object
        module.__file__ = self.path                       # constants/physics_constants.py:243 builds a path from it (animation data: unused headless)
        exec(code, module.__dict__)


class _RefFinder(importlib.abc.MetaPathFinder):
    """Imports `nclone.*` from the compiled modules of oracle/_ref (make_ref.py writes .pyc images as *.refc)."""

    def __init__(self, root):
        self.root = root

    def find_spec(self, name, path=None, target=None):
        if name.split(".")[0] != "nclone":
            return None
        base = os.path.join(self.root, *name.split("."))
        if os.path.isfile(os.path.join(base, "__init__.refc")):
            spec = importlib.util.spec_from_loader(name, _RefLoader(os.path.join(base, "__init__.refc")), origin=os.path.join(base, "__init__.refc"), is_package=True)
            spec.submodule_search_locations = [base]
            return spec
        if os.path.isfile(base + ".refc"):
            return importlib.util.spec_from_loader(name, _RefLoader(base + ".refc"), origin=base + ".refc")
        if os.path.isdir(base):                                  # a package without __init__ (namespace)
            spec = importlib.util.spec_from_loader(name, None, is_package=True)
            spec.submodule_search_locations = [base]
            return spec
        return None


def _import():
    p = ref_path()
    if p is None:
        raise RuntimeError("reference not available: neither oracle/_ref nor /root/reference")
    scratch = os.environ.get("NCLONE_REF_HOME", "/tmp/nclone_ref_home")      # persistent_collision_cache.py:47-50 writes ~/.cache
    os.makedirs(scratch, exist_ok=True)
    os.environ["HOME"] = scratch
    sys.dont_write_bytecode = True
    if os.path.isfile(os.path.join(p, "nclone", "nsim.refc")):
        if not any(isinstance(f, _RefFinder) for f in sys.meta_path):
            sys.meta_path.insert(0, _RefFinder(p))
    elif p not in sys.path:
        sys.path.insert(0, p)
    from nclone.nsim import Simulator
    from nclone.sim_config import SimConfig
    return Simulator, SimConfig


def _as_list(md):
    # map_data the way the reference's loaders hand it over: ints where integral (entity types index lists)
    return [int(v) if float(v).is_integer() else float(v) for v in md]


def _worker(args):
    maps, seed, frame_skip, max_frames, seconds, n_slices = args
    Simulator, SimConfig = _import()
    sims = []
    for md in maps:
        s = Simulator(SimConfig(enable_anim=False, log_data=False, debug=False))
        s.load(_as_list(md))
        sims.append(s)
    rng = random.Random(seed)
    out = []
    for _ in range(n_slices):
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
        out.append((ticks, time.perf_counter() - t0))
    return out


def run_slices(maps, processes, seconds, n_slices, frame_skip=4, max_frames=2000, seed=0, envs_per_process=8):
    """One worker pool; every worker loads `envs_per_process` of `maps` (round-robin) once and then runs `n_slices` timed
    slices of `seconds` each (the envs keep their episode state across slices).
    -> (list over slices of (ticks/s summed over workers, ticks), wall seconds, processes, envs)."""
    processes = max(1, int(processes))
    shares = [[maps[(p * envs_per_process + k) % len(maps)] for k in range(envs_per_process)] for p in range(processes)]
    jobs = [(shares[p], seed * 1000 + p, frame_skip, max_frames, seconds, n_slices) for p in range(processes)]
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(processes) as pool:
        res = pool.map(_worker, jobs)
    wall = time.perf_counter() - t0
    slices = []
    for k in range(n_slices):
        slices.append((sum(r[k][0] / r[k][1] for r in res), sum(r[k][0] for r in res)))
    return slices, wall, processes, processes * envs_per_process


def run(maps, processes, seconds, frame_skip=4, max_frames=2000, seed=0, envs_per_process=8):
    """One slice: -> (ticks/s over all processes, ticks, wall seconds, processes used, envs used)."""
    slices, wall, procs, envs = run_slices(maps, processes, seconds, 1, frame_skip, max_frames, seed, envs_per_process)
    return slices[0][0], slices[0][1], wall, procs, envs


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
