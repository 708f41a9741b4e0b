"""ctypes wrapper around ``oracle/_build/liboracle.so`` (nsim_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of nsim_oracle.c.  Importers:
tests/, __graft_entry__.smoke(), bench.py (cpu_baseline / --impl reference).
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import schema

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build(force=False):
    src = [os.path.join(_HERE, f) for f in ("nsim_oracle.c", "baseline_runner.c", "Makefile")]
    if (not force and os.path.exists(_LIB_PATH)
            and all(os.path.getmtime(_LIB_PATH) >= os.path.getmtime(s) for s in src)):
        return _LIB_PATH
    env = dict(os.environ)
    env.setdefault("CC", "gcc")
    subprocess.run(["make", "-C", _HERE, "-B", "_build/liboracle.so"], check=True,
                   stdout=subprocess.DEVNULL, env=env)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        dp, ip, i8p = C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_int8)
        i32p, u8p = C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
        L.osim_create.restype = C.c_void_p
        L.osim_destroy.argtypes = [C.c_void_p]
        L.osim_load.argtypes = [C.c_void_p, dp, C.c_int]; L.osim_load.restype = C.c_int
        L.osim_load_ex.argtypes = [C.c_void_p, dp, C.c_int, C.c_int]; L.osim_load_ex.restype = C.c_int
        L.osim_reset.argtypes = [C.c_void_p]; L.osim_reset.restype = C.c_int
        L.osim_fast_reset.argtypes = [C.c_void_p]; L.osim_fast_reset.restype = C.c_int
        L.osim_tick.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.osim_apply_curriculum.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double]
        L.osim_apply_curriculum.restype = C.c_int
        L.osim_num_entities.argtypes = [C.c_void_p]; L.osim_num_entities.restype = C.c_int
        L.osim_num_segments.argtypes = [C.c_void_p]; L.osim_num_segments.restype = C.c_int
        L.osim_events.argtypes = [C.c_void_p]; L.osim_events.restype = C.c_int
        L.osim_get_state.argtypes = [C.c_void_p, dp, ip, dp, ip]
        L.osim_run.argtypes = [C.c_void_p, i8p, i8p, C.c_int, dp, ip, dp, ip]
        L.osim_get_cell_counts.argtypes = [C.c_void_p, i32p, i32p]
        L.osim_get_segments.argtypes = [C.c_void_p, dp]
        L.osim_get_grid_edges.argtypes = [C.c_void_p, i32p, i32p]
        L.osim_ninja_state.argtypes = [C.c_void_p, dp]
        L.osim_action_mask.argtypes = [C.c_void_p, C.c_int, u8p]
        L.osim_baseline_run.argtypes = [dp, i32p, C.c_int, C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_int, dp, dp]
        L.osim_baseline_run.restype = C.c_long
        L.osim_rollout_records.argtypes = [dp, i32p, C.c_int, i32p, C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, dp, ip]
        L.osim_rollout_records.restype = C.c_long
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class OracleSim:
    """Single-environment mirror of nsim.Simulator (load / reset / fast_reset / tick)."""

    def __init__(self):
        self.L = lib()
        self.h = C.c_void_p(self.L.osim_create())

    def __del__(self):
        try:
            self.L.osim_destroy(self.h)
        except Exception:
            pass

    def load(self, map_data, keep_counts=False):
        md = np.ascontiguousarray(np.asarray(map_data, dtype=np.float64))
        rc = self.L.osim_load_ex(self.h, _p(md, C.c_double), len(md), int(keep_counts))
        if rc:
            raise ValueError(f"oracle load failed rc={rc}")

    def reset(self):
        rc = self.L.osim_reset(self.h)
        if rc:
            raise ValueError(f"oracle reset failed rc={rc}")

    def fast_reset(self):
        self.L.osim_fast_reset(self.h)

    def apply_curriculum(self, switch_xy, door_xy):
        """IntermediateGoalManager.apply_to_simulator (intermediate_goal_manager.py:698-960)."""
        if self.L.osim_apply_curriculum(self.h, float(switch_xy[0]), float(switch_xy[1]), float(door_xy[0]), float(door_xy[1])):
            raise RuntimeError("No EntityExitSwitch found in entity_dic[3]")

    def tick(self, hor, jump):
        self.L.osim_tick(self.h, int(hor), int(jump))

    @property
    def num_entities(self):
        return self.L.osim_num_entities(self.h)

    def events(self):
        return self.L.osim_events(self.h)

    def ninja_state(self):
        """NPlayHeadless.get_ninja_state: 40 floats."""
        o = np.zeros(40)
        self.L.osim_ninja_state(self.h, _p(o, C.c_double))
        return o

    def action_mask(self, path_dir=0):
        """Ninja.get_valid_action_mask; path_dir 0 None / 1 left / 2 right / 3 down."""
        m = np.zeros(6, dtype=np.uint8)
        self.L.osim_action_mask(self.h, int(path_dir), _p(m, C.c_uint8))
        return m

    def state(self):
        E = self.num_entities
        nf = np.zeros(schema.NF); ni = np.zeros(schema.NI, dtype=np.int64)
        ef = np.zeros((E, schema.EF)); ei = np.zeros((E, schema.EI), dtype=np.int64)
        self.L.osim_get_state(self.h, _p(nf, C.c_double), _p(ni, C.c_int64), _p(ef, C.c_double), _p(ei, C.c_int64))
        return nf, ni, ef, ei

    def run(self, hor, jump, entities=True):
        """T ticks; per-frame records for frames 1..T."""
        hor = np.ascontiguousarray(hor, dtype=np.int8); jump = np.ascontiguousarray(jump, dtype=np.int8)
        T = len(hor); E = self.num_entities
        nf = np.zeros((T, schema.NF)); ni = np.zeros((T, schema.NI), dtype=np.int64)
        if entities:
            ef = np.zeros((T, E, schema.EF)); ei = np.zeros((T, E, schema.EI), dtype=np.int64)
            self.L.osim_run(self.h, _p(hor, C.c_int8), _p(jump, C.c_int8), T, _p(nf, C.c_double), _p(ni, C.c_int64),
                            _p(ef, C.c_double), _p(ei, C.c_int64))
            return nf, ni, ef, ei
        self.L.osim_run(self.h, _p(hor, C.c_int8), _p(jump, C.c_int8), T, _p(nf, C.c_double), _p(ni, C.c_int64), None, None)
        return nf, ni, None, None

    def static_tables(self):
        counts = np.zeros(44 * 25, dtype=np.int32); present = np.zeros(44 * 25, dtype=np.int32)
        self.L.osim_get_cell_counts(self.h, _p(counts, C.c_int32), _p(present, C.c_int32))
        segs = np.zeros((int(counts.sum()), 8))
        self.L.osim_get_segments(self.h, _p(segs, C.c_double))
        hor = np.zeros(89 * 51, dtype=np.int32); ver = np.zeros(89 * 51, dtype=np.int32)
        self.L.osim_get_grid_edges(self.h, _p(hor, C.c_int32), _p(ver, C.c_int32))
        return counts, present, segs, hor, ver


def run_trajectory(map_data, hor, jump, moves=()):
    """Same return shape as refdump.run_trajectory (frame 0 = after load).  moves: (t, sx, sy, dx, dy) tuples,
    apply_curriculum before tick t (t = 0: right after load, i.e. before the frame-0 record)."""
    s = OracleSim()
    s.load(map_data)
    T = len(hor)
    cuts = sorted(set(int(m[0]) for m in moves if 0 < m[0] < T))
    for m in moves:
        if m[0] == 0:
            s.apply_curriculum(m[1:3], m[3:5])
    recs = [tuple(a[None] for a in s.state())]
    t0 = 0
    for t1 in cuts + [T]:
        if t1 > t0:
            recs.append(s.run(hor[t0:t1], jump[t0:t1]))
        for m in moves:
            if m[0] == t1 and t1 < T:
                s.apply_curriculum(m[1:3], m[3:5])
        t0 = t1
    NF, NI, EF, EI = (np.concatenate([r[k] for r in recs]) for k in range(4))
    H = np.array([schema.state_hash(NF[t], NI[t], EF[t], EI[t]) for t in range(T + 1)], dtype=np.uint64)
    return {"nin_f": NF, "nin_i": NI, "ent_f": EF, "ent_i": EI, "hash": H}


def baseline_run(maps, lens, actions, frame_skip, max_frames, n_threads):
    """Multi-threaded CPU baseline (baseline_runner.c). maps [M,stride] f64, actions [M,S] u8.
    Returns (ticks executed, slowest worker's seconds inside its tick loops)."""
    maps = np.ascontiguousarray(maps, dtype=np.float64); lens = np.ascontiguousarray(lens, dtype=np.int32)
    actions = np.ascontiguousarray(actions, dtype=np.uint8)
    cs = C.c_double(0); secs = C.c_double(0)
    ticks = lib().osim_baseline_run(_p(maps, C.c_double), _p(lens, C.c_int32), maps.shape[0], maps.shape[1],
                                    _p(actions, C.c_uint8), actions.shape[1], frame_skip, max_frames, n_threads,
                                    C.byref(cs), C.byref(secs))
    return int(ticks), secs.value


def rollout_records(maps, lens, env_map, actions, frame_skip, max_frames, every, n_threads):
    """Multi-threaded oracle rollout with checkpoints (baseline_runner.c: osim_rollout_records).  maps [M,stride] f64,
    env_map [N], actions [S,N] u8 (step-major, as the GPU takes them) -> (ticks, nin_f [S//every, N, 31], nin_i [S//every, N, 27])
    taken after every `every`-th agent step, before that step's turnover (fresh load on win / death / max_frames)."""
    maps = np.ascontiguousarray(maps, dtype=np.float64); lens = np.ascontiguousarray(lens, dtype=np.int32)
    env_map = np.ascontiguousarray(env_map, dtype=np.int32); actions = np.ascontiguousarray(actions, dtype=np.uint8)
    S, N = actions.shape
    out_f = np.zeros((S // every, N, 31)); out_i = np.zeros((S // every, N, 27), dtype=np.int64)
    ticks = lib().osim_rollout_records(_p(maps, C.c_double), _p(lens, C.c_int32), maps.shape[1], _p(env_map, C.c_int32), N,
                                       _p(actions, C.c_uint8), S, frame_skip, max_frames, every, n_threads,
                                       _p(out_f, C.c_double), _p(out_i, C.c_int64))
    return int(ticks), out_f, out_i
