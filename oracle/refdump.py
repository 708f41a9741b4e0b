"""Run the UNMODIFIED reference (imported read-only from /root/reference) and
extract the canonical state record of ``oracle/schema.py``.

TEST INFRASTRUCTURE ONLY, and usable only in the build container: the GPU box
has no /root/reference.  Used by ``oracle/gen_golden.py`` (fixture generation)
and by ad-hoc differential fuzzing of ``nsim_oracle.c`` (``oracle/fuzz_vs_ref.py``).
"""
import os
import sys

import numpy as np

from . import schema

REF_ROOT = os.environ.get("NCLONE_REF", "/root/reference")


def import_reference():
    """HOME must be writable because
    utils/persistent_collision_cache.py:47-50 writes ~/.cache/nclone."""
    if not os.path.isdir(REF_ROOT):
        raise RuntimeError("reference tree not present (expected on the GPU box)")
    scratch = os.environ.get("NCLONE_REF_HOME", "/tmp/nclone_ref_home")
    os.makedirs(scratch, exist_ok=True)
    os.environ["HOME"] = scratch
    sys.dont_write_bytecode = True
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    from nclone.nsim import Simulator
    from nclone.sim_config import SimConfig
    return Simulator, SimConfig


def new_sim(map_data):
    Simulator, SimConfig = import_reference()
    sim = Simulator(SimConfig(enable_anim=False, log_data=False, debug=False))
    sim.load(list(map_data))
    return sim


def entity_rows(sim):
    ents = [e for lst in sim.entity_dic.values() for e in lst]
    ef = np.zeros((len(ents), schema.EF), dtype=np.float64)
    ei = np.zeros((len(ents), schema.EI), dtype=np.int64)
    for k, e in enumerate(ents):
        name = type(e).__name__
        s0 = s1 = 0
        a0 = a1 = 0.0
        if name == "EntityToggleMine":
            s0 = e.state
        elif name == "EntityExit":
            s0 = int(e.switch_hit)
        elif name in ("EntityDoorRegular", "EntityDoorLocked", "EntityDoorTrap"):
            s0 = int(e.closed)
            if name == "EntityDoorRegular":
                s1 = e.open_timer
        elif name in ("EntityBounceBlock", "EntityDeathBall"):
            a0, a1 = e.xspeed, e.yspeed
        elif name in ("EntityDroneZap", "EntityMiniDrone"):
            s0 = e.dir
            a0, a1 = e.xtarget, e.ytarget
        elif name == "EntityThwump":
            s0 = e.state
        elif name == "EntityBoostPad":
            s0 = int(e.is_touching_ninja)
        elif name == "EntityShoveThwump":
            s0 = e.state
            s1 = int(e.activated)
            a0, a1 = e.xdir, e.ydir
        cell_list = sim.grid_entity[e.cell]
        rank = -1
        for r, other in enumerate(cell_list):
            if other is e:
                rank = r
                break
        ei[k] = (e.type, e.index, int(e.active), e.cell[0], e.cell[1], s0, s1, rank)
        ef[k] = (e.xpos, e.ypos, a0, a1)
    return ef, ei


def ninja_rows(sim):
    n = sim.ninja
    nf = np.array([float(getattr(n, f, 0.0)) for f in schema.NINJA_F], dtype=np.float64)
    vals = []
    for f in schema.NINJA_I:
        if f == "frame":
            vals.append(sim.frame)
        elif f == "death_cause":
            vals.append(schema.DEATH_CAUSE.get(n.death_cause, 3))
        else:
            vals.append(int(getattr(n, f)))
    return nf, np.array(vals, dtype=np.int64)


def canonical(sim):
    nf, ni = ninja_rows(sim)
    ef, ei = entity_rows(sim)
    return nf, ni, ef, ei


_IGM = None


def apply_curriculum(sim, switch_xy, door_xy):
    """Run the reference's own IntermediateGoalManager.apply_to_simulator
    (gym_environment/reward_calculation/intermediate_goal_manager.py:698-960) on `sim` with the given target
    positions.  The module is loaded by path (its package __init__ chain needs gymnasium); `self` is a stand-in
    that only answers the two position getters and the fields the method logs."""
    global _IGM
    if _IGM is None:
        import importlib.util
        import logging
        p = os.path.join(REF_ROOT, "nclone", "gym_environment", "reward_calculation", "intermediate_goal_manager.py")
        spec = importlib.util.spec_from_file_location("_ref_intermediate_goal_manager", p)
        _IGM = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_IGM)
        _IGM.logger.setLevel(logging.CRITICAL)
    from types import SimpleNamespace
    fake = SimpleNamespace(
        get_curriculum_switch_position=lambda: (switch_xy[0], switch_xy[1]),
        get_curriculum_exit_position=lambda: (door_xy[0], door_xy[1]),
        state=SimpleNamespace(unified_stage=0), config=SimpleNamespace(stage_distance_interval=100.0),
        _spawn_to_switch_distance=0.0, _combined_distance=0.0)
    _IGM.IntermediateGoalManager.apply_to_simulator(fake, sim)


def run_trajectory(map_data, hor, jump, full=False, moves=()):
    """Tick the reference through (hor[t], jump[t]) and return per-frame records.
    Frame 0 is the state right after load().  moves: (t, sx, sy, dx, dy) tuples -- the curriculum's
    apply_to_simulator before tick t (t = 0: before the frame-0 record)."""
    sim = new_sim(map_data)
    T = len(hor)
    for m in moves:
        if m[0] == 0:
            apply_curriculum(sim, m[1:3], m[3:5])
    nf0, ni0, ef0, ei0 = canonical(sim)
    NF = np.zeros((T + 1, schema.NF)); NI = np.zeros((T + 1, schema.NI), dtype=np.int64)
    H = np.zeros(T + 1, dtype=np.uint64)
    EFs = []; EIs = []
    NF[0], NI[0] = nf0, ni0
    H[0] = schema.state_hash(nf0, ni0, ef0, ei0)
    if full:
        EFs.append(ef0); EIs.append(ei0)
    for t in range(T):
        for m in moves:
            if m[0] == t and t > 0:
                apply_curriculum(sim, m[1:3], m[3:5])
        sim.tick(int(hor[t]), int(jump[t]))
        nf, ni, ef, ei = canonical(sim)
        NF[t + 1], NI[t + 1] = nf, ni
        H[t + 1] = schema.state_hash(nf, ni, ef, ei)
        if full:
            EFs.append(ef); EIs.append(ei)
    out = {"nin_f": NF, "nin_i": NI, "hash": H}
    if full:
        out["ent_f"] = np.stack(EFs); out["ent_i"] = np.stack(EIs)
    return out
