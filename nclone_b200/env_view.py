"""EnvView: env *i* of a BatchedSimulator behind the reference's single-env interface.

The unmodified callers above the seam (gym_environment/base_environment.py, replay/replay_executor.py,
action_replayer.py) talk to ``NPlayHeadless`` (/root/reference/nclone/nplay_headless.py:195-331, 401-455, 546-720)
and read ``sim.frame``, ``sim.ninja.<attr>``, ``sim.entity_dic[type]``, ``sim.map_data`` and ``sim.tile_dic``
directly (base_environment.py:2950-2989, 3346-3360).  An EnvView answers exactly those calls for one env of a
batch, with the reference's scalar return types.  It is host glue: every read copies that env's record from HBM
(cached until the batch launches again), every ``tick`` launches the batch kernel with the other envs held
(``tick_limit`` one-hot) -- correct, and meant for debugging / gradual migration; a vectorised caller uses the
batched methods of BatchedSimulator, which move nothing to the host.
"""
import numpy as np
import torch

from . import fields as FD

# reference entity.type -> key of sim.entity_dic (the exit switch, type 4, is filed under 3: entity_factory.py:183-200)
_DIC_KEY = {4: 3}
_TYPE_ATTRS = {   # canonical columns (a0, a1, s0, s1) -> the reference's attribute names per entity type
    1: {"state": "s0"}, 21: {"state": "s0"},
    3: {"switch_hit": "s0"},
    5: {"closed": "s0", "open_timer": "s1"}, 6: {"closed": "s0"}, 8: {"closed": "s0"},
    14: {"dir": "s0", "xtarget": "a0", "ytarget": "a1"}, 26: {"dir": "s0", "xtarget": "a0", "ytarget": "a1"},
    17: {"xspeed": "a0", "yspeed": "a1"}, 25: {"xspeed": "a0", "yspeed": "a1"},
    20: {"state": "s0"}, 24: {"is_touching_ninja": "s0"},
    28: {"state": "s0", "activated": "s1", "xdir": "a0", "ydir": "a1"},
}
_BOOL_ATTRS = {"switch_hit", "closed", "is_touching_ninja", "activated"}


class EntityView:
    """Snapshot of one entity with the reference's attribute names (entities.py:209-236 + entity_classes/*)."""

    def __init__(self, ef, ei):
        self.type, self.index, self.active = int(ei[0]), int(ei[1]), bool(ei[2])
        self.xpos, self.ypos = float(ef[0]), float(ef[1])
        self.cell = (int(ei[3]), int(ei[4]))
        col = {"a0": float(ef[2]), "a1": float(ef[3]), "s0": int(ei[5]), "s1": int(ei[6])}
        for name, c in _TYPE_ATTRS.get(self.type, {}).items():
            v = col[c]
            setattr(self, name, bool(v) if name in _BOOL_ATTRS else v)

    def __repr__(self):
        return f"EntityView(type={self.type}, index={self.index}, active={self.active}, pos=({self.xpos}, {self.ypos}))"


class NinjaView:
    """sim.ninja: every field of the canonical record by its reference name (ninja.py:80-196)."""

    def __init__(self, nf, ni):
        for k, name in enumerate(FD.NINJA_F):
            setattr(self, name, float(nf[k]))
        for k, name in enumerate(FD.NINJA_I):
            if name not in ("frame", "death_cause"):
                setattr(self, name, int(ni[k]))
        for name in ("airborn", "airborn_old", "walled", "is_jump_useless", "last_jump_was_buffered", "is_crushable", "terminal_impact"):
            setattr(self, name, bool(getattr(self, name)))
        self.death_cause = FD.DEATH_CAUSES[int(ni[FD.I["death_cause"]])] if int(ni[FD.I["death_cause"]]) < 3 else "other"

    def has_won(self):            # ninja.py:1276-1278
        return self.state == 8

    def has_died(self):           # ninja.py:1280-1282
        return self.state in (6, 7)

    def is_valid_target(self):    # ninja.py:1272-1274
        return self.state not in (6, 8, 9)


class SimView:
    """The attributes callers read from NPlayHeadless.sim."""

    def __init__(self, view):
        self._v = view
        self._action_masking_mode = "hard"            # written by base_environment.py:3100-3105
        self._path_straightness_direction = None

    @property
    def frame(self):
        return int(self._v._record()[1][FD.I["frame"]])

    @property
    def ninja(self):
        nf, ni = self._v._record()[:2]
        return NinjaView(nf, ni)

    @property
    def entity_dic(self):
        _, _, ef, ei = self._v._record()
        dic = {}
        for k in range(len(ei)):
            e = EntityView(ef[k], ei[k])
            dic.setdefault(_DIC_KEY.get(e.type, e.type), []).append(e)
        return dic

    @property
    def map_data(self):
        return self._v.map_data

    @property
    def tile_dic(self):
        """{(x, y): tile id} over the 44 x 25 cell grid, border included (map_loader.py:18-67)."""
        t = self._v._tables().tiles
        return {(x, y): int(t[x * 25 + y]) for x in range(44) for y in range(25)}


class EnvView:
    def __init__(self, batch, env, map_data=None):
        if not 0 <= int(env) < batch.n_envs:
            raise IndexError(env)
        self.batch, self.env = batch, int(env)
        self.map_data = map_data
        self.sim = SimView(self)
        self.current_tick = 0
        self._cache = (None, None)

    # ---- plumbing
    def _record(self):
        key = self.batch.launch_count()
        if self._cache[0] != key:
            self._cache = (key, self.batch.canonical_state([self.env])[0])
        return self._cache[1]

    def _tables(self):
        return self.batch.maps[self.batch.env_map_host[self.env]]

    def _one_hot(self):
        m = np.zeros(self.batch.n_envs, dtype=np.uint8)
        m[self.env] = 1
        return m

    # ---- NPlayHeadless.tick / reset (nplay_headless.py:322-331, 257-269)
    def tick(self, horizontal_input, jump_input):
        self.current_tick += 1
        N = self.batch.n_envs
        hor = np.zeros(N, dtype=np.int8); jump = np.zeros(N, dtype=np.int8)
        hor[self.env], jump[self.env] = int(horizontal_input), int(jump_input)
        self.batch.tick(hor, jump, n_ticks=1, tick_limit=self._one_hot().astype(np.int32))

    def reset(self):
        self.current_tick = 0
        self.batch.reset(self._one_hot())

    def fast_reset(self):
        self.current_tick = 0
        self.batch.fast_reset(self._one_hot())

    def load_map_from_map_data(self, map_data):
        raise NotImplementedError("levels are loaded for the whole batch: BatchedSimulator.load_maps(map_datas, env_to_map)")

    # ---- getters (nplay_headless.py:401-455, 546-616, 735-917)
    def ninja_has_won(self):
        return self.sim.ninja.has_won()

    def ninja_has_died(self):
        return self.sim.ninja.has_died()

    def ninja_death_cause(self):
        return self.sim.ninja.death_cause

    def ninja_position(self):
        n = self.sim.ninja
        return n.xpos, n.ypos

    def ninja_velocity(self):
        n = self.sim.ninja
        return n.xspeed, n.yspeed

    def ninja_velocity_old(self):
        n = self.sim.ninja
        return n.xspeed_old, n.yspeed_old

    def ninja_is_in_air(self):
        return self.sim.ninja.airborn

    def ninja_is_walled(self):
        return self.sim.ninja.walled

    def _sim_exit_switch(self):
        for e in reversed(self.sim.entity_dic.get(3, [])):
            if e.type == 4:
                return e
        return None

    def exit_switch_activated(self):
        sw = self._sim_exit_switch()
        return True if sw is None else not sw.active

    def exit_switch_position(self):
        sw = self._sim_exit_switch()
        return (0, 0) if sw is None else (sw.xpos, sw.ypos)

    def exit_door_position(self):
        """The parent of the last exit switch: doors and switches are created pairwise in file order
        (entity_factory.py:183-200), so it is the door with the switch's rank among type-3 entries."""
        ents = self.sim.entity_dic.get(3, [])
        n_sw = sum(1 for e in ents if e.type == 4)
        doors = [e for e in ents if e.type == 3]
        return (0, 0) if n_sw == 0 else (doors[n_sw - 1].xpos, doors[n_sw - 1].ypos)

    def get_ninja_state(self):
        """40 float32 features, computed on the device (nppsim_observe)."""
        return self.batch.get_ninja_state()[self.env].cpu().numpy()

    def get_action_mask(self):
        pd = {None: 0, "left": 1, "right": 2, "down": 3}[self.sim._path_straightness_direction]
        full = torch.full((self.batch.n_envs,), pd, dtype=torch.uint8, device=self.batch.device)
        return [bool(v) for v in self.batch.get_action_mask(full)[self.env].cpu().numpy()]

    def mines(self):              # nplay_headless.py:618-623
        return [e for e in self.sim.entity_dic.get(1, []) if e.active and e.state == 0]

    def locked_doors(self):
        return list(self.sim.entity_dic.get(6, []))
