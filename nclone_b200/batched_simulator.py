"""BatchedSimulator: N independent nclone simulators advanced by one CUDA launch.

Host-side mirror of the reference's ``nsim.Simulator`` (/root/reference/nclone/nsim.py:11-292)
and of the getters of ``nplay_headless.NPlayHeadless`` (nplay_headless.py:401-455, 546-720),
batched: every method that takes/returns a scalar there takes/returns an array over envs here.
PyTorch only owns device memory and streams; all physics runs in libnppsim.so through the
C ABI of include/nppsim.h.  There is no CPU path: construction fails without a CUDA device.
"""
import ctypes as C

import numpy as np
import torch

from . import fields as FD
from ._lib import lib, check, NppsimState, NppsimInputStream, STREAM_ACTIONS, STREAM_REPLAY_BYTES
from .map_compile import compile_map, content_key, MapTables


class BatchedSimulator:
    def __init__(self, n_envs, device="cuda:0", handoff=None, placement="auto"):
        """handoff: run wedged envs on solo warps (nppsim_state_t.handoff; identical results, shorter step kernel).
        Default (None): by batch size, see _auto_handoff.
        placement: "level" keeps the envs of one level in consecutive state columns, whatever order the caller numbers them
        in -- the lanes of a warp then read the same level tables and run the same entity loops -- and every per-env array of
        this API stays in the caller's order (nppsim_state_t.ext_of); "caller": columns = env ids; "auto" (default): "level"
        when a level has at least a warp's worth of replicas (n_envs >= 32 x distinct levels), else "caller"."""
        if not torch.cuda.is_available():
            raise RuntimeError("BatchedSimulator needs a CUDA device (there is no CPU fallback)")
        self.n_envs = int(n_envs)
        self._handoff_arg = handoff
        if placement not in ("level", "caller", "auto"):
            raise ValueError("placement must be 'level', 'caller' or 'auto'")
        self.placement = placement
        self.slot_env = None        # numpy [N]: caller env id of state column s (None: identity)
        self._slot_of = None        # device int64 [N]: state column of caller env e
        self.device = torch.device(device)
        self.use_handoff = self._auto_handoff(self.n_envs) if handoff is None else bool(handoff)
        self.L = lib()
        self.maps = []          # MapTables per distinct level
        self._keys = {}
        self._state = None
        self.env_map_host = None

    # ------------------------------------------------------------------ Simulator.load (nsim.py:51-56)
    def load_maps(self, map_datas, env_to_map=None):
        """map_datas: sequence of map_data rows (map_loader format, ints or floats, any length
        >= 1235); env_to_map: for each env the index into map_datas (default: env i -> map i % M).
        Identical levels are compiled and stored once.  All envs are reset (mode 'reset')."""
        N = self.n_envs
        M = len(map_datas)
        if env_to_map is None:
            env_to_map = np.arange(N) % M
        env_to_map = np.asarray(env_to_map, dtype=np.int64)
        if env_to_map.shape != (N,) or env_to_map.min() < 0 or env_to_map.max() >= M:
            raise ValueError("env_to_map must hold one valid map index per env")
        self.maps, self._keys = [], {}
        remap = np.zeros(M, dtype=np.int32)
        blobs = []
        for m, md in enumerate(map_datas):
            key = content_key(md)
            if key not in self._keys:
                blob = compile_map(md)
                self._keys[key] = len(blobs)
                blobs.append(blob)
                self.maps.append(MapTables(blob))
            remap[m] = self._keys[key]
        offs = np.zeros(len(blobs), dtype=np.int64)
        total = 0
        for k, b in enumerate(blobs):
            offs[k] = total
            total += (len(b) + 255) & ~255
        host_blob = np.zeros(max(total, 256), dtype=np.uint8)
        for k, b in enumerate(blobs):
            host_blob[offs[k]:offs[k] + len(b)] = np.frombuffer(b, dtype=np.uint8)
        dev = self.device
        self.map_blob = torch.from_numpy(host_blob).to(dev)
        self.map_off = torch.from_numpy(offs).to(dev)
        self.env_map_host = remap[env_to_map].astype(np.int32)
        self._place(self.env_map_host)
        kind_mask = 0
        for t in self.maps:
            kind_mask |= t.kind_mask           # entity-kind census of the level set -> step-kernel variant
        self._alloc_state(len(blobs), max(t.n_ent for t in self.maps), max(t.n_dyn for t in self.maps), kind_mask)
        self.reset(mode="load")

    def load_maps_device(self, map_data, map_len=None, env_to_map=None):
        """Simulator.load for levels that are already in HBM: map_data is a CUDA float64 tensor
        [M, L] of map_loader rows (map_len: optional int32 [M] valid lengths), compiled to blobs by ONE kernel launch
        (nppsim_compile_maps_device) -- no per-level host work, nothing crosses PCIe but a 16-byte summary.  Levels are
        not deduplicated (the caller chooses env_to_map; default env i -> level i % M).  Raises ValueError if any level
        is malformed (first offender reported), like load_maps."""
        N, dev = self.n_envs, self.device
        md = torch.as_tensor(map_data)
        if md.device.type != "cuda" or md.dtype != torch.float64 or md.dim() != 2:
            raise ValueError("map_data must be a CUDA float64 tensor [M, L]")
        md = md.contiguous()
        M, Lrow = md.shape
        if env_to_map is None:
            env_map = torch.arange(N, device=dev, dtype=torch.int32) % M
        else:
            env_map = torch.as_tensor(env_to_map).to(device=dev, dtype=torch.int32).contiguous()
            if env_map.shape != (N,) or int(env_map.min()) < 0 or int(env_map.max()) >= M:
                raise ValueError("env_to_map must hold one valid map index per env")
        ml = None
        if map_len is not None:
            ml = torch.as_tensor(map_len).to(device=dev, dtype=torch.int32).contiguous()
            if ml.shape != (M,):
                raise ValueError("map_len must hold one length per level")
        stride = (int(self.L.nppsim_map_blob_bound(int(Lrow))) + 255) & ~255
        self.map_blob = torch.empty(M * stride, dtype=torch.uint8, device=dev)
        self.map_off = torch.empty(M, dtype=torch.int64, device=dev)
        self.compile_status = torch.empty(M, dtype=torch.int32, device=dev)
        summary = torch.empty(4, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            check(self.L.nppsim_compile_maps_device(C.c_void_p(md.data_ptr()), int(Lrow), C.c_void_p(ml.data_ptr()) if ml is not None else None,
                                                    int(Lrow), int(M), C.c_void_p(self.map_blob.data_ptr()), stride,
                                                    C.c_void_p(self.map_off.data_ptr()), C.c_void_p(self.compile_status.data_ptr()),
                                                    C.c_void_p(summary.data_ptr()), self._stream()))
        ent_cap, dyn_cap, kind_mask, n_bad = (int(v) for v in summary.cpu().numpy().view(np.uint32))
        if n_bad:
            bad = torch.nonzero(self.compile_status)[0].item()
            raise ValueError(f"map compile failed for {n_bad} level(s); first: level {bad}, code {int(self.compile_status[bad])}")
        self._blob_stride = stride
        self.maps = _DeviceMaps(self.map_blob, stride, M)
        self._keys = {}
        self.env_map_host = env_map.cpu().numpy()
        self._place(self.env_map_host)
        self._alloc_state(M, ent_cap, dyn_cap, kind_mask)
        self.reset(mode="load")

    def _place(self, env_map_host):
        """State-column order (placement='level'): envs sorted by level, stable, unless that is the caller's order anyway
        or no two envs share a level.  Sets env_map (device, column order), slot_env / _slot_of / ext_of."""
        N, dev = self.n_envs, self.device
        order = None
        want = self.placement == "level"
        if self.placement == "auto":                    # a level must be able to fill a warp for the grouping to pay (profiles/r02_handoff_ab.md)
            want = N >= 32 * len(np.unique(env_map_host))
        if want and N > 1:
            o = np.argsort(env_map_host, kind="stable")
            if not np.array_equal(o, np.arange(N)) and len(np.unique(env_map_host)) < N:
                order = o
        if order is None:
            self.slot_env, self._slot_of, self.ext_of = None, None, None
            self.env_map = torch.from_numpy(np.ascontiguousarray(env_map_host)).to(dev)
        else:
            self.slot_env = order.astype(np.int64)
            inv = np.empty(N, dtype=np.int64); inv[order] = np.arange(N)
            self._slot_of = torch.from_numpy(inv).to(dev)
            self.ext_of = torch.from_numpy(order.astype(np.int32)).to(dev)
            self.env_map = torch.from_numpy(np.ascontiguousarray(env_map_host[order])).to(dev)

    # state tensors in the CALLER's env order (copies when the engine keeps another column order; the raw column-order
    # tensors are _nin_f ... _dyn_i)
    def _ext(self, t):
        return t if self._slot_of is None else t[..., self._slot_of]

    nin_f = property(lambda self: self._ext(self._nin_f))
    nin_i = property(lambda self: self._ext(self._nin_i))
    env_i = property(lambda self: self._ext(self._env_i))
    ent_flag = property(lambda self: self._ext(self._ent_flag))
    dyn_f = property(lambda self: self._ext(self._dyn_f))
    dyn_i = property(lambda self: self._ext(self._dyn_i))
    frames_executed = property(lambda self: self._frames_executed)         # written by the kernels in caller order

    def _alloc_state(self, n_blobs, ent_cap, dyn_cap, kind_mask):
        N, dev = self.n_envs, self.device
        self._exit_cache = None
        self.ent_cap = max(1, int(ent_cap))
        self.dyn_cap = max(1, int(dyn_cap))
        self._nin_f = torch.zeros((FD.NF, N), dtype=torch.float64, device=dev)
        self._nin_i = torch.zeros((FD.NI, N), dtype=torch.int32, device=dev)
        self._env_i = torch.zeros((FD.NENV, N), dtype=torch.int32, device=dev)
        self._ent_flag = torch.zeros((self.ent_cap, N), dtype=torch.uint8, device=dev)
        self._dyn_f = torch.zeros((self.dyn_cap * 4, N), dtype=torch.float64, device=dev)
        self._dyn_i = torch.zeros((self.dyn_cap * 3, N), dtype=torch.int32, device=dev)
        self._frames_executed = torch.zeros(N, dtype=torch.int32, device=dev)
        self._actions_dev = torch.zeros(N, dtype=torch.uint8, device=dev)
        self._staging = torch.zeros(int(self.L.nppsim_host_block_bytes(N, 1)) + 8, dtype=torch.uint8, device=dev)
        self.stats = torch.zeros(4, dtype=torch.int64, device=dev)
        st = NppsimState()
        st.n_envs, st.ent_cap, st.dyn_cap, st.n_maps = N, self.ent_cap, self.dyn_cap, int(n_blobs)
        st.nin_f, st.nin_i, st.env_i = self._nin_f.data_ptr(), self._nin_i.data_ptr(), self._env_i.data_ptr()
        st.ent_flag, st.dyn_f, st.dyn_i = self._ent_flag.data_ptr(), self._dyn_f.data_ptr(), self._dyn_i.data_ptr()
        st.env_map, st.map_blob, st.map_off = self.env_map.data_ptr(), self.map_blob.data_ptr(), self.map_off.data_ptr()
        st.ext_of = self.ext_of.data_ptr() if self.ext_of is not None else None
        nw = (N + 31) // 32
        self.warp_order = torch.arange(nw, dtype=torch.int32, device=dev)      # longest-first launch order (csrc/sim_variant.inc)
        self.warp_cost = torch.zeros(nw, dtype=torch.int32, device=dev)
        st.warp_order, st.warp_cost = self.warp_order.data_ptr(), self.warp_cost.data_ptr()
        self.kind_mask = int(kind_mask)
        st.kind_mask = self.kind_mask
        # solo warps exist in the basic kernel only (csrc/nppsim_kernels.cu: NPP_SOLO)
        self.handoff = self._new_handoff() if (self.use_handoff and self._is_basic(self.kind_mask)) else None
        self._retired_handoff = None
        st.handoff = self.handoff.data_ptr() if self.handoff is not None else None
        self._adapt_calls, self._adapt_pending, self._adapt_seen, self._adapt_buf = 0, None, (0, 0), None
        self._state = st

    # ------------------------------------------------------------------ sub-batches for asynchronous stepping
    def groups(self, n_groups):
        """Split the batch into n_groups views over contiguous env ranges (multiples of 32 envs).  A view is a
        BatchedSimulator over the SAME device memory (column slices, nppsim_state_t.stride = the parent's n_envs), so it
        can be stepped / reset / observed on its own stream while the other groups do something else -- the policy acting
        on one group while the next one steps, as asynchronous RL loops do.  The kernel of one group then overlaps the
        long tail of another's (a wedged ninja's serial depenetration chain): +30 % env-steps/s at 131,072 envs per GPU.
        Do not mix calls on the parent with calls on its groups without synchronising in between."""
        self._need_maps()
        N = self.n_envs
        per = -(-N // n_groups)
        per = (per + 31) // 32 * 32
        views = []
        for b in range(0, N, per):
            views.append(self._view(b, min(N, b + per)))
        return views

    env_ids = None      # on a group view: the parent's env ids of the view's envs 0..n-1 (numpy)

    def _view(self, b, e):
        v = BatchedSimulator.__new__(BatchedSimulator)
        v.__dict__.update(self.__dict__)
        v.n_envs = e - b
        v._exit_cache = None
        v._adapt_calls, v._adapt_pending, v._adapt_seen, v._adapt_buf, v._retired_handoff = 0, None, (0, 0), None, None      # (_adapt_handoff: per queue)
        # a view is a range of state COLUMNS; its own env numbering is the column order (no permutation inside a view)
        for name in ("_nin_f", "_nin_i", "_env_i", "_ent_flag", "_dyn_f", "_dyn_i"):
            setattr(v, name, getattr(self, name)[:, b:e])
        for name in ("env_map", "_frames_executed", "_actions_dev"):
            setattr(v, name, getattr(self, name)[b:e])
        v._staging = torch.zeros(int(self.L.nppsim_host_block_bytes(e - b, 1)) + 8, dtype=torch.uint8, device=self.device)
        v.env_ids = (self.slot_env if self.slot_env is not None else np.arange(self.n_envs))[b:e]
        v.env_map_host = self.env_map_host[v.env_ids]
        v.slot_env, v._slot_of, v.ext_of = None, None, None
        v.warp_order = self.warp_order[b // 32:(e + 31) // 32]
        v.warp_cost = self.warp_cost[b // 32:(e + 31) // 32]
        v.warp_order.copy_(torch.arange(v.warp_order.numel(), dtype=torch.int32, device=self.device))   # group-local slots
        parent = self._state
        st = NppsimState()
        st.n_envs, st.ent_cap, st.dyn_cap, st.n_maps = v.n_envs, parent.ent_cap, parent.dyn_cap, parent.n_maps
        st.stride = parent.stride or parent.n_envs
        st.nin_f, st.nin_i, st.env_i = v._nin_f.data_ptr(), v._nin_i.data_ptr(), v._env_i.data_ptr()
        st.ent_flag, st.dyn_f, st.dyn_i = v._ent_flag.data_ptr(), v._dyn_f.data_ptr(), v._dyn_i.data_ptr()
        st.env_map, st.map_blob, st.map_off = v.env_map.data_ptr(), parent.map_blob, parent.map_off
        st.warp_order, st.warp_cost = v.warp_order.data_ptr(), v.warp_cost.data_ptr()
        v.use_handoff = self._auto_handoff(v.n_envs) if self._handoff_arg is None else self.use_handoff
        v._handoff_arg = self._handoff_arg
        v.handoff = self._new_handoff() if (v.use_handoff and self._is_basic(parent.kind_mask)) else None      # one queue per concurrently stepped state
        st.handoff = v.handoff.data_ptr() if v.handoff is not None else None
        st.kind_mask = parent.kind_mask
        v._state = st
        return v

    def _auto_handoff(self, n_envs):
        """Solo-warp handoff of wedged envs (solo SMs, mid-tick handoff) is on from 1,024 envs: it pays wherever the wedged
        ninjas are the tail of a launch (config 3: 65,536 envs +8 %, 131,072 +10 %) and loses where they are the bulk of the
        work (config 4's locked-door levels -20 %, config 3 at 262,144 envs -11 %: one warp per env takes the slots the
        regular warps need).  That is a property of the levels, not of the batch size, so the engine measures it: sched_kernel
        counts the launches whose wedged envs overflowed the queue, and _adapt_handoff drops the queue when most do
        (profiles/r02_handoff_ab.md section 11)."""
        return 1024 <= n_envs

    def _adapt_handoff(self):
        """Called after every step: every 16th launch, read the queue's error word and saturation counters (asynchronously --
        the copy is looked at once its event has completed).  A set error word (a solo warp gave up waiting for an entry, so an
        env may have missed ticks: csrc/sim_variant.inc) raises -- such a batch is no longer the reference's.  Under the
        automatic policy the queue is then dropped when more than half of the launches overflowed it, or when no env has used
        it in 128 launches.  The kernels without a queue run every env in its home lane; results are identical."""
        if self.handoff is None:
            return
        self._adapt_calls = getattr(self, "_adapt_calls", 0) + 1
        pending = getattr(self, "_adapt_pending", None)
        if pending is not None:
            ev, buf = pending
            if not ev.query():
                return
            self._adapt_pending = None
            if int(buf[0]) != 0:                                       # word 3, sticky
                raise RuntimeError(f"handoff queue error bits {int(buf[0])}: a solo warp timed out waiting for an env; "
                                   "the batch may have lost ticks (reload, or step with handoff=False)")
            if self._handoff_arg is not None:                          # queue forced on or off by the caller: no policy
                return
            used, clipped, total = int(buf[1]) + int(buf[4]), int(buf[9]), int(buf[10])      # words 4 + 7 | 12 | 13
            seen = getattr(self, "_adapt_seen", (0, 0))
            drop = False
            if total - seen[1] >= 16:
                self._adapt_seen = (clipped, total)
                drop = 2 * (clipped - seen[0]) > total - seen[1]
            # and a queue no env has ever used after 128 launches (levels without corners to wedge into, e.g. config 2's
            # single level) only costs the solo kernel's 3-6 % over the plain one
            if not drop and total >= 128 and used == 0:
                drop = True
            if drop:
                self._retired_handoff = self.handoff              # (launches in flight still read it)
                self.handoff = None
                self._state.handoff = None
            return
        if self._adapt_calls % 16 == 0:
            buf = getattr(self, "_adapt_buf", None)
            if buf is None:
                buf = self._adapt_buf = torch.zeros(11, dtype=torch.int32).pin_memory()
            with torch.cuda.device(self.device):
                buf.copy_(self.handoff[3:14], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record()
            self._adapt_pending = (ev, buf)

    _handoff_tags = 0

    def _new_handoff(self):
        """A zeroed handoff queue (nppsim_state_t.handoff) with a tag no other queue of this process carries."""
        BatchedSimulator._handoff_tags += 1
        q = torch.zeros(FD.HANDOFF_WORDS, dtype=torch.int32, device=self.device)
        q[6] = BatchedSimulator._handoff_tags
        return q

    def force_full_kernel(self, on=True):
        """Test hook: run this batch on the full (all entity kinds) step kernel whatever its level set holds, by
        declaring the kind census unknown (nppsim_state_t.kind_mask = 0)."""
        self._need_maps()
        self._state.kind_mask = 0 if on else self.kind_mask

    @staticmethod
    def _is_basic(kind_mask):
        basic = sum(1 << k for k in (FD.K_NONE, FD.K_MINE, FD.K_GOLD, FD.K_EXIT, FD.K_EXIT_SWITCH, FD.K_DOOR_LOCKED, FD.K_DOOR_TRAP))
        return kind_mask != 0 and (kind_mask & ~basic) == 0

    def kernel_variant(self):
        """Which step-kernel instantiation launch_step picks for this batch (nppsim_kernels.cu: NPP_BASIC_MASK, NPP_SOLO)."""
        self._need_maps()
        if self._is_basic(int(self._state.kind_mask)):
            return "sim_basic_solo" if self.handoff is not None else "sim_basic"
        return "sim_full"

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _need_maps(self):
        if self._state is None:
            raise RuntimeError("load_maps() first")

    # ------------------------------------------------------------------ Simulator.reset / fast_reset (nsim.py:62-140)
    RESET_MODES = {"reset": 0, "fast": 1, "load": 2}

    def reset(self, mask=None, mode="reset"):
        """mode 'reset' = Simulator.reset() as the reference's callers issue it, on the Simulator that has held the level since
        load_maps (NPlayHeadless.reset, ActionReplayer): entities are re-created and their `index` keeps counting
        (entities.py:213-216), so death balls no longer repel each other afterwards (entity_death_ball.py:153-166);
        'fast' = Simulator.fast_reset; 'load' = a fresh Simulator loading the level (what load_maps does)."""
        self._need_maps()
        m = self.RESET_MODES[mode]
        ptr = None
        if mask is not None:
            mask = torch.as_tensor(mask, device=self.device).to(torch.uint8).contiguous()
            ptr = C.c_void_p(mask.data_ptr())
        with torch.cuda.device(self.device):
            check(self.L.nppsim_reset(C.byref(self._state), ptr, m, self._stream()))

    def fast_reset(self, mask=None):
        self.reset(mask, mode="fast")

    def reset_terminal(self, max_frames=0, mode="reset"):
        """Episode turnover for the batch: reset the envs that won, died or reached max_frames.
        Episode statistics accumulate in self.stats (device uint64[4]: episodes, wins, deaths, truncations)."""
        self._need_maps()
        with torch.cuda.device(self.device):
            check(self.L.nppsim_reset_terminal(C.byref(self._state), int(max_frames), self.RESET_MODES[mode],
                                               C.c_void_p(self.stats.data_ptr()), self._stream()))

    # ------------------------------------------------------------------ curriculum mutation from outside tick
    def set_exit_positions(self, switch_xy, door_xy, mask=None):
        """IntermediateGoalManager.apply_to_simulator for the batch
        (gym_environment/reward_calculation/intermediate_goal_manager.py:698-960): move every masked env's exit switch
        to switch_xy[:, i] and its door to door_xy[:, i] (float64 [2, N]), re-arm the switch while frame <= 10, and
        re-home both in the entity grid as the reference does.  reset()/fast_reset() restore the level's own positions.
        Raises RuntimeError if a masked env's level has no exit switch, like the reference."""
        self._need_maps()
        N, dev = self.n_envs, self.device
        sxy = torch.as_tensor(switch_xy).to(device=dev, dtype=torch.float64).contiguous()
        dxy = torch.as_tensor(door_xy).to(device=dev, dtype=torch.float64).contiguous()
        if sxy.shape != (2, N) or dxy.shape != (2, N):
            raise ValueError("switch_xy and door_xy must be [2, n_envs]")
        ptr = None
        if mask is not None:
            mask = torch.as_tensor(mask, device=dev).to(torch.uint8).contiguous()
            ptr = C.c_void_p(mask.data_ptr())
        status = torch.zeros(N, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            check(self.L.nppsim_set_exit_positions(C.byref(self._state), ptr, C.c_void_p(sxy.data_ptr()), C.c_void_p(dxy.data_ptr()),
                                                   C.c_void_p(status.data_ptr()), self._stream()))
        if bool((status != 0).any()):
            bad = torch.nonzero(status)[0].item()
            raise RuntimeError(f"No EntityExitSwitch found in entity_dic[3] (env {bad})")

    def total_ticks(self):
        return int(self._env_i[FD.ENV_TICKS].to(torch.int64).sum().item())

    # ------------------------------------------------------------------ Simulator.tick (nsim.py:221-292)
    def _dev_i8(self, x):
        t = torch.as_tensor(x)
        if t.dtype != torch.int8 or t.device != self.device:
            t = t.to(device=self.device, dtype=torch.int8)
        return t.contiguous()

    def tick(self, hor_input, jump_input, n_ticks=1, tick_limit=None, stop_on_terminal=False):
        """hor_input / jump_input: [N] (held for n_ticks) or [n_ticks, N] (one row per tick)."""
        self._need_maps()
        hor, jump = self._dev_i8(hor_input), self._dev_i8(jump_input)
        if hor.shape != jump.shape:
            raise ValueError("hor_input and jump_input shapes differ")
        if hor.dim() == 1:
            input_ticks = 1
        else:
            input_ticks = hor.shape[0]
            if input_ticks != n_ticks:
                raise ValueError("per-tick inputs need n_ticks rows")
        if hor.shape[-1] != self.n_envs:
            raise ValueError("inputs need one column per env")
        lim = None
        if tick_limit is not None:
            tick_limit = torch.as_tensor(tick_limit).to(device=self.device, dtype=torch.int32).contiguous()
            lim = C.c_void_p(tick_limit.data_ptr())
        with torch.cuda.device(self.device):
            check(self.L.nppsim_step(C.byref(self._state), C.c_void_p(hor.data_ptr()), C.c_void_p(jump.data_ptr()),
                                     input_ticks, int(n_ticks), lim, int(bool(stop_on_terminal)),
                                     C.c_void_p(self._frames_executed.data_ptr()), self._stream()))
        self._adapt_handoff()
        return self.frames_executed

    def step(self, actions, frame_skip=4, stop_on_terminal=True):
        """Discrete(6) actions [N] (device uint8 tensor preferred) held for frame_skip ticks with the
        reference's early exit on win/death (gym_environment/base_environment.py:535-609)."""
        self._need_maps()
        a = torch.as_tensor(actions)
        if a.dtype != torch.uint8 or a.device != self.device:
            a = a.to(device=self.device, dtype=torch.uint8)
        a = a.contiguous()
        with torch.cuda.device(self.device):
            check(self.L.nppsim_step_actions(C.byref(self._state), C.c_void_p(a.data_ptr()), int(frame_skip),
                                             int(bool(stop_on_terminal)), C.c_void_p(self._frames_executed.data_ptr()),
                                             self._stream()))
        self._adapt_handoff()
        return self.frames_executed

    # ------------------------------------------------------------------ per-env input streams
    def make_stream(self, sequences, encoding="actions", repeat=1):
        """Pack one input sequence per env (any lengths, may be empty) into the ragged device layout of
        nppsim_input_stream_t.  encoding: "actions" (Discrete(6)) or "replay" (.replay input bytes)."""
        if len(sequences) != self.n_envs:
            raise ValueError("one sequence per env")
        lens = np.fromiter((len(q) for q in sequences), dtype=np.int64, count=self.n_envs)
        off = np.zeros(self.n_envs + 1, dtype=np.int64)
        np.cumsum(lens, out=off[1:])
        flat = np.zeros(max(int(off[-1]), 1), dtype=np.uint8)
        for k, q in enumerate(sequences):
            if lens[k]:
                flat[off[k]:off[k + 1]] = np.asarray(q, dtype=np.uint8)
        sp = NppsimInputStream()
        data = torch.from_numpy(flat).to(self.device); offset = torch.from_numpy(off).to(self.device)
        sp.data, sp.offset = data.data_ptr(), offset.data_ptr()
        sp.encoding = {"actions": STREAM_ACTIONS, "replay": STREAM_REPLAY_BYTES}[encoding]
        sp.repeat, sp.start_tick = int(repeat), 0
        sp._keep = (data, offset)                  # the struct only holds raw pointers
        sp._lens = lens
        return sp

    def step_stream(self, stream, n_ticks, start_tick=0, stop_on_terminal=False):
        """nppsim_step_stream: every env runs up to n_ticks ticks of its own stream, from stream tick start_tick."""
        self._need_maps()
        stream.start_tick = int(start_tick)
        with torch.cuda.device(self.device):
            check(self.L.nppsim_step_stream(C.byref(self._state), C.byref(stream), int(n_ticks), int(bool(stop_on_terminal)),
                                            C.c_void_p(self._frames_executed.data_ptr()), self._stream()))
        self._adapt_handoff()
        return self.frames_executed

    def replay_to_checkpoint(self, action_sequences, frame_skip=1, mask=None):
        """ActionReplayer.replay_to_checkpoint for the batch (action_replayer.py:79-134): reset the masked envs
        (Simulator.reset), then replay env i's stored Discrete(6) actions, frame_skip ticks each, in ONE launch
        however ragged the sequences are.  Unmasked envs are left untouched (their sequences are ignored).
        Returns the ninja positions (x, y) like _get_ninja_position."""
        self._need_maps()
        if mask is not None:
            m = np.asarray(torch.as_tensor(mask).cpu()).astype(bool)
            action_sequences = [q if m[k] else () for k, q in enumerate(action_sequences)]
        self.reset(mask)
        sp = self.make_stream(action_sequences, "actions", max(1, int(frame_skip)))
        total = int(sp._lens.max()) * sp.repeat
        if total > 0:
            self.step_stream(sp, total)
        return self.ninja_position()

    def host_block(self, with_obs=False):
        """Pinned host block for step_host and numpy views into it: {"pos_vel": f64 [4, N], "state": i32 [N]} and, with
        observations, {"ninja_state": f32 [N, 40], "action_mask": u8 [N, 6]} (layout: include/nppsim.h)."""
        N = self.n_envs
        block = torch.zeros(int(self.L.nppsim_host_block_bytes(N, int(bool(with_obs)))), dtype=torch.uint8).pin_memory()
        raw = block.numpy()
        views = {"pos_vel": raw[:32 * N].view(np.float64).reshape(4, N), "state": raw[32 * N:36 * N].view(np.int32)}
        if with_obs:
            views["ninja_state"] = raw[36 * N:196 * N].view(np.float32).reshape(N, 40)
            views["action_mask"] = raw[196 * N:202 * N].reshape(N, 6)
        return block, views

    def step_host(self, actions_host, block, frame_skip=4, stop_on_terminal=True, with_obs=False, reset_max_frames=None, reset_mode="reset"):
        """Host-buffer entry point (nppsim_step_host): numpy uint8 actions [N] in, `block` (host_block()) out -- ninja
        position / velocity / state and optionally the per-step observations, all in the caller's env order; copies are
        part of the call and the stream is synchronised on return.
        reset_max_frames (not None): nppsim_step_host_reset -- the episode turnover of reset_terminal(reset_max_frames,
        reset_mode) is queued behind the results in the same call (the block still shows the terminal states)."""
        self._need_maps()
        a = np.asarray(actions_host)
        if a.dtype != np.uint8 or a.shape != (self.n_envs,) or not a.flags["C_CONTIGUOUS"]:
            raise ValueError("actions_host must be a contiguous uint8 array with one action per env")
        need = int(self.L.nppsim_host_block_bytes(self.n_envs, int(bool(with_obs))))
        if block.dtype != torch.uint8 or block.device.type != "cpu" or block.numel() < need or not block.is_contiguous():
            raise ValueError(f"block must be a contiguous CPU uint8 tensor of at least {need} bytes (host_block())")
        stg = self._staging.data_ptr()
        stg += (-stg) % 8
        with torch.cuda.device(self.device):
            if reset_max_frames is None:
                check(self.L.nppsim_step_host(C.byref(self._state), C.c_void_p(a.ctypes.data), C.c_void_p(self._actions_dev.data_ptr()),
                                              int(frame_skip), int(bool(stop_on_terminal)), C.c_void_p(stg),
                                              C.c_void_p(block.data_ptr()), int(bool(with_obs)), self._stream()))
            else:
                check(self.L.nppsim_step_host_reset(C.byref(self._state), C.c_void_p(a.ctypes.data), C.c_void_p(self._actions_dev.data_ptr()),
                                                    int(frame_skip), int(bool(stop_on_terminal)), C.c_void_p(stg),
                                                    C.c_void_p(block.data_ptr()), int(bool(with_obs)), int(reset_max_frames),
                                                    self.RESET_MODES[reset_mode], C.c_void_p(self.stats.data_ptr()), self._stream()))
        self._adapt_handoff()

    def step_host_static(self, actions_host, block, frame_skip=4, stop_on_terminal=True, with_obs=False):
        """step_host through nppsim_step_host_static: same inputs, same block, same results, but the two device scratch
        buffers of step_host (this object's _actions_dev / _staging) are not used -- the library's built-in array stands in
        for them, which is what a caller without a device allocator has to rely on (include/nppsim.h).  At most
        STATIC_MAX_ENVS envs per call; larger batches raise (step their groups() one after the other instead).  The array is
        shared by every simulator on the device: do not call this from two threads at once."""
        self._need_maps()
        a = np.asarray(actions_host)
        if a.dtype != np.uint8 or a.shape != (self.n_envs,) or not a.flags["C_CONTIGUOUS"]:
            raise ValueError("actions_host must be a contiguous uint8 array with one action per env")
        need = int(self.L.nppsim_host_block_bytes(self.n_envs, int(bool(with_obs))))
        if block.dtype != torch.uint8 or block.device.type != "cpu" or block.numel() < need or not block.is_contiguous():
            raise ValueError(f"block must be a contiguous CPU uint8 tensor of at least {need} bytes (host_block())")
        with torch.cuda.device(self.device):
            check(self.L.nppsim_step_host_static(C.byref(self._state), C.c_void_p(a.ctypes.data), int(frame_skip), int(bool(stop_on_terminal)),
                                                 C.c_void_p(block.data_ptr()), int(bool(with_obs)), self._stream()))

    STATIC_MAX_ENVS = 16384     # NPPSIM_STATIC_MAX_ENVS

    # ------------------------------------------------------------------ NPlayHeadless getters, batched
    def _f(self, name):
        return self._ext(self._nin_f[FD.F[name]])

    def _i(self, name):
        return self._ext(self._nin_i[FD.I[name]])

    # ------------------------------------------------------------------ observation primitives
    def get_action_mask(self, path_direction=None):
        """Ninja.get_valid_action_mask for every env (nplay_headless.py:433-442, ninja.py:628-839): bool [N, 6].
        path_direction: optional uint8 tensor [N], 0 None / 1 "left" / 2 "right" / 3 "down"
        (sim._path_straightness_direction, masking mode "hard")."""
        self._need_maps()
        out = torch.empty((self.n_envs, 6), dtype=torch.uint8, device=self.device)
        pd = None
        if path_direction is not None:
            pd = torch.as_tensor(path_direction, dtype=torch.uint8, device=self.device).contiguous()
            if pd.shape != (self.n_envs,):
                raise ValueError("path_direction must hold one value per env")
        with torch.cuda.device(self.device):
            check(self.L.nppsim_observe(C.byref(self._state), C.c_void_p(pd.data_ptr()) if pd is not None else None,
                                        C.c_void_p(out.data_ptr()), None, self._stream()))
        return out.bool()

    def get_ninja_state(self):
        """NPlayHeadless.get_ninja_state for every env (nplay_headless.py:735-917): float32 [N, 40]."""
        self._need_maps()
        out = torch.empty((self.n_envs, 40), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            check(self.L.nppsim_observe(C.byref(self._state), None, None, C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def ninja_position(self):            # nplay_headless.py:415-416
        return self._f("xpos"), self._f("ypos")

    def ninja_velocity(self):            # :418-419
        return self._f("xspeed"), self._f("yspeed")

    def ninja_velocity_old(self):        # :421-423
        return self._f("xspeed_old"), self._f("yspeed_old")

    def ninja_state(self):
        return self._i("state")

    def ninja_has_won(self):             # :401-402
        return self._i("state") == 8

    def ninja_has_died(self):            # :404-405
        s = self._i("state")
        return (s == 6) | (s == 7)

    def ninja_death_cause(self):         # :407-413 (0 None, 1 "mine", 2 "terminal_impact")
        return self._i("death_cause")

    def ninja_airborn_old(self):
        return self._i("airborn_old")

    def ninja_floor_normal(self):
        return self._f("floor_normalized_x"), self._f("floor_normalized_y")

    def ninja_ceiling_normal(self):
        return self._f("ceiling_normalized_x"), self._f("ceiling_normalized_y")

    def ninja_last_jump_was_buffered(self):
        return self._i("last_jump_was_buffered")

    def get_ninja_terminal_impact(self):
        return self._i("terminal_impact")

    def frame(self):
        return self._i("frame")

    def events(self, clear=True):
        """Per-env NPPSIM_EV_* bits raised since the last clear (replaces the reference's
        sim.gym_env.invalidate_switch_cache() callback, entity_exit_switch.py:128-129)."""
        ev = self._ext(self._env_i[FD.ENV_EVENTS]).clone()
        if clear:
            self._env_i[FD.ENV_EVENTS].zero_()
        return ev

    def switch_frame(self):
        return self._ext(self._env_i[FD.ENV_SWFRAME])

    def _cols(self):
        """State column of every env, in the caller's env order (device int64 [N])."""
        return self._slot_of if self._slot_of is not None else torch.arange(self.n_envs, device=self.device)

    def _exit_slots(self):
        """Per env: entity id / dyn slot of the LAST exit switch in entity_dic[3]
        (nplay_headless.py:546-564 iterates reversed) and of its parent door; -1 if none."""
        if getattr(self, "_exit_cache", None) is None or self._exit_cache[0] is not self.env_map:
            if isinstance(self.maps, _DeviceMaps):
                # device-compiled levels: read the blob headers and entity tables where they are (no blob crosses PCIe)
                self._exit_cache = (self.env_map,) + self.maps.exit_slots(self.ent_cap, self.env_map[self._cols()])      # caller order
                return self._exit_cache[1:]
            sw_id = np.full(len(self.maps), -1); sw_slot = sw_id.copy(); dr_slot = sw_id.copy()
            for k, t in enumerate(self.maps):
                ids = np.nonzero(t.ents["kind"] == FD.K_EXIT_SWITCH)[0]
                if len(ids):
                    e = t.ents[ids[-1]]
                    sw_id[k], sw_slot[k] = ids[-1], e["dyn"]
                    dr_slot[k] = t.ents[e["aux"]]["dyn"]
            em = self.env_map_host
            dev = self.device
            self._exit_cache = (self.env_map, torch.from_numpy(sw_id[em]).to(dev), torch.from_numpy(sw_slot[em]).to(dev),
                                torch.from_numpy(dr_slot[em]).to(dev))
        return self._exit_cache[1:]

    def exit_switch_activated(self):     # nplay_headless.py:566-575 (no switch -> True)
        sw_id, _, _ = self._exit_slots()
        fl = self._ent_flag[sw_id.clamp(min=0), self._cols()]
        return (sw_id < 0) | ((fl & FD.FL_ACTIVE) == 0)

    def _slot_xy(self, slot):
        ar = self._cols()
        s = slot.clamp(min=0)
        x = self._dyn_f[s * 4, ar]; y = self._dyn_f[s * 4 + 1, ar]
        zero = torch.zeros_like(x)
        return torch.where(slot >= 0, x, zero), torch.where(slot >= 0, y, zero)

    def exit_switch_position(self):      # :577-592 ((0,0) when absent)
        return self._slot_xy(self._exit_slots()[1])

    def exit_door_position(self):        # :601-616
        return self._slot_xy(self._exit_slots()[2])

    # ------------------------------------------------------------------ host-side full state (parity tests)
    def canonical_state(self, env_ids=None):
        """Host copy of the full canonical record (ninja 31 doubles + 27 ints, and one row per
        entity in entity_dic order: doubles xpos,ypos,a0,a1; ints type,index,active,cell_x,cell_y,s0,s1,
        cell_rank) for the given envs -- the record the reference / oracle produce, for bit comparison."""
        self._need_maps()
        torch.cuda.synchronize(self.device)
        ids = np.arange(self.n_envs) if env_ids is None else np.asarray(env_ids)
        tid = self._cols()[torch.from_numpy(ids.astype(np.int64)).to(self.device)]
        nf = self._nin_f[:, tid].cpu().numpy().T.copy()
        ni = self._nin_i[:, tid].cpu().numpy().T.astype(np.int64)
        fl = self._ent_flag[:, tid].cpu().numpy()
        df = self._dyn_f[:, tid].cpu().numpy()
        di = self._dyn_i[:, tid].cpu().numpy()
        order = self._env_i[FD.ENV_ORDER][tid].cpu().numpy()
        reloads = self._env_i[FD.ENV_RELOADS][tid].cpu().numpy()
        out = []
        for k, env in enumerate(ids):
            t = self.maps[self.env_map_host[env]]
            out.append((nf[k], ni[k]) + _entity_rows(t, fl[:, k], df[:, k], di[:, k], int(order[k]), int(reloads[k])))
        return out

    def _queue(self):
        """The handoff queue in use, else the one the automatic policy retired (its counters stay readable)."""
        return self.handoff if self.handoff is not None else getattr(self, "_retired_handoff", None)

    def handoff_status(self):
        """Sticky error bits of the handoff queue (0 = fine; 1 = a solo warp gave up waiting)."""
        q = self._queue()
        return 0 if q is None else int(q[3].item())

    def handoff_count(self):
        """Envs that left their home lane for a solo warp in the middle of a launch since load_maps (0 without a queue)."""
        q = self._queue()
        return 0 if q is None else int(q[4].item())

    def solo_queue_length(self):
        """Envs that will start the next launch on a solo warp (their last tick was wedged)."""
        return 0 if self.handoff is None else int(self.handoff[7].item())

    def solo_sms(self):
        """SMs reserved for solo blocks in the next launch (0: solo blocks sit at the front of the grid; include/nppsim.h, word 10)."""
        return 0 if self.handoff is None else int(self.handoff[10].item())

    def handoff_saturation(self):
        """(launches after which the wedged envs overflowed the queue, launches) -- what _adapt_handoff decides on."""
        q = self._queue()
        return (0, 0) if q is None else (int(q[12].item()), int(q[13].item()))

    def launch_count(self):
        return int(self.L.nppsim_launch_count())


class _DeviceMaps:
    """self.maps for device-compiled levels: MapTables views are pulled from HBM on first use (host getters and
    parity tests only; the step path never needs them)."""

    def __init__(self, blob, stride, n):
        self._blob, self._stride, self._n, self._cache = blob, stride, n, {}

    def __len__(self):
        return self._n

    def __getitem__(self, k):
        k = int(k)
        if k < 0 or k >= self._n:
            raise IndexError(k)
        if k not in self._cache:
            head = self._blob[k * self._stride:k * self._stride + 8].cpu().numpy().view(np.uint32)
            raw = self._blob[k * self._stride:k * self._stride + int(head[1])].cpu().numpy().tobytes()
            self._cache[k] = MapTables(raw)
        return self._cache[k]

    def __iter__(self):
        return (self[k] for k in range(self._n))

    def exit_slots(self, ent_cap, env_map):
        """Per env of `env_map` (device int64): entity id / dyn slot of the last exit switch of its level and the dyn slot
        of that switch's door, -1 if none -- torch ops over the blobs' headers and NppStaticEnt tables in HBM (map_blob.h:
        n_ent at byte 8, off_ents at byte 72; per 32-byte record kind at +16, dyn i16 at +20, aux i16 at +22)."""
        dev = self._blob.device
        B = self._blob[:self._n * self._stride].view(self._n, self._stride)

        def u32(off):
            w = B[:, off:off + 4].to(torch.int64)
            return w[:, 0] | (w[:, 1] << 8) | (w[:, 2] << 16) | (w[:, 3] << 24)

        def i16(pos):                                              # little-endian int16 at byte positions pos [M]
            lo = B.gather(1, pos[:, None]).to(torch.int64)[:, 0]; hi = B.gather(1, pos[:, None] + 1).to(torch.int64)[:, 0]
            v = lo | (hi << 8)
            return torch.where(v >= 32768, v - 65536, v)

        n_ent, off_ents = u32(8), u32(72)
        e = torch.arange(max(1, int(ent_cap)), device=dev, dtype=torch.int64)[None, :]
        valid = e < n_ent[:, None]
        pos = (off_ents[:, None] + 32 * e + 16).clamp(max=self._stride - 1)
        kind = B.gather(1, pos)
        is_sw = valid & (kind == FD.K_EXIT_SWITCH)
        sw_id = torch.where(is_sw, e, torch.full_like(e, -1)).max(dim=1).values          # the LAST one (nplay_headless.py:546-564)
        base = off_ents + 32 * sw_id.clamp(min=0)
        sw_slot = i16(base + 20)
        door = i16(base + 22)
        dr_slot = i16(off_ents + 32 * door.clamp(min=0) + 20)
        none = sw_id < 0
        sw_slot = torch.where(none, torch.full_like(sw_slot, -1), sw_slot)
        dr_slot = torch.where(none | (door < 0), torch.full_like(dr_slot, -1), dr_slot)
        em = env_map.to(torch.int64)
        return sw_id[em], sw_slot[em], dr_slot[em]


def _entity_rows(t, fl, df, di, order_mode, reloads=0):
    """Rebuild the reference's per-entity attributes from one env's flag / slot arrays.  `index` is the entity's index in the
    level plus, per Simulator.reset() since the load, the number of entities of its type (entities.py:213-216)."""
    E = t.n_ent
    per_type = np.bincount(np.asarray(t.ents["type"][:E], dtype=np.int64), minlength=40) if E else np.zeros(40, dtype=np.int64)
    ef = np.zeros((E, 4)); ei = np.zeros((E, 8), dtype=np.int64)
    cell = np.zeros(E, dtype=np.int64); stamp = np.zeros(E, dtype=np.int64); in_grid = np.ones(E, dtype=bool)
    for e in range(E):
        s = t.ents[e]
        kind, slot, f = int(s["kind"]), int(s["dyn"]), int(fl[e])
        x, y, a0, a1, s0, s1 = float(s["x"]), float(s["y"]), 0.0, 0.0, 0, 0
        c = int(s["cell"]); st = e if order_mode else int(s["load_order"])
        active = 1 if f & FD.FL_ACTIVE else 0
        roaming = slot >= 0 and kind != FD.K_DOOR_REGULAR
        if roaming:
            x, y = df[slot * 4], df[slot * 4 + 1]
            c = int(di[slot * 3]); st = int(di[slot * 3 + 1])
        if kind == FD.K_MINE:
            s0 = f & FD.FL_MINE
        elif kind == FD.K_EXIT:
            s0 = 1 if f & FD.FL_CLOSED else 0
            if c < 0:
                in_grid[e] = False
                c = int(di[slot * 3 + 2])
        elif kind == FD.K_DOOR_REGULAR:
            s0 = 1 if f & FD.FL_CLOSED else 0
            s1 = int(di[slot * 3 + 2]) & 255
        elif kind in (FD.K_DOOR_LOCKED, FD.K_DOOR_TRAP):
            s0 = 1 if f & FD.FL_CLOSED else 0
        elif kind in (FD.K_BOUNCE, FD.K_DEATH_BALL):
            a0, a1 = df[slot * 4 + 2], df[slot * 4 + 3]
        elif kind in (FD.K_DRONE_ZAP, FD.K_MINI_DRONE):
            s0 = (f & FD.FL_ST) >> FD.FL_ST_SHIFT
            a0, a1 = df[slot * 4 + 2], df[slot * 4 + 3]
        elif kind == FD.K_THWUMP:
            s0 = (f & FD.FL_ST) >> FD.FL_ST_SHIFT
            s0 = -1 if s0 == 3 else s0
        elif kind == FD.K_BOOST:
            s0 = 1 if f & FD.FL_TOUCH else 0
        elif kind == FD.K_SHWUMP:
            s0 = (f & FD.FL_ST) >> FD.FL_ST_SHIFT
            s1 = 1 if f & FD.FL_ACTIVATED else 0
            a0, a1 = df[slot * 4 + 2], df[slot * 4 + 3]
        cell[e], stamp[e] = c, st
        ef[e] = (x, y, a0, a1)
        ei[e, :7] = (int(s["type"]), int(s["index"]) + reloads * int(per_type[int(s["type"])]), active, c // 25, c % 25, s0, s1)
    for e in range(E):
        if not in_grid[e]:
            ei[e, 7] = -1
        else:
            same = in_grid & (cell == cell[e])
            ei[e, 7] = int(np.count_nonzero(stamp[same] < stamp[e]))
    return ef, ei
