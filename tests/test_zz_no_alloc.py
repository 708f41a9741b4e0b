"""nppsim_step_host_static -- the host-buffer step for callers that cannot allocate device memory (a binding with no CUDA
runtime of its own, a process whose card is already parcelled out to another framework's pool, a loop that must not stall
the device on cudaMalloc) -- against nppsim_step_host, which takes its two scratch buffers from the caller.

The file sorts last on purpose: it is the only test that depends on a process-wide device array keeping state between
calls, and `pytest -x` would hide everything collected after a failure here.  It must pass on its own in a fresh process
(then its first call is also the first step launch of the process: the entry point's own per-device set-up runs before the
ordinary launcher's) and after the rest of the suite (then the array and the kernels have been used by others before).
Bar, as everywhere in this suite: identical bytes."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _static_step(sim, actions, block, n_ticks, with_obs):
    """The raw C-ABI call, nothing of the Python wrapper in between."""
    import torch
    with torch.cuda.device(sim.device):
        return sim.L.nppsim_step_host_static(C.byref(sim._state), C.c_void_p(actions.ctypes.data), n_ticks, 1,
                                             C.c_void_p(block.data_ptr()), int(with_obs), sim._stream())


def test_static_scratch_step_matches_step_host(golden_replays, golden_generated):
    import torch
    from nclone_b200 import BatchedSimulator
    from nclone_b200._lib import NppsimState
    # three levels interleaved over 96 envs (state columns permuted: ext_of); replay level 1 has drones, thwumps, death balls ...
    # (13 entity kinds), so this batch runs on the full step kernel
    maps = [golden_replays.traj(1)["map_data"], golden_generated.traj(0)["map_data"], golden_generated.traj(40)["map_data"]]
    N = 96
    e2m = np.arange(N) % 3
    rng = np.random.default_rng(7)
    a = rng.integers(0, 6, size=(8, N)).astype(np.uint8)
    new = BatchedSimulator(N); new.load_maps(maps, e2m)
    ref = BatchedSimulator(N); ref.load_maps(maps, e2m)
    assert new.ext_of is not None and not new._is_basic(new.kind_mask)
    # a second, smaller batch of curriculum levels only (basic kernel, caller order, N not a multiple of 32) stepped in between:
    # the array then holds a DIFFERENT layout from the call before, which is what stale contents would show up against
    M = 40
    small_maps = [golden_generated.traj(k)["map_data"] for k in (0, 40)]
    b = rng.integers(0, 6, size=(8, M)).astype(np.uint8)
    new_s = BatchedSimulator(M, placement="caller"); new_s.load_maps(small_maps, np.arange(M) % 2)
    ref_s = BatchedSimulator(M, placement="caller"); ref_s.load_maps(small_maps, np.arange(M) % 2)
    assert new_s.ext_of is None and new_s._is_basic(new_s.kind_mask)

    blk_new, _ = new.host_block(with_obs=True)
    blk_ref, v_ref = ref.host_block(with_obs=True)
    blk_new_s = torch.zeros(int(new_s.L.nppsim_host_block_bytes(M, 1)), dtype=torch.uint8)      # pageable on this side
    blk_ref_s, _ = ref_s.host_block(with_obs=True)

    # calls 1 and 2, back to back on the same device array, each checked against the caller-buffer entry point
    for k in range(2):
        assert _static_step(new, a[k], blk_new, 4, True) == 0, new.L.nppsim_last_error()
        ref.step_host(a[k], blk_ref, 4, True, with_obs=True)
        assert np.array_equal(blk_new.numpy(), blk_ref.numpy()), f"call {k + 1}"
    assert len(np.unique(v_ref["pos_vel"][0])) > 3                # the levels really moved apart: equal blocks are not trivial
    # then alternating batch sizes, kernels and block kinds (with / without observations) on the one array
    for k in range(2, 8):
        with_obs = k % 2 == 0
        blk_new_s.zero_(); blk_ref_s.zero_()
        assert _static_step(new_s, b[k], blk_new_s, 4, with_obs) == 0, new_s.L.nppsim_last_error()
        ref_s.step_host(b[k], blk_ref_s, 4, True, with_obs=with_obs)
        assert np.array_equal(blk_new_s.numpy(), blk_ref_s.numpy()), k     # without observations the tail stays zero on both sides
        blk_new.zero_(); blk_ref.zero_()
        new.step_host_static(a[k], blk_new, 4, True, with_obs=not with_obs)   # the wrapper a Python user calls
        ref.step_host(a[k], blk_ref, 4, True, with_obs=not with_obs)
        assert np.array_equal(blk_new.numpy(), blk_ref.numpy()), k
    # the device state behind the blocks is the same too, bit for bit
    for x, y in ((new, ref), (new_s, ref_s)):
        assert torch.equal(x.nin_f.view(torch.int64), y.nin_f.view(torch.int64))
        assert torch.equal(x.nin_i, y.nin_i) and torch.equal(x.ent_flag, y.ent_flag)
        assert torch.equal(x.dyn_f.view(torch.int64), y.dyn_f.view(torch.int64)) and torch.equal(x.dyn_i, y.dyn_i)

    # a batch the array cannot hold is refused before anything runs: the state and the block stay as they are
    big = NppsimState()
    C.memmove(C.byref(big), C.byref(new._state), C.sizeof(NppsimState))
    big.n_envs = new.STATIC_MAX_ENVS + 1
    before = new.nin_f.clone()
    blk_new.fill_(0xAB)
    launches = new.launch_count()
    with torch.cuda.device(new.device):
        rc = new.L.nppsim_step_host_static(C.byref(big), C.c_void_p(a[0].ctypes.data), 4, 1, C.c_void_p(blk_new.data_ptr()), 1, new._stream())
    assert rc == -3 and b"NPPSIM_STATIC_MAX_ENVS" in new.L.nppsim_last_error()
    torch.cuda.synchronize()
    assert new.launch_count() == launches and torch.equal(new.nin_f, before) and bool((blk_new == 0xAB).all())
