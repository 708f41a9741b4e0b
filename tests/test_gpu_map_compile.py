"""Device map compiler (nppsim_compile_maps_device) against the host compiler, whose tables
tests/test_capi.py pins to the oracle / reference.  Bar: the two blobs are byte-identical, error codes equal."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def host_compile(L, md):
    md = np.ascontiguousarray(md, dtype=np.float64)
    cap = int(L.nppsim_map_blob_bound(len(md)))
    out = np.zeros(cap, dtype=np.uint8)
    used = C.c_int64(0); ne = C.c_int32(0); nd = C.c_int32(0)
    rc = L.nppsim_compile_map(md.ctypes.data, len(md), out.ctypes.data, cap, C.byref(used), C.byref(ne), C.byref(nd))
    return rc, out[:used.value].copy(), ne.value, nd.value


def device_compile(L, maps, explicit_len=True):
    import torch
    M = len(maps)
    lens = np.array([len(m) for m in maps], dtype=np.int32)
    Lrow = max(1235, int(lens.max()))
    arr = np.zeros((M, Lrow), dtype=np.float64)
    for k, m in enumerate(maps):
        arr[k, :len(m)] = m
    md = torch.from_numpy(arr).cuda()
    ml = torch.from_numpy(lens).cuda()
    stride = (int(L.nppsim_map_blob_bound(Lrow)) + 255) & ~255
    blobs = torch.full((M * stride,), 0xAB, dtype=torch.uint8, device="cuda")       # poison: every byte of a blob must be written
    off = torch.zeros(M, dtype=torch.int64, device="cuda")
    status = torch.full((M,), 77, dtype=torch.int32, device="cuda")
    summary = torch.zeros(4, dtype=torch.int32, device="cuda")
    rc = L.nppsim_compile_maps_device(md.data_ptr(), Lrow, ml.data_ptr() if explicit_len else None, Lrow, M, blobs.data_ptr(), stride,
                                      off.data_ptr(), status.data_ptr(), summary.data_ptr(), None)
    assert rc == 0, L.nppsim_last_error()
    torch.cuda.synchronize()
    assert np.array_equal(off.cpu().numpy(), np.arange(M) * stride)
    return blobs.cpu().numpy().reshape(M, stride), status.cpu().numpy(), summary.cpu().numpy().view(np.uint32)


def fuzz_map(rng, hard):
    """Random level in the map_loader layout: any tile id 0..37 (plus junk when hard), every entity kind the factory
    instantiates, unknown types, doors with both record strides, non-integer coordinates."""
    n_exit = int(rng.integers(0, 3))
    recs = []
    for _ in range(n_exit):
        recs.append([3, rng.integers(4, 170), rng.integers(4, 96), 0, 0])
    for _ in range(n_exit):
        recs.append([4, rng.integers(4, 170), rng.integers(4, 96), 0, 0])
    kinds = [1, 21, 2, 5, 6, 8, 10, 11, 14, 26, 17, 20, 24, 25, 28, 7, 9, 30]
    for _ in range(int(rng.integers(0, 60))):
        t = int(rng.choice(kinds))
        x, y = float(rng.integers(0, 176)), float(rng.integers(0, 100))
        if rng.random() < 0.2:
            x += float(rng.random()); y -= float(rng.random())
        o, mode = int(rng.integers(0, 8)), int(rng.integers(0, 4))
        if t in (6, 8):
            tail = [0, rng.integers(4, 170), rng.integers(0, 96), 0] + ([0] if rng.random() < 0.5 else [])
            recs.append([t, x, y, o, mode] + tail)
        else:
            recs.append([t, x, y, o, mode])
    if hard and rng.random() < 0.5 and recs:           # an out-of-range orientation / mode somewhere
        r = recs[int(rng.integers(0, len(recs)))]
        r[3] = float(rng.choice([8, 9, -1, 2.5]))
    flat = [v for r in recs for v in r]
    if rng.random() < 0.3:
        flat += [1, 5][: int(rng.integers(0, 3))]      # ragged tail (incomplete record)
    md = np.zeros(1235 + len(flat))
    tiles = rng.integers(0, 38, size=966)
    tiles[rng.random(966) < 0.5] = 0
    md[184:1150] = tiles
    if hard:
        md[184 + rng.integers(0, 966, size=5)] = rng.choice([2.5, -1, 40, 1e9, 33.0000001], size=5)
    md[1156] = n_exit
    md[1200] = float(rng.integers(0, 5))
    md[1231], md[1232] = float(rng.integers(4, 170)), float(rng.integers(4, 96))
    md[1235:] = flat
    return md


def _all_maps(golden_replays, golden_generated):
    maps = [golden_replays.traj(i)["map_data"] for i in range(len(golden_replays))]
    maps += [golden_generated.traj(i)["map_data"] for i in range(len(golden_generated))]
    import bench
    for fx in ("simplest_with_mines", "locked_doors"):
        arr, lens = bench.load_maps(fx)
        maps += [arr[k, :lens[k]] for k in range(0, len(arr), 37)]
    return maps


def _compare(L, maps, explicit_len=True):
    blobs, status, summary = device_compile(L, maps, explicit_len)
    n_bad = 0; max_ent = 0; max_dyn = 0; kinds = 0
    for k, m in enumerate(maps):
        rc, ref, ne, nd = host_compile(L, m)
        assert status[k] == rc, f"level {k}: device status {status[k]}, host {rc}"
        if rc:
            n_bad += 1
            assert blobs[k, :4].view(np.uint32)[0] == 0
            continue
        got = blobs[k, :len(ref)]
        if not np.array_equal(got, ref):
            diff = np.nonzero(got != ref)[0]
            raise AssertionError(f"level {k}: {len(diff)} bytes differ, first at offset {diff[0]} (blob {len(ref)} bytes)")
        max_ent = max(max_ent, ne); max_dyn = max(max_dyn, nd)
        kinds |= int(ref[100:104].view(np.uint32)[0])
    assert tuple(int(v) for v in summary) == (max_ent, max_dyn, kinds, n_bad)
    return n_bad


def test_device_blobs_identical_to_host(golden_replays, golden_generated):
    from nclone_b200._lib import lib
    maps = _all_maps(golden_replays, golden_generated)
    assert len(maps) > 400
    assert _compare(lib(), maps) == 0


def test_device_compile_edge_cases_and_errors():
    from nclone_b200._lib import lib
    L = lib()
    empty = np.zeros(1238); empty[1231] = 22; empty[1232] = 22                          # ragged tail, no entities
    glitch = np.zeros(1245); glitch[1231] = 22; glitch[1232] = 22; glitch[184:200] = 35; glitch[1235] = 99
    bad_pad = np.zeros(1245); bad_pad[1235:1240] = [10, 20, 20, 9, 0]                   # KeyError in the reference
    bad_exit = np.zeros(1245); bad_exit[1235:1240] = [3, 20, 20, 0, 0]; bad_exit[1156] = 5     # IndexError in the reference
    bad_cnt = np.zeros(1245); bad_cnt[1156] = 0.5
    bad_drone = np.zeros(1245); bad_drone[1235:1240] = [14, 20, 20, 0, 7]
    short_door = np.zeros(1241); short_door[1235:1241] = [6, 20, 20, 0, 0, 0]
    maps = [empty, glitch, bad_pad, bad_exit, bad_cnt, bad_drone, short_door]
    assert _compare(L, maps) == 5
    # default_len path (map_len == NULL): rows of one length
    assert _compare(L, [glitch, bad_pad, np.zeros(1245)], explicit_len=False) == 1
    # too short a row is an error, not a crash
    import torch
    md = torch.zeros((2, 1300), dtype=torch.float64, device="cuda")
    ml = torch.tensor([100, 1300], dtype=torch.int32, device="cuda")
    stride = (int(L.nppsim_map_blob_bound(1300)) + 255) & ~255
    blobs = torch.zeros(2 * stride, dtype=torch.uint8, device="cuda")
    status = torch.zeros(2, dtype=torch.int32, device="cuda"); summary = torch.zeros(4, dtype=torch.int32, device="cuda")
    assert L.nppsim_compile_maps_device(md.data_ptr(), 1300, ml.data_ptr(), 1300, 2, blobs.data_ptr(), stride, None, status.data_ptr(), summary.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert status.tolist() == [-2, 0] and summary[3].item() == 1
    # argument errors
    assert L.nppsim_compile_maps_device(None, 1300, None, 1300, 2, blobs.data_ptr(), stride, None, status.data_ptr(), summary.data_ptr(), None) == -1
    assert L.nppsim_compile_maps_device(md.data_ptr(), 1300, None, 1300, 2, blobs.data_ptr(), 1000, None, status.data_ptr(), summary.data_ptr(), None) == -1


def test_device_compile_fuzz():
    from nclone_b200._lib import lib
    rng = np.random.default_rng(20251104)
    maps = [fuzz_map(rng, hard=False) for _ in range(600)] + [fuzz_map(rng, hard=True) for _ in range(400)]
    n_bad = _compare(lib(), maps)
    assert 20 < n_bad < 400          # the hard half really exercises the error paths


def test_capacity_is_an_error():
    from nclone_b200._lib import lib
    md = np.zeros(1235 + 5 * 1100); md[1231] = 22; md[1232] = 22
    md[1235:] = np.tile([2, 20, 20, 0, 0], 1100)       # 1,100 gold > NPPSIM_DEVICE_COMPILE_MAX_ENT
    blobs, status, summary = device_compile(lib(), [md])
    assert status[0] == -3 and summary[3] == 1
    md2 = md[:1235 + 5 * 1024]
    assert _compare(lib(), [md2]) == 0                  # exactly at capacity


def test_load_maps_device_equals_load_maps(golden_generated):
    """Same levels through both loaders, same random actions: identical device state after a rollout."""
    import torch
    from nclone_b200 import BatchedSimulator
    maps = [golden_generated.traj(i)["map_data"] for i in range(len(golden_generated))]
    N = 4 * len(maps)
    a = BatchedSimulator(N); a.load_maps(maps, np.arange(N) % len(maps))
    Lrow = max(len(m) for m in maps)
    arr = np.zeros((len(maps), Lrow)); lens = np.zeros(len(maps), dtype=np.int32)
    for k, m in enumerate(maps):
        arr[k, :len(m)] = m; lens[k] = len(m)
    b = BatchedSimulator(N); b.load_maps_device(torch.from_numpy(arr).cuda(), lens)
    assert a.kind_mask == b.kind_mask and a.dyn_cap == b.dyn_cap and a.ent_cap == b.ent_cap
    gen = torch.Generator(device="cuda"); gen.manual_seed(5)
    for _ in range(100):
        act = torch.randint(0, 6, (N,), generator=gen, device="cuda", dtype=torch.uint8)
        a.step(act); b.step(act)
        a.reset_terminal(500); b.reset_terminal(500)
    for name in ("nin_f", "nin_i", "ent_flag", "dyn_f", "dyn_i"):
        x, y = getattr(a, name), getattr(b, name)
        assert torch.equal(x.view(torch.uint8), y.view(torch.uint8)), name
    assert torch.equal(a.env_i[:5], b.env_i[:5])
    # the exit getters of a device-compiled batch read the entity tables in HBM: same answers, no blob pulled to the host
    for fa, fb in ((a.exit_switch_position(), b.exit_switch_position()), (a.exit_door_position(), b.exit_door_position())):
        assert torch.equal(fa[0], fb[0]) and torch.equal(fa[1], fb[1])
    assert torch.equal(a.exit_switch_activated(), b.exit_switch_activated())
    assert a.exit_switch_position()[0].abs().sum() > 0 and len(b.maps._cache) == 0
    ra, rb = a.canonical_state([0, 7, N - 1]), b.canonical_state([0, 7, N - 1])
    from oracle import schema
    assert [schema.state_hash(*r) for r in ra] == [schema.state_hash(*r) for r in rb]
    with pytest.raises(ValueError):
        arr[3, 1235:1240] = [10, 20, 20, 9, 0]
        BatchedSimulator(8).load_maps_device(torch.from_numpy(arr).cuda(), lens)
