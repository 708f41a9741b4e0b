"""C-ABI library: loads, exports every symbol include/nppsim.h declares, argument checking,
and the host-side map compiler (no GPU needed for any of this)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from nclone_b200._lib import lib, EXPORTS, NppsimState
from nclone_b200.map_compile import compile_map, MapTables
from nclone_b200 import fields as FD
from oracle import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    hdr = open(os.path.join(ROOT, "include", "nppsim.h")).read()
    declared = set(re.findall(r"\b(nppsim_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations found"
    assert declared == set(EXPORTS)
    L = lib()
    for name in declared:
        assert getattr(L, name) is not None
    assert L.nppsim_version() == 102
    assert L.nppsim_sq_mode() == 1


def test_constants_match_header():
    hdr = open(os.path.join(ROOT, "include", "nppsim.h")).read()
    defs = dict(re.findall(r"#define (NPPSIM_[A-Z_]+)\s+\(?(-?\d+)\)?", hdr))
    assert int(defs["NPPSIM_NF"]) == FD.NF == 31 and int(defs["NPPSIM_NI"]) == FD.NI == 27
    assert int(defs["NPPSIM_NENV"]) == FD.NENV
    assert int(defs["NPPSIM_ENV_TICKS"]) == FD.ENV_TICKS
    assert (int(defs["NPPSIM_EV_EXIT_SWITCH"]), int(defs["NPPSIM_EV_DEATH"])) == (FD.EV_EXIT_SWITCH, FD.EV_DEATH)
    from oracle import schema
    assert schema.NINJA_F == FD.NINJA_F and schema.NINJA_I == FD.NINJA_I


def test_argument_errors_without_gpu():
    L = lib()
    assert L.nppsim_reset(None, None, 0, None) == -1
    assert b"NULL" in L.nppsim_last_error()
    st = NppsimState()
    assert L.nppsim_step(C.byref(st), None, None, 1, 4, None, 0, None, None) == -1
    st.n_envs = 4
    assert L.nppsim_step_actions(C.byref(st), None, 4, 0, None, None) == -1


def _maps(G, n):
    return [G.traj(i)["map_data"] for i in range(0, len(G), max(1, len(G) // n))]


def test_map_compiler_matches_oracle_tables(golden_replays, golden_generated):
    """Segment lists per cell (order included), grid-edge tables and entity counts against the
    oracle's own loader, which is pinned to the reference."""
    for md in _maps(golden_replays, 40) + _maps(golden_generated, 20):
        t = MapTables(compile_map(md))
        s = oracle.OracleSim(); s.load(md)
        counts, present, segs, hor, ver = s.static_tables()
        c2, rows = t.decoded_segments()
        assert np.array_equal(counts, c2)
        assert np.array_equal(segs[:, :6], rows)
        hb = t.edge_bits(False).astype(np.int32).copy(); vb = t.edge_bits(True).astype(np.int32).copy()
        for d in t.doors:
            if t.ents[d["ent"]]["kind"] in (FD.K_DOOR_REGULAR, FD.K_DOOR_LOCKED):      # closed at load
                tgt = vb if d["vertical"] else hb
                tgt[d["x0"], d["y0"]] += 1; tgt[d["x1"], d["y1"]] += 1
        assert np.array_equal(hb.reshape(-1), hor) and np.array_equal(vb.reshape(-1), ver)
        assert t.n_ent == s.num_entities
        _, _, ef, ei = s.state()
        assert np.array_equal(t.ents["type"], ei[:, 0]) and np.array_equal(t.ents["index"], ei[:, 1])
        assert np.array_equal(t.ents["x"], ef[:, 0]) and np.array_equal(t.ents["y"], ef[:, 1])
        assert np.array_equal(t.ents["cell"] // 25, ei[:, 3]) and np.array_equal(t.ents["cell"] % 25, ei[:, 4])


def test_map_compiler_rejects_malformed():
    with pytest.raises(ValueError):
        compile_map(np.zeros(100))
    md = np.zeros(1245); md[1231] = 10; md[1232] = 10
    md[1235:1240] = [10, 20, 20, 9, 0]        # launch pad with orientation 9 -> KeyError in the reference
    with pytest.raises(ValueError):
        compile_map(md)
    md[1235:1240] = [3, 20, 20, 0, 0]; md[1156] = 5      # exit door whose switch record lies past the end -> IndexError
    with pytest.raises(ValueError):
        compile_map(md)


def test_map_compiler_edge_cases():
    # empty level: only the border tiles; ragged tail (incomplete record) is ignored (map_loader.py:90-92)
    md = np.zeros(1238); md[1231] = 22; md[1232] = 22
    t = MapTables(compile_map(md))
    assert t.n_ent == 0 and t.n_dyn == 0
    s = oracle.OracleSim(); s.load(md)
    assert np.array_equal((t.cell_hdr & 15), s.static_tables()[0])
    # unknown entity types and glitched tiles (34..37) are skipped
    md = np.zeros(1245); md[1231] = 22; md[1232] = 22; md[184:200] = 35; md[1235] = 99
    t = MapTables(compile_map(md))
    assert t.n_ent == 0 and (t.tiles.reshape(44, 25)[1:17, 1] == 0).all()


def test_map_compiler_capacity_is_an_error_not_truncation():
    """include/nppsim.h: "capacity exceeded: an error at load, never truncation".  NppStaticEnt.dyn / .aux are int16, so a
    level with more than 32,767 entities must be refused (NPPSIM_E_CAPACITY), not wrapped to negative slot numbers."""
    n = 32768
    md = np.zeros(1235 + 5 * n); md[1231] = 22; md[1232] = 22
    rec = md[1235:].reshape(n, 5)
    rec[:, 0] = 17; rec[:, 1] = 40 + (np.arange(n) % 100); rec[:, 2] = 40 + (np.arange(n) // 100 % 60)     # bounce blocks: one slot each
    with pytest.raises(ValueError, match=r"\(-3\).*entities"):
        compile_map(md)
    t = MapTables(compile_map(md[:1235 + 5 * 32767]))
    assert t.n_ent == 32767 and t.n_dyn == 32767 and int(t.ents["dyn"].min()) == 0 and int(t.ents["dyn"].max()) == 32766


def test_state_stride_is_validated():
    L = lib()
    st = NppsimState()
    st.n_envs, st.stride = 64, 32
    for name in ("nin_f", "nin_i", "env_i", "env_map", "map_blob", "map_off"):
        setattr(st, name, 0x1000)                       # never dereferenced: check_state runs before any launch
    assert L.nppsim_reset(C.byref(st), None, 0, None) == -1
    assert b"stride" in L.nppsim_last_error()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from nclone_b200 import BatchedSimulator
    with pytest.raises(RuntimeError):
        BatchedSimulator(4)
