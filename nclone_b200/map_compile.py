"""Host-side map compilation: map_data (map_loader format) -> blob (csrc/map_blob.h).

Wraps nppsim_compile_map; also parses a blob back into numpy views so that host code
(getters, parity tests) can read the static tables without a GPU."""
import ctypes as C
import hashlib

import numpy as np

from ._lib import lib, check

HEADER = np.dtype([
    ("magic", "<u4"), ("total_bytes", "<u4"), ("n_ent", "<u4"), ("n_dyn", "<u4"), ("n_doors", "<u4"),
    ("n_static_cell", "<u4"), ("n_segments", "<u4"), ("flags", "<u4"),
    ("spawn_x", "<f8"), ("spawn_y", "<f8"), ("db_count", "<f8"),
    ("off_cell_hdr", "<u4"), ("off_segs", "<u4"), ("off_hedge", "<u4"), ("off_vedge", "<u4"), ("off_ents", "<u4"),
    ("off_ecell_off", "<u4"), ("off_ecell_ids", "<u4"), ("off_ecell_ids_dic", "<u4"), ("off_dyn_ent", "<u4"),
    ("off_doors", "<u4"), ("off_tiles", "<u4"), ("kind_mask", "<u4"),
])
STATIC_ENT = np.dtype([
    ("x", "<f8"), ("y", "<f8"), ("kind", "u1"), ("type", "u1"), ("p0", "u1"), ("p1", "u1"),
    ("dyn", "<i2"), ("aux", "<i2"), ("cell", "<u2"), ("index", "<u2"), ("load_order", "<u2"), ("pad", "<u2"),
])
DOOR = np.dtype([("ent", "<u2"), ("vertical", "u1"), ("x0", "u1"), ("y0", "u1"), ("x1", "u1"), ("y1", "u1"), ("pad", "u1")])
MAGIC = 0x4E505042
assert HEADER.itemsize == 104 and STATIC_ENT.itemsize == 32 and DOOR.itemsize == 8


def compile_map(map_data):
    """-> bytes blob.  Raises ValueError on malformed map_data (the reference raises
    KeyError / IndexError / ValueError from the entity constructors)."""
    md = np.ascontiguousarray(np.asarray(map_data, dtype=np.float64))
    L = lib()
    cap = int(L.nppsim_map_blob_bound(len(md)))
    out = np.zeros(cap, dtype=np.uint8)
    used = C.c_int64(0); n_ent = C.c_int32(0); n_dyn = C.c_int32(0)
    rc = L.nppsim_compile_map(md.ctypes.data, len(md), out.ctypes.data, cap, C.byref(used), C.byref(n_ent), C.byref(n_dyn))
    if rc:
        msg = L.nppsim_last_error()
        raise ValueError(f"map compile failed ({rc}): {msg.decode() if msg else '?'}")
    return out[:used.value].tobytes()


class MapTables:
    """numpy views over one blob."""

    def __init__(self, blob):
        self.blob = blob
        b = np.frombuffer(blob, dtype=np.uint8)
        self.h = np.frombuffer(blob, dtype=HEADER, count=1)[0]
        if int(self.h["magic"]) != MAGIC:
            raise ValueError("not a map blob")
        h = self.h
        self.n_ent, self.n_dyn, self.n_doors = int(h["n_ent"]), int(h["n_dyn"]), int(h["n_doors"])
        self.kind_mask = int(h["kind_mask"])
        self.cell_hdr = np.frombuffer(blob, "<u4", 1100, int(h["off_cell_hdr"]))
        self.segs = np.frombuffer(blob, "<u4", 8800, int(h["off_segs"])).reshape(1100, 8)
        self.hedge = np.frombuffer(blob, "<u4", 142, int(h["off_hedge"]))
        self.vedge = np.frombuffer(blob, "<u4", 142, int(h["off_vedge"]))
        self.ents = np.frombuffer(blob, STATIC_ENT, self.n_ent, int(h["off_ents"]))
        self.ecell_off = np.frombuffer(blob, "<u2", 1101, int(h["off_ecell_off"]))
        ns = int(h["n_static_cell"])
        self.ecell_ids = np.frombuffer(blob, "<u2", ns, int(h["off_ecell_ids"]))
        self.ecell_ids_dic = np.frombuffer(blob, "<u2", ns, int(h["off_ecell_ids_dic"]))
        self.dyn_ent = np.frombuffer(blob, "<u4", self.n_dyn, int(h["off_dyn_ent"]))
        self.doors = np.frombuffer(blob, DOOR, self.n_doors, int(h["off_doors"]))
        self.tiles = np.frombuffer(blob, "u1", 1100, int(h["off_tiles"]))
        del b

    def edge_bits(self, vertical):
        w = self.vedge if vertical else self.hedge
        bits = np.unpackbits(w.view(np.uint8), bitorder="little")[:89 * 51]
        return bits.reshape(89, 51)

    def decoded_segments(self):
        """Per cell (x-major), the reference-order list as rows
        (kind, a, b, c, d, convex): linear x1,y1,x2,y2 | arc cx,cy,hor,ver,convex; pixels."""
        rows = []
        counts = (self.cell_hdr & 15).astype(np.int32)
        for cell in range(1100):
            for k in range(counts[cell]):
                s = int(self.segs[cell, k])
                if s & 0x80000000:
                    rows.append((1, 12 * (s & 127), 12 * ((s >> 7) & 63), -1 if (s >> 13) & 1 else 1,
                                 -1 if (s >> 14) & 1 else 1, (s >> 15) & 1))
                else:
                    rows.append((0, 12 * (s & 127), 12 * ((s >> 7) & 63), 12 * ((s >> 13) & 127), 12 * ((s >> 20) & 63), 0))
        return counts, np.array(rows, dtype=np.float64).reshape(-1, 6)


def content_key(map_data):
    md = np.ascontiguousarray(np.asarray(map_data, dtype=np.float64))
    return hashlib.blake2b(md.tobytes(), digest_size=16).digest()
