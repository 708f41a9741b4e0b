"""tests/golden/doors.npz -- goldens for the entity rows that no reference replay exercises, plus the reference's own
scenario maps.  Produced by the UNMODIFIED reference (imported read-only from /root/reference), build container only:

    python -m oracle.gen_golden_doors

No file under /root/reference contains a regular door (type 5), and trap doors / shove thwumps appear in one replay
each, so these levels are synthesised the way a level editor would: a 5-value record `[5, x, y, orientation, 0]`
(entity_factory.py:200-203), a 10-value trap-door record `[8, x, y, orientation, 0, 9, sx, sy, 0, 0]`
(map_loader.py:126-141, entity_factory.py:204-216) or `[28, x, y, 0, 0]` appended to a real level, placed ON the path the
ninja walks under a seeded random policy (found with a first run of the reference on the unmodified level) or on the
path of a drone / thwump of that level, so that the door's grid-edge toggling (entity_door_base.py:70-108) changes
what is_empty_row / is_empty_column sees.

Groups (names prefix):
  door5/...     injected regular doors, orientations 0..7, on the ninja's path and on drone paths
  trap8/...     injected trap doors whose switch lies on the ninja's path
  shwump28/...  injected shove thwumps on the ninja's path
  testmap/...   the 17 scenario maps of nclone/test_maps/ under random held actions
  official/...  the 5 maps of nclone/maps/test-maps/
  tmp/...       the 8 bc_replays_tmp replays: their own inputs and random held actions
Same per-frame record as the other goldens (oracle/gen_golden.pack)."""
import glob
import os
import random
import sys
import zlib

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import schema, refdump  # noqa: E402
from oracle.fuzz_vs_ref import generated_map, random_inputs  # noqa: E402
from oracle.gen_golden import pack, OUT  # noqa: E402

T_EP = 320


def _entity_end(md):
    """Index just behind the last non-empty entity record (records parsed like map_loader.py:84-141)."""
    i, end = 1230, 1235
    while i + 4 < len(md):
        t = md[i]
        step = 5
        if t in (6, 8):
            step = 10 if (i + 9 < len(md) and md[i + 7] != 0 and md[i + 8] == 0 and md[i + 9] == 0) else 9
        if t != 0:
            end = i + step
        i += step
    return end


def inject(md, records):
    md = list(md)                       # entity types index lists in the reference: ints stay ints
    end = min(_entity_end(md), len(md))
    out = md[:end]
    for r in records:
        out += [int(v) for v in r]
    return out


def units(px):
    """pixel -> map units (x6), on the quarter-tile lattice the editor uses"""
    return int(round(px / 6.0))


def path_point(rng, rec, lo=15, hi=None):
    T = len(rec["nin_f"]) - 1
    alive = np.nonzero(rec["nin_i"][:, 0] < 6)[0]
    hi = min(hi or T, int(alive.max()) if len(alive) else T)
    t = rng.randrange(min(lo, hi), hi + 1)
    return float(rec["nin_f"][t, 0]), float(rec["nin_f"][t, 1]), t


def drone_paths(md, hor, jump):
    """positions over time of drones / thwumps of the unmodified level (entity rows of the full record)"""
    rec = refdump.run_trajectory(md, hor, jump, full=True)
    types = rec["ent_i"][0][:, 0]
    rows = [k for k, t in enumerate(types) if t in (14, 26, 20)]
    return rec, [(rec["ent_f"][:, k, 0], rec["ent_f"][:, k, 1]) for k in rows]


def door_events(rec, kind_type):
    """per entity row of `kind_type`: number of closed 1->0 and 0->1 transitions over the episode"""
    types = rec["ent_i"][0][:, 0]
    ev = []
    for k in np.nonzero(types == kind_type)[0]:
        s = rec["ent_i"][:, k, 5]
        ev.append((int(((s[:-1] == 1) & (s[1:] == 0)).sum()), int(((s[:-1] == 0) & (s[1:] == 1)).sum())))
    return ev


def main():
    refdump.import_reference()
    from nclone.map_generation.generator_configs import CATEGORIES  # noqa
    files = sorted(glob.glob(os.path.join(refdump.REF_ROOT, "bc_replays", "*.replay")))
    replay_maps = {os.path.basename(f)[:-7]: list(schema.parse_replay(open(f, "rb").read())[0]) for f in files}

    def has(md, types):
        return any(md[i] in types for i in range(1235, len(md) - 4, 5))
    movers = [n for n, md in replay_maps.items() if has(md, (14, 26, 20))]
    plain = [n for n in replay_maps if n not in movers]
    trajs = []
    stats = {"door_reclosed": 0, "door_opened": 0, "door_on_mover_path": 0, "trap_closed": 0, "shwump_activated": 0}

    # ---- (a) regular doors ----------------------------------------------------------------------------------
    bases = [("rp/" + n, replay_maps[n]) for n in movers[:14]] + [("rp/" + n, replay_maps[n]) for n in plain[:6]]
    bases += [(f"gen/{cat}/{52000 + s}", generated_map(cat, 52000 + s)) for cat in list(CATEGORIES)[:6] for s in range(1)]
    for bi, (bname, md) in enumerate(bases):
        for ep in range(2):
            rng = random.Random(zlib.crc32(f"door5:{bname}:{ep}".encode()))
            hor, jump = random_inputs(rng, T_EP)
            base, paths = drone_paths(md, hor, jump)
            recs = []
            for d in range(3):                                  # up to three doors on the ninja's path, all orientations over the set
                x, y, _ = path_point(rng, base)
                recs.append([5, units(x), units(y), (bi + ep * 3 + d * 2 + rng.randrange(2)) % 8, 0])
            on_mover = 0
            for px, py in paths[:2]:                            # and doors across the way of the level's drones / thwumps
                t = rng.randrange(10, T_EP)
                recs.append([5, units(float(px[t])), units(float(py[t])), rng.choice((0, 2, 4, 6)), 0])
                on_mover += 1
            md2 = inject(md, recs)
            rec = refdump.run_trajectory(md2, hor, jump, full=True)
            ev = door_events(rec, 5)
            stats["door_opened"] += int(any(o > 0 for o, c in ev))
            stats["door_reclosed"] += int(any(o > 0 and c > 0 for o, c in ev))
            stats["door_on_mover_path"] += int(on_mover > 0)
            trajs.append(dict(name=f"door5/{bname}/{ep}", map_data=np.asarray(md2, dtype=np.float64), hor=hor, jump=jump, rec=rec))

    # ---- (b) trap doors: switch on the ninja's path --------------------------------------------------------
    bases_t = [("rp/" + n, replay_maps[n]) for n in (movers[14:20] + plain[6:12])]
    bases_t += [(f"gen/{cat}/{53000}", generated_map(cat, 53000)) for cat in list(CATEGORIES)[:4]]
    for bi, (bname, md) in enumerate(bases_t):
        rng = random.Random(zlib.crc32(f"trap8:{bname}".encode()))
        hor, jump = random_inputs(rng, T_EP)
        base, paths = drone_paths(md, hor, jump)
        recs = []
        for d in range(2):
            sx, sy, t = path_point(rng, base)
            if paths:                                           # the door itself across a drone's way: closing it changes the drone's route
                px, py = paths[d % len(paths)]
                tt = rng.randrange(t, T_EP + 1)
                dx, dy = float(px[tt]), float(py[tt])
            else:
                dx, dy, _ = path_point(rng, base, lo=t)
            recs.append([8, units(dx), units(dy), rng.choice((0, 2, 4, 6)), 0, 9, units(sx), max(1, units(sy)), 0, 0])
        md2 = inject(md, recs)
        rec = refdump.run_trajectory(md2, hor, jump, full=True)
        stats["trap_closed"] += int(any(c > 0 for o, c in door_events(rec, 8)))
        trajs.append(dict(name=f"trap8/{bname}", map_data=np.asarray(md2, dtype=np.float64), hor=hor, jump=jump, rec=rec))

    # ---- (c) shove thwumps on the ninja's path --------------------------------------------------------------
    bases_s = [("rp/" + n, replay_maps[n]) for n in (plain[12:20] + movers[20:24])]
    bases_s += [(f"gen/{cat}/{54000}", generated_map(cat, 54000)) for cat in list(CATEGORIES)[:4]]
    for bname, md in bases_s:
        rng = random.Random(zlib.crc32(f"shwump28:{bname}".encode()))
        hor, jump = random_inputs(rng, T_EP)
        base = refdump.run_trajectory(md, hor, jump)
        recs = []
        for d in range(3):
            x, y, t = path_point(rng, base, lo=25)
            # next to the path, not on it: the ninja must walk INTO the block (entity_shove_thwump.py:173-202)
            recs.append([28, units(x + rng.choice((-26, 26, 0, 0))), units(y + rng.choice((-26, 26, 0))), 0, 0])
        md2 = inject(md, recs)
        rec = refdump.run_trajectory(md2, hor, jump, full=True)
        types = rec["ent_i"][0][:, 0]
        rows = np.nonzero(types == 28)[0]
        stats["shwump_activated"] += int(any((rec["ent_i"][:, k, 5] > 0).any() for k in rows))
        trajs.append(dict(name=f"shwump28/{bname}", map_data=np.asarray(md2, dtype=np.float64), hor=hor, jump=jump, rec=rec))

    # ---- (d) the reference's own scenario maps --------------------------------------------------------------
    def raw_map(path):
        return list(open(path, "rb").read())
    scen = [("testmap/" + os.path.basename(p), raw_map(p)) for p in sorted(glob.glob(os.path.join(refdump.REF_ROOT, "nclone", "test_maps", "*")))
            if not p.endswith(".json")]
    scen += [("official/" + os.path.basename(p), raw_map(p)) for p in sorted(glob.glob(os.path.join(refdump.REF_ROOT, "nclone", "maps", "test-maps", "*")))]
    for name, md in scen:
        for ep in range(3):
            hor, jump = random_inputs(random.Random(zlib.crc32(f"{name}:{ep}".encode())), T_EP)
            rec = refdump.run_trajectory(md, hor, jump, full=True)
            trajs.append(dict(name=f"{name}/{ep}", map_data=np.asarray(md, dtype=np.float64), hor=hor, jump=jump, rec=rec))
    for f in sorted(glob.glob(os.path.join(refdump.REF_ROOT, "bc_replays_tmp", "*.replay"))):
        md, inp, ok = schema.parse_replay(open(f, "rb").read())
        hj = [schema.decode_input_byte(b) for b in inp]
        hor = np.array([h for h, _ in hj], dtype=np.int8); jump = np.array([j for _, j in hj], dtype=np.int8)
        short = os.path.basename(f)[:15]
        rec = refdump.run_trajectory(list(md), hor, jump, full=True)
        trajs.append(dict(name=f"tmp/{short}/own", map_data=np.asarray(list(md), dtype=np.float64), hor=hor, jump=jump, rec=rec))
        hor, jump = random_inputs(random.Random(zlib.crc32(short.encode())), T_EP)
        rec = refdump.run_trajectory(list(md), hor, jump, full=True)
        trajs.append(dict(name=f"tmp/{short}/rand", map_data=np.asarray(list(md), dtype=np.float64), hor=hor, jump=jump, rec=rec))

    d = pack(trajs)
    d["stats_names"] = np.array(sorted(stats))
    d["stats"] = np.array([stats[k] for k in sorted(stats)], dtype=np.int64)
    path = os.path.join(OUT, "doors.npz")
    np.savez_compressed(path, **d)
    print(f"doors: {len(trajs)} episodes, {len(d['hor'])} frames, {os.path.getsize(path) // 1024} KiB;", stats)


if __name__ == "__main__":
    main()
