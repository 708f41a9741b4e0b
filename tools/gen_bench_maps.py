"""Generate the benchmark level sets with the reference's own generators (build container only;
/root/reference does not exist on the GPU box, so the result is committed as a fixture).

  python tools/gen_bench_maps.py simplest_with_mines 8192 2000      -> bench_data/simplest_with_mines.npz
  python tools/gen_bench_maps.py locked_doors 2048 50000            -> bench_data/locked_doors.npz

`simplest_with_mines` is BASELINE.json config 3 ("curriculum 3: exit switch + 5-7 toggle mines"):
GeneratorRegistry(category, seed=base+i) as in gym_environment/generator_registry.py:31-58.
`locked_doors` is config 4: maps from the `simplest*` categories that contain a type-6 record
(map_generation/map_horizontal_corridor.py:576-615), kept until `count` distinct maps are found.
map_data rows are stored as float64 (generators emit non-integer coordinates), zlib-compressed by npz.
"""
import multiprocessing as mp
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.fuzz_vs_ref import generated_map  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench_data")


def gen_simple(seed):
    return np.asarray(generated_map("simplest_with_mines", seed), dtype=np.float64)


def gen_locked(seed):
    for cat in ("simplest", "simplest_few_mines", "simplest_with_mines"):
        md = np.asarray(generated_map(cat, seed), dtype=np.float64)
        i = 1235
        while i + 4 < len(md):
            if md[i] == 6:
                return md
            i += 9 if md[i] in (6, 8) else 5
    return None


def main():
    kind, count, base = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    os.makedirs(OUT, exist_ok=True)
    maps = []
    with mp.Pool(os.cpu_count()) as pool:
        if kind == "simplest_with_mines":
            maps = pool.map(gen_simple, range(base, base + count), chunksize=16)
        else:
            seed = base
            while len(maps) < count:
                got = pool.map(gen_locked, range(seed, seed + 1024), chunksize=16)
                maps += [m for m in got if m is not None]
                seed += 1024
            maps = maps[:count]
    # distinct only
    seen, uniq = set(), []
    for m in maps:
        k = m.tobytes()
        if k not in seen:
            seen.add(k); uniq.append(m)
    lens = np.array([len(m) for m in uniq], dtype=np.int32)
    stride = int(lens.max())
    arr = np.zeros((len(uniq), stride), dtype=np.float64)
    for k, m in enumerate(uniq):
        arr[k, :len(m)] = m
    path = os.path.join(OUT, kind + ".npz")
    # tiles and most records are small integers: an int16 image + the few non-integer values as a sparse list
    frac = arr != np.floor(arr)
    rows, cols = np.nonzero(frac)
    np.savez_compressed(path, lens=lens, ints=np.where(frac, 0, arr).astype(np.int16),
                        frac_row=rows.astype(np.int32), frac_col=cols.astype(np.int32), frac_val=arr[rows, cols])
    print(kind, "maps", len(uniq), "of", len(maps), "stride", stride, "file", os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
