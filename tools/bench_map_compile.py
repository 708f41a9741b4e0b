#!/usr/bin/env python
"""Throughput of the device map compiler against the host compiler on the same levels.

    python tools/bench_map_compile.py [n_maps] [fixture]

Prints one JSON line: levels/s on the device (CUDA events, L2 flushed by the workload itself: 65,536 blobs = 2.8 GB),
the achieved HBM rate on algorithmic bytes (row read + blob written) against MEASURED_PEAKS.json, and the host
compiler's single-thread rate on a bounded sample."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from nclone_b200._lib import lib  # noqa: E402

n_maps = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
fixture = sys.argv[2] if len(sys.argv) > 2 else "simplest_with_mines"
L = lib()
arr, lens = bench.load_maps(fixture)
reps = (n_maps + len(arr) - 1) // len(arr)
md = torch.from_numpy(arr).cuda().repeat(reps, 1)[:n_maps].contiguous()
ml = torch.from_numpy(lens).cuda().repeat(reps)[:n_maps].contiguous()
Lrow = md.shape[1]
stride = (int(L.nppsim_map_blob_bound(Lrow)) + 255) & ~255
blobs = torch.empty(n_maps * stride, dtype=torch.uint8, device="cuda")
off = torch.empty(n_maps, dtype=torch.int64, device="cuda")
status = torch.empty(n_maps, dtype=torch.int32, device="cuda")
summary = torch.empty(4, dtype=torch.int32, device="cuda")
stream = torch.cuda.current_stream().cuda_stream


def launch():
    rc = L.nppsim_compile_maps_device(md.data_ptr(), Lrow, ml.data_ptr(), Lrow, n_maps, blobs.data_ptr(), stride, off.data_ptr(),
                                      status.data_ptr(), summary.data_ptr(), stream)
    assert rc == 0, L.nppsim_last_error()


for _ in range(3):
    launch()
torch.cuda.synchronize()
assert summary[3].item() == 0
K = 10
ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
ev[0].record()
for k in range(K):
    launch()
    ev[k + 1].record()
torch.cuda.synchronize()
ms = np.array([ev[k].elapsed_time(ev[k + 1]) for k in range(K)])
blob_bytes = blobs.view(-1, stride)[:, 4:8].contiguous().view(torch.int32).to(torch.int64).sum().item()
row_bytes = int(ml.to(torch.int64).sum().item()) * 8
alg = blob_bytes + row_bytes
peaks = {}
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    pass
peak = None
for key in ("hbm_gbs",):
    if key in peaks:
        peak = float(peaks[key]); break
# host compiler, one thread, bounded sample
sample = min(2000, n_maps)
out = np.zeros(int(L.nppsim_map_blob_bound(Lrow)), dtype=np.uint8)
used = C.c_int64(0); ne = C.c_int32(0); nd = C.c_int32(0)
t0 = time.perf_counter()
for k in range(sample):
    row = arr[k % len(arr)]
    L.nppsim_compile_map(row.ctypes.data, int(lens[k % len(arr)]), out.ctypes.data, len(out), C.byref(used), C.byref(ne), C.byref(nd))
host_s = time.perf_counter() - t0
gbps = alg / (ms.mean() * 1e-3) / 1e9
print(json.dumps({
    "metric": "levels compiled/sec", "value": n_maps / (ms.mean() * 1e-3), "unit": "levels/s", "n_maps": n_maps, "fixture": fixture,
    "ms_per_launch": float(ms.mean()), "ms_min": float(ms.min()),
    "roofline": {"bound": "hbm", "achieved": gbps, "peak": peak, "unit": "GB/s", "frac": (gbps / peak) if peak else None,
                 "algorithmic_bytes_per_level": alg / n_maps},
    "cpu_baseline": {"value": sample / host_s, "unit": "levels/s", "cores": 1, "kind": "host compiler (map_compiler.cpp)", "sample": f"{sample} levels"},
}))
