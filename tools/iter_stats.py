"""Depenetration-loop statistics of the bench workload on the CPU oracle (build container or GPU box).
Builds a -DOSIM_STATS variant of the oracle; usage: python tools/iter_stats.py [n_maps] [agent_steps]"""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
so = "/tmp/liboracle_stats.so"
subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-std=c11", "-ffp-contract=off", "-fno-builtin-pow", "-DOSIM_STATS", "-pthread", "-o", so,
                os.path.join(ROOT, "oracle", "nsim_oracle.c"), os.path.join(ROOT, "oracle", "baseline_runner.c"), "-lm"], check=True)
L = C.CDLL(so)
L.osim_create.restype = C.c_void_p
dp = C.POINTER(C.c_double)
L.osim_load.argtypes = [C.c_void_p, dp, C.c_int]; L.osim_tick.argtypes = [C.c_void_p, C.c_int, C.c_int]; L.osim_reset.argtypes = [C.c_void_p]
L.osim_get_state.argtypes = [C.c_void_p, dp, C.POINTER(C.c_int64), dp, C.POINTER(C.c_int64)]
L.osim_num_entities.argtypes = [C.c_void_p]
n_maps = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 400
arr, lens = bench.load_maps(os.environ.get("FIXTURE", "simplest_with_mines"))
rng = np.random.default_rng(0)
H = [0, -1, 1, 0, -1, 1]; J = [0, 0, 0, 1, 1, 1]
calls = (C.c_long).in_dll(L, "osim_stat_calls")
per_map = []
nf = np.zeros(31); ni = np.zeros(27, dtype=np.int64)
heavy = []
for m in range(n_maps):
    S = C.c_void_p(L.osim_create()); L.osim_stat_new_env()
    md = np.ascontiguousarray(arr[m, :lens[m]])
    L.osim_load(S, md.ctypes.data_as(dp), len(md))
    frames = 0; c0 = calls.value; ticks = 0
    for s in range(n_steps):
        a = int(rng.integers(0, 6))
        for k in range(4):
            c1 = calls.value
            L.osim_tick(S, H[a], J[a]); frames += 1; ticks += 1
            L.osim_get_state(S, nf.ctypes.data_as(dp), ni.ctypes.data_as(C.POINTER(C.c_int64)), None, None)
            if calls.value - c1 > 40:
                heavy.append((m, frames, calls.value - c1, nf[0], nf[1], nf[2], nf[3], int(ni[0])))
            if 6 <= ni[0] <= 8: break
        if 6 <= ni[0] <= 8 or frames >= 2000:
            L.osim_reset(S); frames = 0
    per_map.append(((calls.value - c0) / ticks, L.osim_num_entities(S)))
it = np.array(list((C.c_long * 34).in_dll(L, "osim_stat_iters")))
ns = np.array(list((C.c_long * 64).in_dll(L, "osim_stat_nseg")))
print("collide_vs_tiles calls:", it.sum(), " closest-point calls:", calls.value, " mean per substep:", calls.value / it.sum())
print("iteration histogram (k closest-point calls -> share):")
for k in range(34):
    if it[k]: print(f"  {k:2d}: {it[k] / it.sum():.5f}   share of closest-point calls {k * it[k] / calls.value:.4f}")
print("gathered segments per substep: mean", (ns * np.arange(64)).sum() / ns.sum(), "hist", {k: int(v) for k, v in enumerate(ns) if v})
pm = np.array(per_map)
print("closest-point calls per tick by map: mean %.2f p50 %.2f p90 %.2f p99 %.2f max %.2f" % (pm[:, 0].mean(), *np.percentile(pm[:, 0], [50, 90, 99]), pm[:, 0].max()))
print("entities per map: mean %.1f max %d" % (pm[:, 1].mean(), pm[:, 1].max()))
print("corr(calls/tick, n_ent) = %.3f" % np.corrcoef(pm[:, 0], pm[:, 1])[0, 1])
print("heavy ticks (>40 closest-point calls):", len(heavy), "of", int(it.sum() / 4))
for h in heavy[:20]: print("  map %d frame %d calls %d pos (%.6f, %.6f) v (%.4f, %.4f) state %d" % h)
per = np.array(list((C.c_long * (5 * 34)).in_dll(L, "osim_stat_period"))).reshape(5, 34)
print("full-length (32-call) loops:", int(per.sum()), " never periodic (p <= 4):", int(per[0].sum()))
for p in range(1, 5):
    if per[p].sum(): print(f"  period {p}: {int(per[p].sum())}  first repeat at call k: " + str({k: int(v) for k, v in enumerate(per[p]) if v}))
tr = np.array(list((C.c_long * 34).in_dll(L, "osim_stat_traj")))
print("full-length loops whose trajectory rejoins the previous full-length loop's of the env at call k:", {k: int(v) for k, v in enumerate(tr) if v}, "(33 = never)")
pr = list((C.c_long * 4).in_dll(L, "osim_stat_pred"))
print("calls 2..31: total %d, winner == winner two calls back %d (%.3f), else == previous winner %d, neither %d" % (pr[0], pr[1], pr[1] / max(pr[0], 1), pr[2], pr[3]))
ar = list((C.c_long * 3).in_dll(L, "osim_stat_arc"))
print("calls 2..31 with an arc in the list %d, inside the query box %d, arc wins %d" % tuple(ar))
