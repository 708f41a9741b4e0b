"""Per-phase cycle breakdown of step_kernel (debug build with -DNPP_PROFILE; GPU box).
usage: python tools/phase_profile.py [envs]"""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nclone_b200 import build as b
lib = next((a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--lib=")), os.path.join(ROOT, "variants", "profile.so"))
os.makedirs(os.path.dirname(lib), exist_ok=True)
if not os.path.exists(lib) or "--rebuild" in sys.argv:
    cmd = [b.nvcc_path()] + b.NVCC_FLAGS + ["-DNPP_PROFILE", "-o", lib] + [os.path.join(b.CSRC, s) for s in b.SOURCES]
    subprocess.run(cmd, check=True)
os.environ["NPPSIM_LIB"] = lib
import torch
import bench
from nclone_b200 import BatchedSimulator
from nclone_b200._lib import lib as L
n_envs = next((int(a) for a in sys.argv[1:] if a.isdigit()), 65536)
fixture = next((a for a in sys.argv[1:] if a in ("simplest_with_mines", "locked_doors", "bc_replays", "first_replay")), "simplest_with_mines")
arr, lens = bench.load_maps(fixture)
handoff = {"--handoff=on": True, "--handoff=off": False}.get(next((a for a in sys.argv[1:] if a.startswith("--handoff=")), ""), None)
placement = "caller" if "--engine-order=caller" in sys.argv else "level"
sim = BatchedSimulator(n_envs, handoff=handoff, placement=placement)
sim.load_maps([arr[k, :lens[k]] for k in range(len(arr))], np.arange(n_envs) % len(arr))
print(f"fixture {fixture}, engine order {placement}, handoff {'on' if sim.handoff is not None else 'off'}, kernel {sim.kernel_variant()}")
gen = torch.Generator(device="cuda"); gen.manual_seed(1)
acts = torch.randint(0, 6, (64, n_envs), generator=gen, device="cuda", dtype=torch.uint8)
for s in range(100):
    sim.step(acts[s % 64], 4, True); sim.reset_terminal(2000)
lib_ = L()
lib_.nppsim_debug_phase_cycles.argtypes = [C.c_void_p, C.c_int]
out = (C.c_ulonglong * 16)()
lib_.nppsim_debug_phase_cycles(None, 1)
K = 50
for s in range(K):
    sim.step(acts[s % 64], 4, True); sim.reset_terminal(2000)
lib_.nppsim_debug_phase_cycles(out, 0)
v = np.array(list(out), dtype=np.float64)
names = {0: "dyn_move", 1: "static mines/boost", 2: "dyn_think", 3: "integrate+pre_collision gather", 4: "(substep loop residue)",
         5: "pooled sweep", 6: "closest-point loop", 7: "(post_collision residue)", 8: "think", 9: "post: entity gather+logical",
         10: "post: pooled wall", 15: "kernel total per warp"}
nw = (n_envs + 31) // 32
tot = v[15]
print(f"envs {n_envs}, warps {nw}, launches {K}; per-warp cycles per launch: {tot / nw / K:,.0f}")
for k in sorted(names):
    print(f"  {names[k]:34s} {v[k] / nw / K:12,.0f} cycles/warp/launch  {100 * v[k] / tot:5.1f}%")
print(f"  segment queries, cycles/warp/launch: gather {v[11] / nw / K:,.0f}, scan+queue {v[12] / nw / K:,.0f}, evaluate {v[13] / nw / K:,.0f}, fold {v[14] / nw / K:,.0f}")
print(f"  {'(load/store/loop overhead)':34s} {(tot - v[:11].sum()) / nw / K:12,.0f}  {100 * (tot - v[:11].sum()) / tot:5.1f}%")

# ---- distribution over warps of the last launch
nw = min(nw, 1 << 16)
wc = np.zeros(nw, dtype=np.uint64); ph = np.zeros((nw, 4), dtype=np.uint64)
lib_.nppsim_debug_warp_cycles.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
# phases accumulate over all launches since load; take one more launch with a fresh delta
ph0 = ph.copy(); lib_.nppsim_debug_warp_cycles(wc.ctypes.data, ph0.ctypes.data, nw, 0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
lib_.nppsim_debug_solo_log.argtypes = [C.c_void_p, C.c_int, C.c_int]
lib_.nppsim_debug_solo_log(None, 0, 1)
e0.record(); sim.step(acts[7], 4, True); e1.record(); torch.cuda.synchronize()
solo_log = np.zeros((8192, 12), dtype=np.uint64)
n_solo_log = min(8192, lib_.nppsim_debug_solo_log(solo_log.ctypes.data, 8192, 0))
lib_.nppsim_debug_warp_cycles(wc.ctypes.data, ph.ctypes.data, nw, 0)
d = (ph - ph0).astype(np.float64); w = wc.astype(np.float64)
print(f"last launch: {e0.elapsed_time(e1):.3f} ms = {e0.elapsed_time(e1) * 1.965e6:,.0f} cycles @1.965GHz")
print("per-warp kernel cycles: mean %.0f p50 %.0f p90 %.0f p99 %.0f max %.0f" % (w.mean(), *np.percentile(w, [50, 90, 99]), w.max()))
order = np.argsort(-w)[:8]
col_env = sim.slot_env if sim.slot_env is not None else np.arange(n_envs)      # state column -> caller's env
n_ent = np.array([sim.maps[m].n_ent for m in np.asarray(sim.env_map_host)[col_env]])
for o in order:
    envs = np.arange(o * 32, min(o * 32 + 32, n_envs))
    print(f"  warp {o}: total {w[o]:,.0f} entities/sweep/closest/post {d[o]} max n_ent in warp {n_ent[envs].max()}")

# ---- timeline of the last launch: who ends the kernel?
span = np.zeros((nw, 2), dtype=np.uint64)
lib_.nppsim_debug_warp_span.argtypes = [C.c_void_p, C.c_int]
lib_.nppsim_debug_warp_span(span.ctypes.data, nw)
cost = sim.warp_cost.cpu().numpy().astype(np.int64)[:nw] - 1
t0 = span[:, 0].min()
st = (span[:, 0] - t0).astype(np.float64) / 1e3; en = (span[:, 1] - t0).astype(np.float64) / 1e3
print(f"timeline (us from the first warp's start): kernel span {en.max():.1f}; all warps started by {st.max():.1f}; "
      f"50% / 90% / 99% of warps finished by {np.percentile(en, 50):.1f} / {np.percentile(en, 90):.1f} / {np.percentile(en, 99):.1f}")
busy = np.sort(en)
for frac in (0.5, 0.75, 0.9, 0.95, 0.99):
    t = np.percentile(en, 100 * frac); print(f"  resident warps at t={t:.0f} us: {int(((st <= t) & (en > t)).sum())} of {148 * 8} slots")
print("  last finishers: (warp group, start us, end us, duration us, cooperative iterations)")
for o in np.argsort(-en)[:12]:
    print(f"    {o:6d} {st[o]:8.1f} {en[o]:8.1f} {en[o] - st[o]:8.1f} {cost[o]:6d}")
long_ = np.argsort(-cost)[:12]
print("  most cooperative iterations: (warp group, start us, end us, duration us, iterations; cycles in entities / sweep / closest-point loop / post_collision)")
for o in long_:
    print(f"    {o:6d} {st[o]:8.1f} {en[o]:8.1f} {en[o] - st[o]:8.1f} {cost[o]:6d}   {d[o].astype(np.int64)}")
quiet = np.nonzero(cost == 0)[0]
print("  warps without cooperative iterations: mean cycles entities / sweep / closest / post:", d[quiet].mean(0).astype(np.int64), "duration us mean", (en - st)[quiet].mean())

# ---- envs served by solo warps in the last launch (handoff): who ends the kernel now?
if n_solo_log:
    lg = solo_log[:n_solo_log]
    env = (lg[:, 0] & 0xffffff).astype(np.int64); tf = ((lg[:, 0] >> 24) & 31).astype(np.int64)
    conv = ((lg[:, 0] >> 32) & 1).astype(bool); init = ((lg[:, 0] >> 33) & 1).astype(bool); sres = ((lg[:, 0] >> 34) & 3).astype(np.int64)
    pick = (lg[:, 1].astype(np.float64) - float(t0)) / 1e3; end = (lg[:, 2].astype(np.float64) - float(t0)) / 1e3
    pub = np.where(lg[:, 3] > 0, (lg[:, 3].astype(np.float64) - float(t0)) / 1e3, np.nan)
    print(f"solo warps: {n_solo_log} envs served, {int(init.sum())} queued at launch (a), {int((~init).sum())} handed over mid-launch (b), "
          f"{int(conv.sum())} of them by warps of finished regular blocks")
    print(f"  regular warps: last one finished at {en.max():.1f} us; solo services: last one finished at {end.max():.1f} us")
    if init.any():
        print(f"  (a) service duration us: mean {np.mean((end - pick)[init]):.1f} max {np.max((end - pick)[init]):.1f}; ends: p50 {np.percentile(end[init], 50):.1f} p90 {np.percentile(end[init], 90):.1f} max {end[init].max():.1f}")
    b = ~init
    if b.any():
        print(f"  (b) published at us: p10 {np.nanpercentile(pub[b], 10):.1f} p50 {np.nanpercentile(pub[b], 50):.1f} p90 {np.nanpercentile(pub[b], 90):.1f} max {np.nanmax(pub[b]):.1f}; "
              f"wait for a warp us: mean {np.nanmean((pick - pub)[b]):.1f} max {np.nanmax((pick - pub)[b]):.1f}; service us: mean {np.mean((end - pick)[b]):.1f} max {np.max((end - pick)[b]):.1f}")
        print("  (b) first tick on the solo warp: " + ", ".join(f"t={k}: {int((tf[b] == k).sum())}" for k in range(0, 4)) +
              f"; handed over mid-tick (resumed at substep 1 / 2 / 3): {int((sres[b] == 1).sum())} / {int((sres[b] == 2).sum())} / {int((sres[b] == 3).sum())}, at a tick boundary: {int((sres[b] == 0).sum())}")
    its = lg[:, 4].astype(np.int64); smid = lg[:, 5].astype(np.int64)
    long_ = its >= 300
    if long_.any():
        cyc = (end - pick)[long_] * 1965.0 / its[long_]
        print(f"  services of >= 300 cooperative iterations: {int(long_.sum())}; cycles per iteration (whole service / iterations): mean {cyc.mean():.0f} min {cyc.min():.0f} max {cyc.max():.0f}")
    print(f"  SMs that ran solo services: {len(np.unique(smid))} ({np.unique(smid)[:40].tolist()})")
    print("  last solo services: (column, kind, first tick, published us, picked up us, ended us, cooperative iterations, smid, us in entities / sweep / closest-point loop / post_collision)")
    for o in np.argsort(-end)[:24]:
        print(f"    {env[o]:7d} {'a' if init[o] else ('b/conv' if conv[o] else 'b'):7s} {tf[o]:2d} {pub[o]:9.1f} {pick[o]:9.1f} {end[o]:9.1f} {its[o]:6d} {smid[o]:4d}   phases us {np.round(lg[o, 6:10].astype(np.float64) / 1965.0, 1).tolist()} in-loop cycles/iteration {lg[o, 10] / max(its[o], 1):.0f} park/take us {lg[o, 11] / 1965.0:.1f}")
