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
n_envs = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 65536
fixture = next((a for a in sys.argv[1:] if a in ("simplest_with_mines", "locked_doors", "bc_replays", "first_replay")), "simplest_with_mines")
arr, lens = bench.load_maps(fixture)
sim = BatchedSimulator(n_envs)
sim.load_maps([arr[k, :lens[k]] for k in range(len(arr))], np.arange(n_envs) % len(arr))
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
e0.record(); sim.step(acts[7], 4, True); e1.record(); torch.cuda.synchronize()
lib_.nppsim_debug_warp_cycles(wc.ctypes.data, ph.ctypes.data, nw, 0)
d = (ph - ph0).astype(np.float64); w = wc.astype(np.float64)
print(f"last launch: {e0.elapsed_time(e1):.3f} ms = {e0.elapsed_time(e1) * 1.965e6:,.0f} cycles @1.965GHz")
print("per-warp kernel cycles: mean %.0f p50 %.0f p90 %.0f p99 %.0f max %.0f" % (w.mean(), *np.percentile(w, [50, 90, 99]), w.max()))
order = np.argsort(-w)[:8]
n_ent = np.array([sim.maps[m].n_ent for m in sim.env_map_host])
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
