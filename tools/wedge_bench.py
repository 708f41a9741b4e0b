"""Critical-path microbenchmark: one wedged ninja alone on the GPU.

A ninja pushed into a corner takes all 32 depenetration iterations in all 4 substeps of every tick (ninja.py:303-379);
about 8 of 65,536 envs are in that state in any agent step of the bench workload, and the step kernel cannot end before
their 4 x 4 x 30 cooperative iterations have run one after the other.  bench_data/wedges.npz holds levels + action
prefixes (found on the CPU oracle) that put an env into such a state; this tool replays the prefix, snapshots the state
and times the wedged agent step by itself (1 env, 1 block), against a free-running step of the same env.

    python tools/wedge_bench.py [reps]      -> us per step, cooperative iterations, cycles per iteration
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nclone_b200 import BatchedSimulator  # noqa: E402


def main():
    reps = next((int(a) for a in sys.argv[1:] if a.isdigit()), 50)
    handoff = True if "--handoff" in sys.argv else None      # --handoff: the wedged step runs on a solo warp from t = 0 (the env is in the queue)
    d = np.load(os.path.join(ROOT, "bench_data", "wedges.npz"))
    sm_mhz = 1965.0
    for k in range(len(d["lens"])):
        md = d["maps"][k, :d["lens"][k]]
        acts = d["acts"][d["act_off"][k]:d["act_off"][k + 1]]
        sim = BatchedSimulator(1, handoff=handoff)
        sim.load_maps([md])
        free = torch.tensor([0], dtype=torch.uint8, device="cuda")
        sim.replay_to_checkpoint([acts[:-1]], frame_skip=4)
        snap = [t.clone() for t in (sim.nin_f, sim.nin_i, sim.env_i, sim.ent_flag, sim.dyn_f, sim.dyn_i)]

        def restore():
            for t, s in zip((sim.nin_f, sim.nin_i, sim.env_i, sim.ent_flag, sim.dyn_f, sim.dyn_i), snap):
                t.copy_(s)

        a = torch.tensor([int(acts[-1])], dtype=torch.uint8, device="cuda")
        times = []
        for r in range(reps + 3):
            restore()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); sim.step(a, 4, False); e1.record()
            torch.cuda.synchronize()
            if r >= 3:
                times.append(e0.elapsed_time(e1) * 1000)
        coop = int(sim.warp_cost[0].item()) - 1
        if handoff:
            coop = [354, 428, 428][k]       # (a solo warp's count is not in warp_cost; the numbers of the run without handoff)
            print(f"  handoff: queue length at the timed step {sim.solo_queue_length()}, status {sim.handoff_status()}")
        # the same env at rest (fresh reset, standing on the floor): the non-wedge part of a step
        base = []
        for r in range(reps + 3):
            sim.reset()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); sim.step(free, 4, False); e1.record()
            torch.cuda.synchronize()
            if r >= 3:
                base.append(e0.elapsed_time(e1) * 1000)
        if hasattr(sim.L, "nppsim_debug_coop_cycles"):          # -DNPP_PROFILE build (NPPSIM_LIB=variants/profile.so)
            import ctypes as C
            out = (C.c_ulonglong * 8)()
            ph = (C.c_ulonglong * 16)()
            sim.L.nppsim_debug_coop_cycles(None, 1); sim.L.nppsim_debug_phase_cycles(None, 1)
            restore(); sim.step(a, 4, False); torch.cuda.synchronize()
            sim.L.nppsim_debug_coop_cycles(out, 0); sim.L.nppsim_debug_phase_cycles(ph, 0)
            names = {0: "dyn_move", 1: "mines/boost", 2: "dyn_think", 3: "integrate+pre", 4: "substep residue", 5: "sweep", 6: "closest loop",
                     7: "post residue", 8: "think", 9: "post gather+logical", 10: "post wall", 15: "kernel"}
            print("  phase cycles (4 warps of the block add up; /4 for the wedge warp's share when the others idle at barriers):")
            print("   ", ", ".join(f"{names[k]} {ph[k]:,}" for k in sorted(names)))
            v = list(out); it = max(v[3], 1)
            print(f"  batches {v[6]}, failed {v[7] & 0xffffffff} (sum of first bad lanes {v[7] >> 32})")
            print(f"  per iteration (clock64): evaluate {v[0] / it:.0f}, select {v[1] / it:.0f}, push {v[2] / it:.0f}; park/take/back {v[5] / max(v[4], 1):.0f} per service, {v[4]} services, {v[3]} iterations")
        t, b = float(np.median(times)), float(np.median(base))
        print(f"wedge {k}: step {t:.1f} us (rest step {b:.1f} us), {coop} cooperative iterations of {d['cp_calls'][k]} closest-point calls"
              f" -> {(t - b) * sm_mhz / max(coop, 1):.0f} cycles per cooperative iteration at {sm_mhz:.0f} MHz")


if __name__ == "__main__":
    main()
