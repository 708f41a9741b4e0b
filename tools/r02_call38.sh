#!/bin/bash
# round 2, GPU call 38 (2 GPUs): end build -- GPU suite, smoke, soak, default bench at 1 and 2 GPUs, reference arm at 2 ranks
mkdir -p gpurun_out/r02al
O=gpurun_out/r02al
echo "== GPU suite"; timeout 1700 python -m pytest tests -x -q -m gpu > $O/gpu_tests.log 2>&1; tail -2 $O/gpu_tests.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== soak"; ( time timeout 900 python tools/soak_parity.py 32768 200 77 > $O/soak.log 2>&1 ) 2>&1 | grep real; tail -3 $O/soak.log
echo "== default bench, 1 GPU"; ( time timeout 900 python bench.py > $O/bench_n1.json 2>$O/bench_n1.err ) 2>&1 | grep real
echo "== default bench, 2 GPUs"; ( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 > $O/bench_n2.json 2>$O/bench_n2.err ) 2>&1 | grep real
python - <<'PY'
import json
for n in (1, 2):
    try:
        l = json.loads(open(f"gpurun_out/r02al/bench_n{n}.json").read().strip().split("\n")[-1])
    except Exception as e:
        print(n, "no line", e); continue
    keep = ("value", "ms_per_step", "kernel", "frac", "achieved", "sm_mhz", "reasons", "frac_issue")
    print(n, {k: (l[k] if not isinstance(l[k], dict) else {a: b for a, b in l[k].items() if a in keep}) for k in ("value", "ms_per_step", "e2e", "e2e_obs", "async_groups", "config2", "config4", "full_kernel", "roofline", "cpu_baseline", "clocks", "gpu_launches") if k in l})
    if "per_rank" in l: print("  per_rank", json.dumps(l["per_rank"])[:600])
PY
tail -2 $O/bench_n2.err
