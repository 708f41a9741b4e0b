#!/bin/bash
# round 2, GPU call 41 (4 GPUs): final tree -- GPU suite, smoke, default bench at 1 and 4 GPUs, reference arm under torchrun
mkdir -p gpurun_out/r02ao
O=gpurun_out/r02ao
echo "== GPU suite"; timeout 1700 python -m pytest tests -x -q -m gpu > $O/gpu_tests.log 2>&1; tail -2 $O/gpu_tests.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== default bench, 1 GPU"; ( time timeout 900 python bench.py > $O/bench_n1.json 2>$O/bench_n1.err ) 2>&1 | grep real
echo "== default bench, 4 GPUs"; ( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 4 > $O/bench_n4.json 2>$O/bench_n4.err ) 2>&1 | grep real
echo "== reference arm, 4 ranks"; ( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29545 bench.py --impl reference --gpus 4 --steps 20 --warmup 3 > $O/ref_n4.json 2>$O/ref_n4.err ) 2>&1 | grep real; tail -1 $O/ref_n4.json | cut -c1-300
python - <<'PY'
import json
for n in (1, 4):
    try:
        l = json.loads(open(f"gpurun_out/r02ao/bench_n{n}.json").read().strip().split("\n")[-1])
    except Exception as e:
        print(n, "no line", e); continue
    keep = ("value", "ms_per_step", "kernel", "frac", "sm_mhz", "reasons")
    print(n, {k: (l[k] if not isinstance(l[k], dict) else {a: b for a, b in l[k].items() if a in keep}) for k in ("value", "ms_per_step", "e2e", "e2e_obs", "async_groups", "config2", "config4", "full_kernel", "roofline", "cpu_baseline", "clocks", "gpu_launches") if k in l})
PY
