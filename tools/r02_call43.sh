#!/bin/bash
# round 2, GPU call 43: final tree -- GPU suite, smoke, default bench
mkdir -p gpurun_out/r02ap
O=gpurun_out/r02ap
echo "== GPU suite"; timeout 1700 python -m pytest tests -x -q -m gpu > $O/gpu_tests.log 2>&1; tail -2 $O/gpu_tests.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== default bench"; ( time timeout 900 python bench.py > $O/bench_n1.json 2>$O/bench_n1.err ) 2>&1 | grep real
python - <<'PY'
import json
l = json.loads(open("gpurun_out/r02ap/bench_n1.json").read().strip().split("\n")[-1])
keep = ("value", "ms_per_step", "kernel", "frac", "sm_mhz", "reasons", "frac_issue")
print({k: (l[k] if not isinstance(l[k], dict) else {a: b for a, b in l[k].items() if a in keep}) for k in ("value", "ms_per_step", "e2e", "e2e_obs", "async_groups", "config2", "config4", "full_kernel", "roofline", "cpu_baseline", "clocks", "gpu_launches") if k in l})
print({k: l[k]["e2e"]["value"] for k in ("config2", "config4", "full_kernel")})
PY
