#!/bin/bash
# round 2, GPU call 9: full suite + smoke + default bench, soak, ncu launch list and --set full of step_kernel
mkdir -p gpurun_out/r02i
O=gpurun_out/r02i
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -x -q -m gpu --durations=8 > $O/pytest_gpu.log 2>&1; tail -14 $O/pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee $O/smoke.log
echo "== default bench"; ( time timeout 900 python bench.py > $O/bench_default.json 2>$O/bench_default.err ) 2>&1 | grep real; python - <<'PY'
import json
l = json.loads(open("gpurun_out/r02i/bench_default.json").read().strip().split("\n")[-1])
print({k: (l[k] if not isinstance(l[k], dict) else {a: b for a, b in l[k].items() if a in ("value", "ms_per_step", "kernel", "handoff", "frac", "achieved")}) for k in ("value", "ms_per_step", "e2e", "e2e_obs", "async_groups", "config2", "config4", "full_kernel", "roofline", "handoff", "clocks")})
PY
tail -3 $O/bench_default.err
echo "== soak"; timeout 900 python tools/soak_parity.py 32768 200 77 2>&1 | tail -3 | tee $O/soak.log
CMD="python bench.py --steps 30 --warmup 5 --spread 60 --no-cpu-baseline --no-secondary --async-groups 0"
echo "== ncu launch list"; $CMD > $O/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 200 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1; tail -2 $O/ncu_launches.log | cut -c1-300
echo "== ncu full"; ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 70 -c 2 -o $O/r02_step $CMD > $O/ncu_full.log 2>&1; tail -2 $O/ncu_full.log | cut -c1-300; ls -la $O/*.ncu-rep
