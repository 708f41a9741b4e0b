#!/bin/bash
# round 2, GPU call 15: final build -- GPU suite, smoke, default bench, ncu launch list + --set full, warp timelines
mkdir -p gpurun_out/r02o
O=gpurun_out/r02o
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; tail -3 $O/pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee $O/smoke.log
echo "== default bench"; ( time timeout 900 python bench.py > $O/bench_default.json 2>$O/bench_default.err ) 2>&1 | grep real; python - <<'PY'
import json
l = json.loads(open("gpurun_out/r02o/bench_default.json").read().strip().split("\n")[-1])
print({k: (l[k] if not isinstance(l[k], dict) else {a: b for a, b in l[k].items() if a in ("value", "ms_per_step", "kernel", "frac", "achieved", "sm_mhz", "reasons")}) for k in ("value", "ms_per_step", "e2e", "e2e_obs", "async_groups", "config2", "config4", "full_kernel", "roofline", "cpu_baseline", "cpu_baseline_port", "clocks")})
PY
tail -3 $O/bench_default.err
CMD="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-secondary --async-groups 0"
echo "== ncu launch list"; $CMD > $O/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 200 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1; tail -1 $O/plain.log | cut -c1-200
echo "== ncu full"; ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 320 -c 2 -o $O/r02_step $CMD > $O/ncu_full.log 2>&1; ls -la $O/*.ncu-rep
echo "== ncu full c4"; ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 140 -c 1 -o $O/r02_step_c4 python bench.py --workload c4 --steps 10 --warmup 3 --spread 120 --no-cpu-baseline --no-secondary --async-groups 0 > $O/ncu_full_c4.log 2>&1; ls -la $O/*.ncu-rep
P="python tools/phase_profile.py --lib=$PWD/variants/main_prof.so"
timeout 300 $P 65536 simplest_with_mines --engine-order=caller > $O/timeline_c3_65k.txt 2>&1; grep -A8 "^timeline" $O/timeline_c3_65k.txt; grep -A6 "^solo warps" $O/timeline_c3_65k.txt
timeout 300 $P 131072 locked_doors > $O/timeline_c4_131k.txt 2>&1; grep -A6 "^timeline" $O/timeline_c4_131k.txt
timeout 300 $P 65536 bc_replays > $O/timeline_c5_65k.txt 2>&1; grep -A6 "^timeline" $O/timeline_c5_65k.txt
