#!/bin/bash
# round 2, GPU call 10: grouped loop with prediction (groups2) vs without (groups1) vs none; instruction-cache metrics per variant
mkdir -p gpurun_out/r02j
O=gpurun_out/r02j
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('handoff'))"; }
echo "== parity (groups2 = libnppsim.so)"; timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_map_compile.py -x -q -m gpu > $O/parity.log 2>&1; tail -3 $O/parity.log
for so in groups2 groups1 nogroups; do
  for cfg in "c3 65536 on" "c3 65536 off" "c4 131072 off" "c5 65536 off" "c3 131072 off"; do
    set -- $cfg
    NPPSIM_LIB=$PWD/variants/$so.so timeout 300 python bench.py --workload $1 --envs $2 --handoff $3 --steps 60 --warmup 5 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_${so}_$1_$2_$3.log | tail -1 > $O/bench_${so}_$1_$2_$3.json
    summ "$so $1 $2 handoff=$3" < $O/bench_${so}_$1_$2_$3.json
  done
done
M="gpu__time_duration.sum,sm__icc_request_hit_rate.pct,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__issue_active.avg.pct,smsp__inst_executed.sum"
for v in "groups2 on level" "groups2 off level" "nogroups on level" "nogroups off level" "nogroups off caller"; do
  set -- $v
  NPPSIM_LIB=$PWD/variants/$1.so ncu --metrics $M --clock-control none -k regex:step_kernel -s 70 -c 2 --csv --log-file $O/icc_$1_$2_$3.csv python bench.py --steps 30 --warmup 5 --spread 60 --no-cpu-baseline --no-secondary --async-groups 0 --handoff $2 --engine-order $3 > /dev/null 2>&1
  echo "icc $v"; python - "$O/icc_$1_$2_$3.csv" <<'PY'
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]; mi, vi = h.index("Metric Name"), h.index("Metric Value")
idi = h.index("ID")
out = {}
for r in rows[1:]:
    out.setdefault(r[idi], {})[r[mi]] = r[vi]
for k, d in out.items():
    print("  ", {m.split("__")[-1][:40]: v for m, v in d.items()})
PY
done
