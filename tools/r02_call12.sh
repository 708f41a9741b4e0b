#!/bin/bash
# round 2, GPU call 12: state-column order at fewer than 32 replicas per level (config 3 at 65,536 envs), 200-step runs
mkdir -p gpurun_out/r02l
O=gpurun_out/r02l
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('kernel'), l.get('handoff'), l.get('async_groups', {}).get('value'))"; }
for rep in 1 2; do
for cfg in "c3 65536 auto caller" "c3 65536 auto level" "c3 65536 off caller" "c3 32768 auto caller" "c3 32768 auto level"; do
  set -- $cfg
  timeout 300 python bench.py --workload $1 --envs $2 --handoff $3 --engine-order $4 --steps 200 --warmup 20 --no-cpu-baseline --no-secondary 2>$O/err_$1_$2_$3_$4_$rep.log | tail -1 > $O/bench_$1_$2_$3_$4_$rep.json
  summ "$1 $2 handoff=$3 order=$4 rep=$rep" < $O/bench_$1_$2_$3_$4_$rep.json
done
done
