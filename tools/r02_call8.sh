#!/bin/bash
# round 2, GPU call 8: residency A/B with the grouped service in place (3 / 4 blocks per SM at 168 / 128 registers vs 2 at 255)
mkdir -p gpurun_out/r02h
O=gpurun_out/r02h
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('handoff'))"; }
for so in groups bps3 bps4; do
  for cfg in "c3 65536 on" "c3 65536 off" "c4 131072 off" "c5 65536 off" "c3 131072 off" "c2 4096 off"; do
    set -- $cfg
    NPPSIM_LIB=$PWD/variants/$so.so timeout 300 python bench.py --workload $1 --envs $2 --handoff $3 --steps 40 --warmup 5 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_${so}_$1_$2_$3.log | tail -1 > $O/bench_${so}_$1_$2_$3.json
    summ "$so $1 $2 handoff=$3" < $O/bench_${so}_$1_$2_$3.json
  done
done
