#!/bin/bash
# round 2, GPU call 16: phase barriers re-measured per workload (config 4: 37 % of the stall samples are barrier waits)
mkdir -p gpurun_out/r02p
O=gpurun_out/r02p
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('kernel'), l.get('handoff', {}).get('enabled'))"; }
for so in main nosync nosync_tiles nosync_it0; do
  for cfg in "c4 131072 auto" "c3 131072 auto" "c3 65536 auto" "c5 65536 auto" "c2 4096 off"; do
    set -- $cfg
    lib=$PWD/variants/$so.so; [ $so = main ] && lib=$PWD/nclone_b200/libnppsim.so
    NPPSIM_LIB=$lib timeout 300 python bench.py --workload $1 --envs $2 --handoff $3 --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_${so}_$1_$2_$3.log | tail -1 > $O/bench_${so}_$1_$2_$3.json
    summ "$so $1 $2 handoff=$3" < $O/bench_${so}_$1_$2_$3.json
  done
done
