#!/bin/bash
# round 2, GPU call 25: 9 / 10 resident warps per SM (3 x 96 threads at 224 registers, 2 x 160 at 200) vs 8 (2 x 128 at 255)
mkdir -p gpurun_out/r02y
O=gpurun_out/r02y
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('kernel'))"; }
for so in main b96x3 b160x2; do
  for cfg in "c3 65536 off" "c3 65536 on" "c3 131072 off" "c4 131072 off" "c5 65536 off" "c2 4096 off"; do
    set -- $cfg
    lib=$PWD/variants/$so.so; [ $so = main ] && lib=$PWD/nclone_b200/libnppsim.so
    NPPSIM_SOLO_SM=0 NPPSIM_LIB=$lib timeout 300 python bench.py --workload $1 --envs $2 --handoff $3 --steps 60 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_${so}_$1_$2_$3.log | tail -1 > $O/bench_${so}_$1_$2_$3.json
    summ "$so $1 $2 handoff=$3" < $O/bench_${so}_$1_$2_$3.json
  done
done
