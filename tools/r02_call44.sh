#!/bin/bash
# round 2, GPU call 44: block size on config 4 (barrier waits = 37 % of its stall samples): 128 (kept) vs 64 vs 32 threads
mkdir -p gpurun_out/r02aq
O=gpurun_out/r02aq
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel']), l.get('kernel'))"; }
for so in main b64 b32; do
  for cfg in "c4 131072" "c4 65536" "c3 65536" "c5 65536"; do
    set -- $cfg
    lib=$PWD/variants/$so.so; [ $so = main ] && lib=$PWD/nclone_b200/libnppsim.so
    NPPSIM_LIB=$lib timeout 300 python bench.py --workload $1 --envs $2 --steps 60 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_${so}_$1_$2.log | tail -1 > $O/bench_${so}_$1_$2.json
    summ "$so $1 $2" < $O/bench_${so}_$1_$2.json
  done
done
