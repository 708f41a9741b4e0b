#!/bin/bash
# round 2, GPU call 36: L1 prefetch of the next roaming entities' state in dyn_move / dyn_think (sim_full, c5)
mkdir -p gpurun_out/r02aj
O=gpurun_out/r02aj
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "bc_replays or doors or basic_variant" > $O/parity.log 2>&1; tail -2 $O/parity.log
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel']), l.get('kernel'))"; }
for so in main ahead0 ahead1 ahead4; do
  lib=$PWD/variants/$so.so; [ $so = main ] && lib=$PWD/nclone_b200/libnppsim.so
  NPPSIM_LIB=$lib timeout 300 python bench.py --workload c5 --steps 60 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$so.log | tail -1 > $O/bench_$so.json
  summ "$so c5" < $O/bench_$so.json
done
