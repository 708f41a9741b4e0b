#!/bin/bash
# round 2, GPU call 17: detachable warps (a warp that serves a wedged ninja leaves its block's phase barriers) vs the same build without
mkdir -p gpurun_out/r02q
O=gpurun_out/r02q
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('kernel'), l.get('handoff', {}).get('enabled'))"; }
echo "== parity"; timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/parity.log 2>&1; tail -3 $O/parity.log
for so in main nodetach; do
  for cfg in "c4 131072 auto" "c3 131072 auto" "c3 65536 off" "c5 65536 auto" "c2 4096 off" "c3 262144 auto"; do
    set -- $cfg
    lib=$PWD/variants/$so.so; [ $so = main ] && lib=$PWD/nclone_b200/libnppsim.so
    NPPSIM_LIB=$lib timeout 300 python bench.py --workload $1 --envs $2 --handoff $3 --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_${so}_$1_$2_$3.log | tail -1 > $O/bench_${so}_$1_$2_$3.json
    summ "$so $1 $2 handoff=$3" < $O/bench_${so}_$1_$2_$3.json
  done
done
