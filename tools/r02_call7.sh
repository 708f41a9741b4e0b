#!/bin/bash
# round 2, GPU call 7: grouped cooperative service (several wedged lanes of a warp served at once) -- parity, then A/B vs the same build without it
mkdir -p gpurun_out/r02g
O=gpurun_out/r02g
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('handoff'))"; }
echo "== parity (groups.so)"; NPPSIM_LIB=$PWD/variants/groups.so timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/parity_groups.log 2>&1; tail -3 $O/parity_groups.log
for so in groups nogroups; do
  for cfg in "c3 65536" "c4 131072" "c5 65536" "c3 131072"; do
    set -- $cfg
    for ho in on off; do
      NPPSIM_LIB=$PWD/variants/$so.so timeout 300 python bench.py --workload $1 --envs $2 --handoff $ho --steps 40 --warmup 5 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_${so}_$1_$2_$ho.log | tail -1 > $O/bench_${so}_$1_$2_$ho.json
      summ "$so $1 $2 handoff=$ho" < $O/bench_${so}_$1_$2_$ho.json
    done
  done
done
P="python tools/phase_profile.py --lib=$PWD/variants/groups_prof.so"
timeout 300 $P 131072 locked_doors --handoff=off > $O/timeline_c4_off.txt 2>&1; grep -A12 "^timeline" $O/timeline_c4_off.txt
timeout 300 $P 131072 locked_doors --handoff=on > $O/timeline_c4_on.txt 2>&1; grep -A6 "^timeline" $O/timeline_c4_on.txt; grep -A6 "^solo warps" $O/timeline_c4_on.txt
