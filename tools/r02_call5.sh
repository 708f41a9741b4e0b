#!/bin/bash
# round 2, GPU call 5: full GPU suite + smoke + default bench line, then A/B of engine order (level / caller) x handoff (on / off)
mkdir -p gpurun_out/r02e
O=gpurun_out/r02e
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('handoff'))"; }
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; tail -5 $O/pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee $O/smoke.log
echo "== default bench"; ( time timeout 900 python bench.py > $O/bench_default.json 2>$O/bench_default.err ) 2>&1 | grep real; tail -c 6000 $O/bench_default.json; tail -3 $O/bench_default.err
for cfg in "c3 65536" "c3 131072" "c4 131072" "c5 65536" "c2 4096"; do
  set -- $cfg
  for eo in level caller; do for ho in on off; do
    timeout 300 python bench.py --workload $1 --envs $2 --engine-order $eo --handoff $ho --steps 40 --warmup 5 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$1_$2_${eo}_$ho.log | tail -1 > $O/bench_$1_$2_${eo}_$ho.json
    summ "$1 $2 order=$eo handoff=$ho" < $O/bench_$1_$2_${eo}_$ho.json
  done; done
done
