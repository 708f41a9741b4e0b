#!/bin/bash
# round 2, GPU call 30: automatic handoff policy (on from 1,024 envs, dropped when the queue overflows in most launches) + whole GPU suite + smoke
mkdir -p gpurun_out/r02ad
O=gpurun_out/r02ad
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('kernel'), l.get('handoff'))"; }
run() {  # name workload envs handoff [env...]
  local name=$1 w=$2 n=$3 h=$4; shift 4
  env "$@" timeout 300 python bench.py --workload $w --envs $n --handoff $h --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$name.log | tail -1 > $O/bench_$name.json
  summ "$name" < $O/bench_$name.json
}
echo "== GPU suite"; timeout 1700 python -m pytest tests -x -q -m gpu > $O/gpu_tests.log 2>&1; tail -3 $O/gpu_tests.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
run c3_65k_auto   c3 65536 auto
run c3_131k_auto  c3 131072 auto
run c3_262k_auto  c3 262144 auto
run c4_131k_auto  c4 131072 auto
run c4_65k_auto   c4 65536 auto
run c2_4k_auto    c2 4096 auto
run c5_65k_auto   c5 65536 auto
