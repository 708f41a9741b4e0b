#!/bin/bash
# round 2, GPU call 35: column order re-measured on the end build (c3 65,536: 16 replicas per level)
mkdir -p gpurun_out/r02ai
O=gpurun_out/r02ai
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('kernel'), l.get('engine_order'))"; }
run() {  # name workload envs order [env...]
  local name=$1 w=$2 n=$3 o=$4; shift 4
  env "$@" timeout 300 python bench.py --workload $w --envs $n --engine-order $o --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$name.log | tail -1 > $O/bench_$name.json
  summ "$name" < $O/bench_$name.json
}
run c3_65k_caller  c3 65536 caller
run c3_65k_level   c3 65536 level
run c3_65k_level_zc c3 65536 level NPPSIM_ZEROCOPY=2
run c3_32k_caller  c3 32768 caller
run c3_32k_level   c3 32768 level
run c3_131k_caller c3 131072 caller
run c3_131k_level  c3 131072 level
