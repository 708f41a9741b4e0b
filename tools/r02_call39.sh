#!/bin/bash
# round 2, GPU call 39: solo-SM cap at 131,072 envs of config 3 (148 queued envs for 144 solo warps at the default cap of 18 SMs)
mkdir -p gpurun_out/r02am
O=gpurun_out/r02am
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel']), l.get('handoff'))"; }
run() {  # name workload envs [env...]
  local name=$1 w=$2 n=$3; shift 3
  env "$@" timeout 300 python bench.py --workload $w --envs $n --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$name.log | tail -1 > $O/bench_$name.json
  summ "$name" < $O/bench_$name.json
}
run c3_131k_cap18      c3 131072
run c3_131k_cap22      c3 131072 NPPSIM_SOLO_SM_MAX=22
run c3_131k_cap26      c3 131072 NPPSIM_SOLO_SM_MAX=26
run c3_131k_cap26_q320 c3 131072 NPPSIM_SOLO_SM_MAX=26 NPPSIM_SOLO_SM_Q8=320
run c3_131k_cap32_q384 c3 131072 NPPSIM_SOLO_SM_MAX=32 NPPSIM_SOLO_SM_Q8=384
run c3_98k_cap18       c3 98304
run c3_98k_cap24_q320  c3 98304 NPPSIM_SOLO_SM_MAX=24 NPPSIM_SOLO_SM_Q8=320
