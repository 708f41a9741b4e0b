#!/bin/bash
# round 2, GPU call 29: handoff (solo SMs, mid-tick) at the larger batches; the solo kernel with the grouped service compiled in
mkdir -p gpurun_out/r02ac
O=gpurun_out/r02ac
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('kernel'), l.get('handoff'))"; }
run() {  # name workload envs handoff [env...]
  local name=$1 w=$2 n=$3 h=$4; shift 4
  env "$@" timeout 300 python bench.py --workload $w --envs $n --handoff $h --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$name.log | tail -1 > $O/bench_$name.json
  summ "$name" < $O/bench_$name.json
}
G=NPPSIM_LIB=$PWD/variants/sologroups.so
run c3_65k_on        c3 65536 on
run c3_65k_on_g      c3 65536 on $G
run c3_131k_on       c3 131072 on
run c3_131k_on_g     c3 131072 on $G
run c3_131k_off      c3 131072 off
run c3_262k_on       c3 262144 on
run c3_262k_on_g     c3 262144 on $G
run c3_262k_off      c3 262144 off
run c4_131k_on       c4 131072 on
run c4_131k_on_g     c4 131072 on $G
run c4_131k_on_g36   c4 131072 on $G NPPSIM_SOLO_SM_MAX=36
run c4_131k_off      c4 131072 off
run c4_65k_on        c4 65536 on
run c4_65k_on_g      c4 65536 on $G
run c4_65k_off       c4 65536 off
run c3_32k_on        c3 32768 on
run c3_32k_on2       c3 32768 on NPPSIM_SOLO_SM=2
run c3_32k_off       c3 32768 off
