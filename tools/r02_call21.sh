#!/bin/bash
# round 2, GPU call 21: fp64 pipe cost vs active lanes; solo SMs with one solo block per SM (the other slot held empty)
mkdir -p gpurun_out/r02u
O=gpurun_out/r02u
./tools/microbench/fp64_lanes > $O/fp64_lanes.txt 2>&1; cat $O/fp64_lanes.txt
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('kernel'), l.get('handoff'))"; }
run() {  # name workload envs handoff [env...]
  local name=$1 w=$2 n=$3 h=$4; shift 4
  env "$@" timeout 300 python bench.py --workload $w --envs $n --handoff $h --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$name.log | tail -1 > $O/bench_$name.json
  summ "$name" < $O/bench_$name.json
}
run c3_65k_sm        c3 65536 on NPPSIM_SOLO_SM=1
run c3_65k_half      c3 65536 on NPPSIM_SOLO_SM=1 NPPSIM_SOLO_SM_HALF=1
run c3_65k_half_q512 c3 65536 on NPPSIM_SOLO_SM=1 NPPSIM_SOLO_SM_HALF=1 NPPSIM_SOLO_SM_Q8=512 NPPSIM_SOLO_SM_MAX=36
run c3_65k_half_q384 c3 65536 on NPPSIM_SOLO_SM=1 NPPSIM_SOLO_SM_HALF=1 NPPSIM_SOLO_SM_Q8=384 NPPSIM_SOLO_SM_MAX=36
P="python tools/phase_profile.py --lib=$PWD/variants/profile.so"
NPPSIM_SOLO_SM=1 NPPSIM_SOLO_SM_HALF=1 NPPSIM_SOLO_SM_Q8=512 NPPSIM_SOLO_SM_MAX=36 timeout 600 $P 65536 --handoff=on --engine-order=caller > $O/timeline_c3_65k_half.txt 2>&1; tail -34 $O/timeline_c3_65k_half.txt
