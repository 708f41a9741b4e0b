#!/bin/bash
# round 2, GPU call 26: solo warps enter the cooperative loop at iteration 0
mkdir -p gpurun_out/r02z
O=gpurun_out/r02z
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('kernel'), l.get('handoff'))"; }
run() {  # name workload envs handoff [env...]
  local name=$1 w=$2 n=$3 h=$4; shift 4
  env "$@" timeout 300 python bench.py --workload $w --envs $n --handoff $h --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$name.log | tail -1 > $O/bench_$name.json
  summ "$name" < $O/bench_$name.json
}
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "handoff or full_size or soak or cooperative" > $O/parity.log 2>&1; tail -3 $O/parity.log
run c3_65k_sm        c3 65536 on
run c3_65k_front     c3 65536 on NPPSIM_SOLO_SM=0
run c3_131k_sm       c3 131072 on
run c3_131k_off      c3 131072 off
echo "== alone, solo warp"; timeout 300 python tools/wedge_bench.py 30 --handoff > $O/wedge_alone_solo.txt 2>&1; cat $O/wedge_alone_solo.txt
P="python tools/phase_profile.py --lib=$PWD/variants/profile.so"
timeout 600 $P 65536 --handoff=on --engine-order=caller > $O/timeline_c3_65k_sm.txt 2>&1; tail -36 $O/timeline_c3_65k_sm.txt | cut -c1-250
