#!/bin/bash
# round 2, GPU call 34: batched speculation in the cooperative loop (four iterations on the predicted segments, one verification pass) vs NPP_COOP_BATCH=0
mkdir -p gpurun_out/r02ah
O=gpurun_out/r02ah
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "handoff or full_size or soak or cooperative or basic_variant or bc_replays" > $O/parity.log 2>&1; tail -3 $O/parity.log
echo "== alone, solo warp (batch)"; timeout 300 python tools/wedge_bench.py 30 --handoff 2>&1 | grep wedge
echo "== alone, solo warp (nobatch)"; NPPSIM_LIB=$PWD/variants/nobatch.so timeout 300 python tools/wedge_bench.py 30 --handoff 2>&1 | grep wedge
echo "== alone, home warp (batch)"; timeout 300 python tools/wedge_bench.py 30 2>&1 | grep wedge
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel']), l.get('kernel'), l.get('handoff'))"; }
run() {  # name workload envs [env...]
  local name=$1 w=$2 n=$3; shift 3
  env "$@" timeout 300 python bench.py --workload $w --envs $n --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$name.log | tail -1 > $O/bench_$name.json
  summ "$name" < $O/bench_$name.json
}
NB=NPPSIM_LIB=$PWD/variants/nobatch.so
run c3_65k_batch    c3 65536
run c3_65k_nobatch  c3 65536 $NB
run c3_131k_batch   c3 131072
run c3_131k_nobatch c3 131072 $NB
run c4_131k_batch   c4 131072
run c4_131k_nobatch c4 131072 $NB
run c5_65k_batch    c5 65536
run c5_65k_nobatch  c5 65536 $NB
