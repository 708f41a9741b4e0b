#!/bin/bash
# round 2, GPU call 4: per-lane loop v3 (prefilter to <= 8 candidates, 4 doubles each) vs cooperative loop, both with handoff
mkdir -p gpurun_out/r02d
O=gpurun_out/r02d
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('handoff'))"; }
for so in solo lane lane_prof; do echo "== wedge $so"; NPPSIM_LIB=$PWD/variants/$so.so timeout 300 python tools/wedge_bench.py 20 2>&1 | grep -v "phase cycles\|dyn_move" | tee $O/wedge_$so.txt; done
echo "== parity (lane.so)"; NPPSIM_LIB=$PWD/variants/lane.so timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "wedged or large_random or full_size or basic_variant or doors" > $O/parity_lane.log 2>&1; tail -3 $O/parity_lane.log
for so in lane; do
  for cfg in "c3 65536 scattered" "c3 131072 adjacent" "c4 131072 scattered" "c4 131072 adjacent" "c5 65536 adjacent"; do
    set -- $cfg
    NPPSIM_LIB=$PWD/variants/$so.so timeout 300 python bench.py --workload $1 --envs $2 --placement $3 --steps 40 --warmup 5 --no-cpu-baseline --async-groups 0 2>$O/err_${so}_$1_$2_$3.log | tail -1 > $O/bench_${so}_$1_$2_$3.json
    summ "$so $1 $2 $3" < $O/bench_${so}_$1_$2_$3.json
  done
done
