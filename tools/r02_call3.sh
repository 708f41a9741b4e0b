#!/bin/bash
# round 2, GPU call 3: solo blocks + in-launch handoff of wedged envs
mkdir -p gpurun_out/r02c
O=gpurun_out/r02c
echo "== parity (default lib, handoff on from 1,024 envs)"; timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/parity.log 2>&1; tail -3 $O/parity.log
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM step_kernel %.3f ms reset %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel'], l['kernel_ms']['reset_kernel']), l.get('handoff'))"; }
for so in solo; do
  for cfg in "c3 65536 scattered" "c3 65536 adjacent" "c3 131072 scattered" "c3 131072 adjacent" "c4 131072 scattered" "c4 131072 adjacent" "c5 65536 adjacent" "c2 4096 scattered"; do
    set -- $cfg
    NPPSIM_LIB=$PWD/variants/$so.so timeout 300 python bench.py --workload $1 --envs $2 --placement $3 --steps 40 --warmup 5 --no-cpu-baseline --async-groups 0 2>$O/err_${so}_$1_$2_$3.log | tail -1 > $O/bench_${so}_$1_$2_$3.json
    summ "$so $1 $2 $3" < $O/bench_${so}_$1_$2_$3.json
  done
done
echo "== fp64 issue"; timeout 60 tools/microbench/fp64_issue | tee $O/compute_peaks.json
echo "== phase profile c3 solo"; timeout 300 python tools/phase_profile.py 65536 --lib=$PWD/variants/solo_prof.so > $O/phase_c3_solo.txt 2>&1; head -16 $O/phase_c3_solo.txt; grep -A22 "timeline" $O/phase_c3_solo.txt
