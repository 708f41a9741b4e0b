#!/bin/bash
# round 2, GPU call 33: nppsim_step_host with zero-copy result stores (pinned block) vs the staged path
mkdir -p gpurun_out/r02ag
O=gpurun_out/r02ag
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "step_host or handoff or env_view or groups" > $O/parity.log 2>&1; tail -3 $O/parity.log
summ() { python -c "
import sys, json
l = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('$1', 'value %.1fM e2e %.1fM (%.3f ms) e2e_obs %.1fM step_kernel %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['e2e']['ms_per_step'], l.get('e2e_obs', {}).get('value', 0) / 1e6, l['kernel_ms']['step_kernel']), l.get('kernel'))"; }
run() {  # name workload envs [env...]
  local name=$1 w=$2 n=$3; shift 3
  env "$@" timeout 300 python bench.py --workload $w --envs $n --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>$O/err_$name.log | tail -1 > $O/bench_$name.json
  summ "$name" < $O/bench_$name.json
}
run c3_65k_zc      c3 65536
run c3_65k_staged  c3 65536 NPPSIM_NO_ZEROCOPY=1
run c3_131k_zc     c3 131072
run c3_131k_staged c3 131072 NPPSIM_NO_ZEROCOPY=1
run c4_131k_zc     c4 131072
run c4_131k_staged c4 131072 NPPSIM_NO_ZEROCOPY=1
run c5_65k_zc      c5 65536
run c5_65k_staged  c5 65536 NPPSIM_NO_ZEROCOPY=1
run c2_4k_zc       c2 4096
run c2_4k_staged   c2 4096 NPPSIM_NO_ZEROCOPY=1
