#!/bin/bash
# A/B of kernel builds on the GPU box: tools/ab.sh variants/a.so variants/b.so ...   (bench line summary per build)
for so in "$@"; do
  for envs in 65536 131072; do
    NPPSIM_LIB=$PWD/$so python bench.py --steps 40 --warmup 5 --no-cpu-baseline --envs $envs 2>&1 | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.read())
print('$so', $envs, 'value %.1fM e2e %.1fM step_kernel %.3f ms' % (l['value'] / 1e6, l['e2e']['value'] / 1e6, l['kernel_ms']['step_kernel']))"
  done
done
