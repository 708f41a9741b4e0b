#!/bin/bash
# round 2, GPU call 42: observation row + mask stored by the step kernel into the pinned block (vs obs kernel + copy)
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "step_host or groups or observation" 2>&1 | tail -3
for z in 1 0; do NPPSIM_ZEROCOPY=$z python bench.py --no-cpu-baseline --no-secondary 2>/dev/null | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.read())
print($z, {k: (l[k] if not isinstance(l[k], dict) else {a: b for a, b in l[k].items() if a in ('value', 'ms_per_step')}) for k in ('value', 'ms_per_step', 'e2e', 'e2e_obs', 'gpu_launches')})"; done
