#!/bin/bash
# round 2, GPU call 6: warp timelines with the solo-service log (profile build), reference arm on the box
mkdir -p gpurun_out/r02f
O=gpurun_out/r02f
ls oracle/_ref/nclone | head -3; nproc
P="python tools/phase_profile.py --lib=$PWD/variants/solo_prof.so"
timeout 300 $P 65536 simplest_with_mines > $O/timeline_c3_65k_on.txt 2>&1; tail -42 $O/timeline_c3_65k_on.txt
timeout 300 $P 65536 simplest_with_mines --handoff=off > $O/timeline_c3_65k_off.txt 2>&1
timeout 300 $P 131072 locked_doors > $O/timeline_c4_on.txt 2>&1; tail -34 $O/timeline_c4_on.txt
timeout 300 $P 131072 locked_doors --handoff=off > $O/timeline_c4_off.txt 2>&1; tail -34 $O/timeline_c4_off.txt | head -20
timeout 300 $P 65536 bc_replays > $O/timeline_c5_on.txt 2>&1; tail -34 $O/timeline_c5_on.txt | head -50
echo "== reference arm"; ( time timeout 600 python bench.py --impl reference --steps 10 --warmup 2 > $O/bench_reference.json 2>$O/bench_reference.err ) 2>&1 | grep real; cut -c1-1800 $O/bench_reference.json; tail -3 $O/bench_reference.err
echo "== default bench"; ( time timeout 900 python bench.py > $O/bench_default.json 2>$O/bench_default.err ) 2>&1 | grep real; python - <<'PY'
import json
l = json.loads(open("gpurun_out/r02f/bench_default.json").read().strip().split("\n")[-1])
print({k: l[k] for k in ("value", "ms_per_step", "cpu_baseline", "cpu_baseline_port") if k in l})
PY
tail -3 $O/bench_default.err
