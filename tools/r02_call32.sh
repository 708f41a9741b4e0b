#!/bin/bash
# round 2, GPU call 32: profiles of the final build -- ncu launch list + --set full of the step kernel (c3 65,536), timelines, c2 policy check
mkdir -p gpurun_out/r02af
O=gpurun_out/r02af
CMD="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-secondary --async-groups 0"
echo "== plain"; $CMD > $O/plain.log 2>$O/plain.err; tail -1 $O/plain.log | cut -c1-260
echo "== ncu launch list"; ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 200 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1; tail -3 $O/launches.csv | cut -c1-200
echo "== ncu full"; ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 320 -c 2 -f -o $O/r02b_step $CMD > $O/ncu_full.log 2>&1; ls -la $O/*.ncu-rep
ncu -i $O/r02b_step.ncu-rep --page raw --csv > $O/r02b_step_raw.csv 2>/dev/null; wc -c $O/r02b_step_raw.csv
python tools/ncu_lines.py $O/r02b_step.ncu-rep nclone_b200/libnppsim.so 60 > $O/r02b_step_lines.txt 2>&1; head -12 $O/r02b_step_lines.txt | cut -c1-200
P="python tools/phase_profile.py --lib=$PWD/variants/profile.so"
timeout 300 $P 65536 simplest_with_mines --engine-order=caller > $O/timeline_c3_65k.txt 2>&1; grep -A8 "^timeline" $O/timeline_c3_65k.txt; grep -A8 "^solo warps" $O/timeline_c3_65k.txt | cut -c1-220
timeout 300 $P 131072 simplest_with_mines > $O/timeline_c3_131k.txt 2>&1; grep -A6 "^timeline" $O/timeline_c3_131k.txt; grep -A5 "^solo warps" $O/timeline_c3_131k.txt | cut -c1-220
echo "== c2 policy"; timeout 300 python bench.py --workload c2 --steps 100 --warmup 10 --no-cpu-baseline --no-secondary --async-groups 0 2>/dev/null | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.read()); print('c2 value %.1fM' % (l['value'] / 1e6), l['kernel'], l['handoff'])"
echo "== wedge alone"; timeout 300 python tools/wedge_bench.py 30 --handoff 2>&1 | grep wedge; NPPSIM_LIB=$PWD/variants/profile.so timeout 300 python tools/wedge_bench.py 10 2>&1 | grep -v "^  phase\|^    dyn" > $O/wedge_cycles.txt; cat $O/wedge_cycles.txt
