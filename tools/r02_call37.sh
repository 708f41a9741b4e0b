#!/bin/bash
# round 2, GPU call 37: timelines of the end build (c3 65,536 / 131,072 with the corrected solo-log flags; c5)
mkdir -p gpurun_out/r02ak
O=gpurun_out/r02ak
P="python tools/phase_profile.py --lib=$PWD/variants/profile.so"
timeout 300 $P 65536 simplest_with_mines --engine-order=caller > $O/timeline_c3_65k.txt 2>&1; grep -A7 "^solo warps" $O/timeline_c3_65k.txt | cut -c1-250
timeout 300 $P 131072 simplest_with_mines > $O/timeline_c3_131k.txt 2>&1; grep -A5 "^solo warps" $O/timeline_c3_131k.txt | cut -c1-250
timeout 300 $P 65536 bc_replays > $O/timeline_c5_65k.txt 2>&1; head -40 $O/timeline_c5_65k.txt | cut -c1-200
