#!/bin/bash
# round 2, GPU call 19: per-warp timeline + solo-service log with solo SMs on / off
mkdir -p gpurun_out/r02s
O=gpurun_out/r02s
P="python tools/phase_profile.py --lib=$PWD/variants/profile.so"
NPPSIM_SOLO_SM=1 timeout 600 $P 65536 --handoff=on --engine-order=caller > $O/timeline_c3_65k_sm.txt 2>&1; tail -45 $O/timeline_c3_65k_sm.txt
NPPSIM_SOLO_SM=0 timeout 600 $P 65536 --handoff=on --engine-order=caller > $O/timeline_c3_65k_front.txt 2>&1; tail -40 $O/timeline_c3_65k_front.txt
