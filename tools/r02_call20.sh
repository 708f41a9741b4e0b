#!/bin/bash
# round 2, GPU call 20: solo-service phases on solo SMs vs the same wedges alone on the GPU
mkdir -p gpurun_out/r02t
O=gpurun_out/r02t
P="python tools/phase_profile.py --lib=$PWD/variants/profile.so"
NPPSIM_SOLO_SM=1 timeout 600 $P 65536 --handoff=on --engine-order=caller > $O/timeline_c3_65k_sm.txt 2>&1; tail -36 $O/timeline_c3_65k_sm.txt
NPPSIM_LIB=$PWD/variants/profile.so timeout 300 python tools/wedge_bench.py 30 > $O/wedge_alone.txt 2>&1; cat $O/wedge_alone.txt
