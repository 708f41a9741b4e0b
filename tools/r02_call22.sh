#!/bin/bash
# round 2, GPU call 22: in-loop cycles of solo services on solo SMs; the same wedges alone on the GPU in their home warp and on a solo warp
mkdir -p gpurun_out/r02v
O=gpurun_out/r02v
P="python tools/phase_profile.py --lib=$PWD/variants/profile.so"
NPPSIM_SOLO_SM=1 timeout 600 $P 65536 --handoff=on --engine-order=caller > $O/timeline_c3_65k_sm.txt 2>&1; tail -26 $O/timeline_c3_65k_sm.txt
echo "== alone, home warp"; timeout 300 python tools/wedge_bench.py 30 > $O/wedge_alone_home.txt 2>&1; cat $O/wedge_alone_home.txt
echo "== alone, solo warp"; timeout 300 python tools/wedge_bench.py 30 --handoff > $O/wedge_alone_solo.txt 2>&1; cat $O/wedge_alone_solo.txt
echo "== alone, solo warp, profile"; NPPSIM_LIB=$PWD/variants/profile.so timeout 300 python tools/wedge_bench.py 30 --handoff > $O/wedge_alone_solo_prof.txt 2>&1; cat $O/wedge_alone_solo_prof.txt
