#!/bin/bash
# round 2, GPU call 24: source-level stall samples of a lone wedged chain (arc wedge 0, line wedge 1) on a solo warp
mkdir -p gpurun_out/r02x
O=gpurun_out/r02x
for k in 0 1; do
  timeout 600 ncu --profile-from-start off --set full --import-source on --warp-sampling-interval 0 --clock-control none -k regex:step_kernel -c 1 -f -o $O/wedge$k python tools/wedge_ncu.py $k > $O/ncu_wedge$k.log 2>&1
  tail -2 $O/ncu_wedge$k.log
  python tools/ncu_lines.py $O/wedge$k.ncu-rep nclone_b200/libnppsim.so 90 > $O/wedge${k}_lines.txt 2>&1
  head -30 $O/wedge${k}_lines.txt
done
ls -la $O
