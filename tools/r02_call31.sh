#!/bin/bash
# round 2, GPU call 31: whole GPU suite, smoke, default bench line, reference arm
mkdir -p gpurun_out/r02ae
O=gpurun_out/r02ae
echo "== GPU suite"; timeout 1700 python -m pytest tests -x -q -m gpu > $O/gpu_tests.log 2>&1; tail -3 $O/gpu_tests.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== default bench"; ( time timeout 1200 python bench.py > $O/bench_default.json 2>$O/bench_default.err ) 2>&1 | grep real; tail -c 3000 $O/bench_default.json
echo "== reference arm"; ( time timeout 1200 python bench.py --impl reference > $O/bench_reference.json 2>$O/bench_reference.err ) 2>&1 | grep real; tail -c 1500 $O/bench_reference.json
