#!/bin/bash
# round 2, GPU call 13 (2 GPUs): full suite on one GPU, then the bench under torchrun on 2 ranks (per-rank attribution, secondary blocks)
mkdir -p gpurun_out/r02m
O=gpurun_out/r02m
echo "== pytest -m gpu"; CUDA_VISIBLE_DEVICES=0 timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; tail -3 $O/pytest_gpu.log
echo "== 2-GPU bench"; ( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 10 > $O/bench_2gpu.json 2>$O/bench_2gpu.err ) 2>&1 | grep real; tail -c 3000 $O/bench_2gpu.json; tail -5 $O/bench_2gpu.err
echo "== 2-GPU reference arm"; ( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > $O/bench_2gpu_ref.json 2>$O/bench_2gpu_ref.err ) 2>&1 | grep real; cut -c1-400 $O/bench_2gpu_ref.json
