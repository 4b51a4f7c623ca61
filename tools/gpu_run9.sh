#!/bin/bash
# 2 GPUs: multi-GPU parity test + tensor-parallel bench (both arms)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -q -x -s -p no:cacheprovider > gpurun_out/pytest_multi.txt 2>&1
echo "multi rc=$?" >> gpurun_out/pytest_multi.txt; tail -40 gpurun_out/pytest_multi.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_tp2.json 2> gpurun_out/bench_tp2.err
echo "tp2 rc=$?"; tail -c 1500 gpurun_out/bench_tp2.json; tail -5 gpurun_out/bench_tp2.err
