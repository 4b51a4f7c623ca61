#!/bin/bash
# 8 GPUs: multi-GPU parity test (4 ranks) + tensor-parallel bench at 8 and 4 ranks
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -q -x -s -p no:cacheprovider > gpurun_out/pytest_multi8.txt 2>&1
echo "multi rc=$?" >> gpurun_out/pytest_multi8.txt; tail -6 gpurun_out/pytest_multi8.txt
for n in 8 4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_tp$n.json 2> gpurun_out/bench_tp$n.err
echo "tp$n rc=$?"; tail -c 900 gpurun_out/bench_tp$n.json; tail -3 gpurun_out/bench_tp$n.err
done
