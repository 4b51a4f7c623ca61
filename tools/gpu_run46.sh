#!/bin/bash
mkdir -p gpurun_out
{
for m in 1 4 16; do
  python tools/chain.py --m $m --layers 8 --reps 5 --combo ""
  for sh in qkv o gate_up down; do python tools/chain.py --m $m --layers 8 --reps 5 --only $sh --combo ""; done
done
} > gpurun_out/chain46.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain46.txt | cut -c1-110
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "split_column or decode or contract or gated" 2>&1 | tail -3
