#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tcgen05 or prefill or splitk or pair or linear or postop" 2>&1 | tail -5
P="tc/pair256"
{
for m in 1024 4096 354; do
  python tools/chain.py --m $m --layers 4 --reps 5 --combo "" --combo "gate_up=$P,down=$P" --combo "qkv=$P,o=$P,gate_up=$P,down=$P"
  for sh in gate_up down; do python tools/chain.py --m $m --layers 8 --reps 5 --only $sh --combo "" --combo "$sh=$P" ; done
done
python tools/chain.py --m 64 --layers 4 --reps 5 --combo ""
python tools/chain.py --m 128 --layers 4 --reps 5 --combo ""
python tools/chain.py --m 256 --layers 4 --reps 5 --combo ""
} > gpurun_out/chain20.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain20.txt | cut -c1-120
