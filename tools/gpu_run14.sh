#!/bin/bash
mkdir -p gpurun_out
P="tc/pair256"
{
for m in 1024 4096 354 2048; do
  python tools/chain.py --m $m --layers 4 --reps 5 --combo "" --combo "qkv=$P,o=$P,gate_up=$P,down=$P" --combo "gate_up=$P,down=$P"
  for sh in qkv o gate_up down; do python tools/chain.py --m $m --layers 8 --reps 5 --only $sh --combo "" --combo "$sh=$P"; done
done
} > gpurun_out/chain14.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain14.txt | cut -c1-120
