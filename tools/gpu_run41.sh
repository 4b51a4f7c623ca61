#!/bin/bash
mkdir -p gpurun_out
{
for m in 16 32 12; do
  nt=2; [ $m -gt 16 ] && nt=4
  P="__half/nt$nt/tn4x64/kb8x16/w8/s5/true/tmem"
  python tools/chain.py --m $m --layers 8 --reps 5 --combo "" --combo "gate_up=$P" --combo "gate_up=$P,down=$P"
done
} > gpurun_out/chain41.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain41.txt | cut -c1-110
