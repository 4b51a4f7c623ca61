#!/bin/bash
mkdir -p gpurun_out
{
for m in 16 32 9; do
  nt=2; [ $m -gt 16 ] && nt=4
  P2="__half/nt$nt/tn2x64/kb16x16/w8/s5/true/tmem"
  P4="__half/nt$nt/tn4x64/kb8x16/w8/s5/true/tmem"
  python tools/chain.py --m $m --layers 8 --reps 5 --combo "" --combo "qkv=$P2,o=$P2,gate_up=$P4,down=$P2" --combo "qkv=$P2,o=$P2,gate_up=$P2,down=$P2"
done
} > gpurun_out/chain51.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain51.txt | cut -c1-130
