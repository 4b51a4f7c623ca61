#!/bin/bash
mkdir -p gpurun_out
{
for m in 1 4 8 16 32; do
  nt=1; [ $m -gt 8 ] && nt=2; [ $m -gt 16 ] && nt=4
  P="__half/nt$nt/tn4x64/kb8x16/w8/s5/true/tmem"
  python tools/chain.py --m $m --layers 8 --reps 5 --combo "" --combo "qkv=$P,o=$P,gate_up=$P,down=$P"
  for sh in qkv gate_up down; do python tools/chain.py --m $m --layers 8 --reps 5 --only $sh --combo "" --combo "$sh=$P"; done
done
} > gpurun_out/chain39.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain39.txt | cut -c1-110
