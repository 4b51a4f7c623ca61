#!/bin/bash
mkdir -p gpurun_out
TC_TILE=tn2x64/kb16x16 timeout 300 python tools/check_tc.py > gpurun_out/check_tc2.txt 2>&1; tail -18 gpurun_out/check_tc2.txt | cut -c1-160
{
for m in 16 32 12; do
  nt=2; [ $m -gt 16 ] && nt=4
  P="__half/nt$nt/tn2x64/kb16x16/w8/s5/true/tmem"
  python tools/chain.py --m $m --layers 8 --reps 5 --combo "" --combo "qkv=$P,o=$P,gate_up=$P,down=$P"
  for sh in qkv o gate_up down; do python tools/chain.py --m $m --layers 8 --reps 5 --only $sh --combo "" --combo "$sh=$P"; done
done
} > gpurun_out/chain50.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain50.txt | cut -c1-100
