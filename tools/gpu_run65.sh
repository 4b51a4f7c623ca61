#!/bin/bash
mkdir -p gpurun_out
{
for m in 1 16; do python tools/chain.py --m $m --layers 8 --reps 5 --combo ""; done
P4="__half/nt2/tn4x64/kb8x16/w8/s5/true/tmem"
P2="__half/nt2/tn2x64/kb16x16/w8/s5/true/tmem"
python tools/chain.py --m 16 --layers 8 --reps 5 --combo "gate_up=$P4" --combo "qkv=$P2,o=$P2,gate_up=$P4,down=$P2" --combo "qkv=$P2,o=$P2,gate_up=$P2,down=$P2" --combo "qkv=$P2,o=$P2,gate_up=$P4"
} > gpurun_out/chain65.txt 2>&1
awk '{ printf "%s %s us %s\n", $1,$2,$6 }' gpurun_out/chain65.txt | cut -c1-150
