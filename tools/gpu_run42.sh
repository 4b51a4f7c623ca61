#!/bin/bash
mkdir -p gpurun_out
{
P="__half/nt2/tn4x64/kb8x16/w8/s5/true/tmem"
Q="__half/nt2/tn4x64/kb8x16/w8/s5/true"
echo "== pdl"; python tools/chain.py --m 16 --layers 8 --reps 5 --combo "" --combo "gate_up=$P" --combo "gate_up=$Q" --combo "qkv=$P,o=$P,gate_up=$P,down=$P"
echo "== no pdl"; python tools/chain.py --m 16 --layers 8 --reps 5 --no-pdl --combo "" --combo "gate_up=$P" --combo "gate_up=$Q" --combo "qkv=$P,o=$P,gate_up=$P,down=$P"
} > gpurun_out/chain42.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain42.txt | cut -c1-110
