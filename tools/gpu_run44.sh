#!/bin/bash
mkdir -p gpurun_out
{
for sh in qkv o gate_up down; do python tools/chain.py --m 64 --layers 8 --reps 5 --only $sh --combo "" --combo "$sh=tc/bn128/bm64/splitk/s1" --combo "$sh=tc/bn128/bm64/splitk/s2" --combo "$sh=tc/bn128/bm64/splitk/s4" --combo "$sh=tc/bn128/bm64/splitk/s8" --combo "$sh=tc/bn128/bm128"; done
python tools/chain.py --m 64 --layers 8 --reps 5 --combo ""
} > gpurun_out/chain44.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s %s\n", $1,$2,$4,$6,$7 }' gpurun_out/chain44.txt | cut -c1-150
