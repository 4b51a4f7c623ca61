#!/bin/bash
mkdir -p gpurun_out
{
timeout 400 python tools/stress_prefill_chain.py --m 354 --passes 4000 --force 1013
timeout 400 python tools/stress_prefill_chain.py --m 1024 --passes 2500
timeout 400 python tools/stress_prefill_chain.py --m 354 --passes 3000
} > gpurun_out/stress57.txt 2>&1
cut -c1-330 gpurun_out/stress57.txt | tail -30
P="tc/pair256"
{
for m in 1024 4096; do
  python tools/chain.py --m $m --layers 4 --reps 5 --combo "" --combo "qkv=tc/bn128/bm256,o=tc/bn128/bm256,gate_up=tc/bn128/bm256,down=tc/bn128/bm256"
  for sh in gate_up down; do python tools/chain.py --m $m --layers 8 --reps 5 --only $sh --combo "$sh=$P" --combo "$sh=tc/bn128/bm256"; done
done
} > gpurun_out/chain57.txt 2>&1
awk '{ printf "%s %s us %s\n", $1,$2,$6 }' gpurun_out/chain57.txt | cut -c1-100
