#!/bin/bash
mkdir -p gpurun_out
{
for m in 1024 354 4096 2048; do timeout 400 python tools/repro_chain.py --m $m --reps 150 2>&1 | grep -v "^frame\|^\[W\|^tiles" | head -30; done
} > gpurun_out/repro29.txt 2>&1
cut -c1-300 gpurun_out/repro29.txt | tail -40
