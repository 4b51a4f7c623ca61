#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool racecheck --racecheck-report all --print-limit 40 python tools/sanitize_case.py > gpurun_out/sanitize_racecheck.log 2>&1; echo "racecheck rc=$?"
grep -c "ok " gpurun_out/sanitize_racecheck.log; grep -i "RACECHECK SUMMARY\|hazard" gpurun_out/sanitize_racecheck.log | sort | uniq -c | sort -rn | head -12 | cut -c1-200
P="tc/pair256"
{
for m in 300 384 512 768 1024 2048 4096; do
  for sh in qkv o gate_up down; do python tools/chain.py --m $m --layers 8 --reps 5 --only $sh --combo "" --combo "$sh=$P" --combo "$sh=tc/bn128/bm256"; done
done
} > gpurun_out/chain31.txt 2>&1
awk '{ printf "%s %s us %s\n", $1,$2,$6 }' gpurun_out/chain31.txt | cut -c1-100 | tail -30
