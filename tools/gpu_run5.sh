#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider > gpurun_out/pytest5.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest5.txt; tail -3 gpurun_out/pytest5.txt
{
for m in 1 4 16 32; do
  nt=$([ $m -le 8 ] && echo 1 || ([ $m -le 16 ] && echo 2 || echo 4))
  H="__half/nt${nt}"
  for sh in qkv o gate_up down; do
    python tools/chain.py --m $m --layers 32 --only $sh --combo "" --combo "$sh=$H/tn2x64/kb16x16/w8" --combo "$sh=$H/tn1x64/kb32x16/w8" --combo "$sh=$H/tn4x64/kb8x16/w16"
  done
  python tools/chain.py --m $m --layers 32 --combo "" --combo "qkv=$H/tn2x64/kb16x16/w8,o=$H/tn2x64/kb16x16/w8,gate_up=$H/tn2x64/kb16x16/w8,down=$H/tn2x64/kb16x16/w8"
done
} > gpurun_out/chain5.txt 2>&1
python tools/timeline.py --m 1 --layers 2 > gpurun_out/timeline5_m1.txt 2>&1
python tools/timeline.py --m 1 --layers 2 --cfg-name __half/nt1/tn2x64/kb16x16/w8 > gpurun_out/timeline5_m1_tn2.txt 2>&1
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain5.txt | cut -c1-150
cat gpurun_out/timeline5_m1.txt gpurun_out/timeline5_m1_tn2.txt
