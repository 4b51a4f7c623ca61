#!/bin/bash
mkdir -p gpurun_out
python tools/timeline.py --m 1 --layers 2 > gpurun_out/timeline4_m1.txt 2>&1
python tools/timeline.py --m 16 --layers 2 > gpurun_out/timeline4_m16.txt 2>&1
{
for m in 1 16; do
  nt=$([ $m -le 8 ] && echo 1 || echo 2)
  H="__half/nt${nt}"
  for sh in qkv o gate_up down; do
    python tools/chain.py --m $m --layers 32 --only $sh --combo "" --combo "$sh=$H/tn4x64/kb4x16/w8" --combo "$sh=$H/tn1x64/kb32x16/w8" --combo "$sh=$H/tn2x64/kb16x16/w8"
  done
done
} > gpurun_out/chain4.txt 2>&1
python -m pytest tests/test_gpu_cpmcu_tokens.py -m gpu -x -q -s -p no:cacheprovider > gpurun_out/pytest_cpmcu.txt 2>&1
echo "cpmcu rc=$?"; tail -30 gpurun_out/pytest_cpmcu.txt
cat gpurun_out/timeline4_m1.txt gpurun_out/timeline4_m16.txt
awk '{ printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain4.txt
