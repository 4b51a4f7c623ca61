#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider > gpurun_out/pytest8.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest8.txt; tail -3 gpurun_out/pytest8.txt
{
for fx in 1 0; do
  echo "== FIXUP=$fx"
  for m in 1 4 16 32; do
    for sh in qkv o gate_up down; do
      B200_W4A16_FIXUP=$fx python tools/chain.py --m $m --layers 32 --only $sh --combo ""
    done
    B200_W4A16_FIXUP=$fx python tools/chain.py --m $m --layers 32 --combo ""
  done
done
} > gpurun_out/chain8.txt 2>&1
python tools/timeline.py --m 1 --layers 2 > gpurun_out/timeline8_m1.txt 2>&1
awk '{ if ($1=="==") print; else printf "%s %s us %s GB/s\n", $1,$2,$4 }' gpurun_out/chain8.txt
cat gpurun_out/timeline8_m1.txt
