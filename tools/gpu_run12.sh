#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider > gpurun_out/pytest12.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest12.txt; tail -3 gpurun_out/pytest12.txt
H=__half/nt2
{
python tools/chain.py --m 16 --layers 32 --combo "" \
  --combo "qkv=$H/tn4x64/kb8x16/w8/s5,o=$H/tn4x64/kb8x16/w8/s5,gate_up=$H/tn4x64/kb8x16/w8/s5,down=$H/tn4x64/kb8x16/w8/s5" \
  --combo "qkv=$H/tn2x64/kb16x16/w8,o=$H/tn2x64/kb16x16/w8,gate_up=$H/tn2x64/kb16x16/w8,down=$H/tn2x64/kb16x16/w8" \
  --combo "qkv=$H/tn2x64/kb16x16/w8,o=$H/tn2x64/kb16x16/w8,down=$H/tn2x64/kb16x16/w8" \
  --combo "down=$H/tn2x64/kb16x16/w8" --combo "qkv=$H/tn2x64/kb16x16/w8,o=$H/tn2x64/kb16x16/w8"
export B200_W4A16_LIB=$PWD/cpmcu_b200/libb200_lab.so
for lab in 0 8; do
  echo "== LAB=$lab (8: release-only ticket)"
  for m in 1 16; do
    B200_LAB=$lab python tools/chain.py --m $m --layers 32 --combo ""
    for sh in qkv down; do B200_LAB=$lab python tools/chain.py --m $m --layers 32 --only $sh --combo ""; done
  done
done
} > gpurun_out/chain12.txt 2>&1
awk '{ if ($1=="==") print; else printf "%s %s us %s GB/s  %s\n", $1,$2,$4,$6 }' gpurun_out/chain12.txt | cut -c1-200
