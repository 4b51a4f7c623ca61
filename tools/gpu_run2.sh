#!/bin/bash
# round-2 GPU call 2: tail fix check + what bounds the loop (lab build: wrong results on purpose, timing only)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "not full_size and not exhaustive" > gpurun_out/pytest2.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest2.txt; tail -3 gpurun_out/pytest2.txt
H=__half/nt1
{
echo "== product build, m=1 per-shape chains"
for sh in qkv o gate_up down; do
  python tools/chain.py --m 1 --layers 32 --only $sh --combo "" --combo "$sh=$H/tn4x64/kb8x16/w8/s10" --combo "$sh=$H/tn4x64/kb16x16/w16"
done
python tools/chain.py --m 1 --layers 32 --combo ""
python tools/chain.py --m 16 --layers 32 --combo "" --combo "qkv=__half/nt2/tn2x64/kb16x16/w8,o=__half/nt2/tn2x64/kb16x16/w8,gate_up=__half/nt2/tn4x64/kb8x16/w16,down=__half/nt2/tn2x64/kb16x16/w8"
export B200_W4A16_LIB=$PWD/cpmcu_b200/libb200_lab.so
for lab in 0 1 2 3 4 7; do
  echo "== LAB=$lab (1: no per-stage A, 2: no per-stage S, 4: no math)"
  for sh in gate_up down qkv; do
    B200_LAB=$lab python tools/chain.py --m 1 --layers 32 --only $sh --combo "" --combo "$sh=$H/tn4x64/kb8x16/w8/s10" --combo "$sh=$H/tn4x64/kb16x16/w16"
  done
done
} > gpurun_out/lab.txt 2>&1
unset B200_W4A16_LIB
python tools/timeline.py --m 1 --layers 2 > gpurun_out/timeline2_m1.txt 2>&1
python tools/timeline.py --m 16 --layers 2 > gpurun_out/timeline2_m16.txt 2>&1
cut -c1-110 gpurun_out/lab.txt
