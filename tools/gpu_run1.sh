#!/bin/bash
# round-2 GPU call 1: correctness of the new decode reduction + first numbers + geometry experiments
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --maxfail=15 -p no:cacheprovider > gpurun_out/pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest.txt
tail -5 gpurun_out/pytest.txt
python bench.py --no-tp1 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err
echo "bench ours rc=$?"
python bench.py --impl reference --no-tp1 --no-cpu-baseline > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "bench ref rc=$?"
for m in 1 16; do
  nt=$([ $m -le 8 ] && echo 1 || echo 2)
  H="__half/nt${nt}"
  python tools/chain.py --m $m --layers 16 \
    --combo "" \
    --combo "qkv=$H/tn4x64/kb8x16/w16,o=$H/tn4x64/kb8x16/w16,gate_up=$H/tn4x64/kb8x16/w16,down=$H/tn4x64/kb8x16/w16" \
    --combo "qkv=$H/tn4x64/kb16x16/w16,o=$H/tn4x64/kb16x16/w16,gate_up=$H/tn4x64/kb16x16/w16,down=$H/tn4x64/kb16x16/w16" \
    --combo "gate_up=$H/tn4x64/kb8x16/w16,down=$H/tn4x64/kb8x16/w16" \
    --combo "gate_up=$H/tn4x64/kb16x16/w16,down=$H/tn4x64/kb16x16/w16" \
    --combo "qkv=$H/tn2x64/kb16x16/w8,o=$H/tn2x64/kb16x16/w8,gate_up=$H/tn2x64/kb16x16/w8,down=$H/tn2x64/kb16x16/w8" \
    >> gpurun_out/chain.txt 2>&1
  for sh in qkv o gate_up down; do
    python tools/chain.py --m $m --layers 32 --only $sh --combo "" \
      --combo "$sh=$H/tn4x64/kb8x16/w16" --combo "$sh=$H/tn4x64/kb16x16/w16" --combo "$sh=$H/tn2x64/kb16x16/w8" >> gpurun_out/chain.txt 2>&1
  done
done
python tools/timeline.py --m 1 --layers 2 > gpurun_out/timeline_m1.txt 2>&1
python tools/timeline.py --m 16 --layers 2 > gpurun_out/timeline_m16.txt 2>&1
cat gpurun_out/chain.txt
