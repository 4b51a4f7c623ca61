#!/bin/bash
# round-2 GPU call 3: ncu on the decode loop (large shape: loop-dominated), 8 vs 16 consumer warps
mkdir -p gpurun_out
for cfg in "__half/nt1/tn4x64/kb8x16/w8/s5" "__half/nt1/tn4x64/kb8x16/w16/s5" "__half/nt2/tn4x64/kb8x16/w8/s5"; do
  tag=$(echo $cfg | tr '/' '_')
  m=1; case $cfg in *nt2*) m=16;; esac
  CMD="python tools/prof_one.py --k 8192 --n 57344 --m $m --copies 3 --cfg-name $cfg"
  $CMD > gpurun_out/plain_$tag.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:w4a16_decode -s 3 -c 1 -o gpurun_out/prof_$tag $CMD > gpurun_out/ncu_$tag.log 2>&1
  echo "$cfg rc=$?"
done
ls -la gpurun_out/*.ncu-rep
