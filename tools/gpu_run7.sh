#!/bin/bash
mkdir -p gpurun_out
python tools/timeline.py --m 1 --layers 2 > gpurun_out/timeline7_m1.txt 2>&1
python tools/timeline.py --m 16 --layers 2 > gpurun_out/timeline7_m16.txt 2>&1
CMD="python tools/prof_one.py --k 4096 --n 32768 --m 4096 --copies 2"
$CMD > gpurun_out/plain_prefill.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:w4a16_prefill -s 2 -c 1 -o gpurun_out/prof_prefill_m4096 $CMD > gpurun_out/ncu_prefill.log 2>&1
echo "ncu rc=$?"
cat gpurun_out/timeline7_m1.txt gpurun_out/timeline7_m16.txt
