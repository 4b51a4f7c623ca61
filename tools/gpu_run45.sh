#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/prof_one.py --k 4096 --n 32768 --m 64 --cfg 0 --copies 3"
timeout 300 $CMD > gpurun_out/prof45_plain.log 2>&1 || { echo plain run failed; tail gpurun_out/prof45_plain.log; exit 1; }
timeout 800 ncu --set full --clock-control none --import-source on -k regex:prefill -o gpurun_out/splitk_m64_gate_up -f $CMD > gpurun_out/prof45_ncu.log 2>&1
tail -3 gpurun_out/prof45_ncu.log
