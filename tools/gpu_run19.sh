#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/prof_one.py --k 4096 --n 32768 --m 1024 --cfg 1002,1013 --copies 2"
timeout 300 $CMD > gpurun_out/prof19_plain.log 2>&1 || { echo plain run failed; tail gpurun_out/prof19_plain.log; exit 1; }
timeout 800 ncu --set full --clock-control none --import-source on -k regex:prefill -o gpurun_out/lean_issuer_gate_up_m1024 -f $CMD > gpurun_out/prof19_ncu.log 2>&1
tail -3 gpurun_out/prof19_ncu.log
