#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/check_tc.py > gpurun_out/check_tc.txt 2>&1; tail -20 gpurun_out/check_tc.txt | cut -c1-200
