#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "cta_pair" > gpurun_out/pytest13.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest13.txt; tail -25 gpurun_out/pytest13.txt
