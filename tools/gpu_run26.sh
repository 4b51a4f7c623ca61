#!/bin/bash
mkdir -p gpurun_out
{
timeout 120 python tools/repro_prefill.py --launches 1
timeout 120 python tools/repro_prefill.py --launches 3
timeout 120 python tools/repro_prefill.py --launches 3 --no-pdl
timeout 120 python tools/repro_prefill.py --launches 3 --k 4096 --n 4096
timeout 120 python tools/repro_prefill.py --launches 3 --cfg 1002
timeout 120 python tools/repro_prefill.py --launches 3 --m 768
} > gpurun_out/repro26.txt 2>&1
cut -c1-400 gpurun_out/repro26.txt | tail -40
