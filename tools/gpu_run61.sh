#!/bin/bash
mkdir -p gpurun_out
{
timeout 600 python tools/stress_prefill_chain.py --m 64 --passes 3000
timeout 600 python tools/stress_prefill_chain.py --m 128 --passes 2000
timeout 600 python tools/stress_prefill_chain.py --m 256 --passes 1500
timeout 600 python tools/stress_prefill_chain.py --m 4096 --passes 300
} > gpurun_out/stress61.txt 2>&1
cut -c1-300 gpurun_out/stress61.txt | tail -16
