#!/bin/bash
mkdir -p gpurun_out
{
timeout 600 python tools/stress_prefill_chain.py --m 64 --passes 3000 --ref 0
timeout 600 python tools/stress_prefill_chain.py --m 128 --passes 2000 --ref 0
timeout 600 python tools/stress_prefill_chain.py --m 256 --passes 1500 --ref 0
} > gpurun_out/stress62.txt 2>&1
cut -c1-300 gpurun_out/stress62.txt | tail -12
