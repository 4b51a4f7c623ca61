#!/bin/bash
mkdir -p gpurun_out
{
for i in 1 2 3; do timeout 300 python tools/stress_prefill_chain.py --m 354 --passes 150; done
timeout 300 python tools/stress_prefill_chain.py --m 1024 --passes 100
timeout 300 python tools/stress_prefill_chain.py --m 354 --passes 150 --no-pdl
} > gpurun_out/stress54.txt 2>&1
cut -c1-330 gpurun_out/stress54.txt | tail -40
