#!/bin/bash
mkdir -p gpurun_out
{
timeout 400 python tools/stress_prefill_chain.py --m 354 --passes 3000
timeout 400 python tools/stress_prefill_chain.py --m 354 --passes 3000 --force 1013
timeout 400 python tools/stress_prefill_chain.py --m 354 --passes 3000 --force 1004
timeout 400 python tools/stress_prefill_chain.py --m 1024 --passes 1500
} > gpurun_out/stress55.txt 2>&1
cut -c1-330 gpurun_out/stress55.txt | tail -40
