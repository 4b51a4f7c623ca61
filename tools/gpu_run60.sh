#!/bin/bash
mkdir -p gpurun_out
{
timeout 600 python tools/stress_prefill_chain.py --m 354 --passes 4000 --force 1013
timeout 600 python tools/stress_prefill_chain.py --m 1024 --passes 3000
timeout 600 python tools/stress_decode_repeat.py --m 32 --passes 3000
timeout 600 python tools/stress_decode_repeat.py --m 20 --passes 2000
timeout 600 python tools/stress_decode_repeat.py --m 1 --passes 2000
} > gpurun_out/stress60.txt 2>&1
cut -c1-300 gpurun_out/stress60.txt | tail -16
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/pytest60.txt; tail -4 gpurun_out/pytest60.txt
timeout 900 python bench.py > gpurun_out/bench60_ours.json 2> gpurun_out/bench60_ours.err; echo "ours rc=$?"
timeout 900 python bench.py --impl reference > gpurun_out/bench60_ref.json 2> gpurun_out/bench60_ref.err; echo "ref rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
