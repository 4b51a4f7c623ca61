#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/pytest_final.txt; tail -3 gpurun_out/pytest_final.txt
timeout 900 python bench.py > gpurun_out/bench_final_ours.json 2> gpurun_out/bench_final_ours.err; echo "ours rc=$?"
timeout 900 python bench.py --impl reference > gpurun_out/bench_final_ref.json 2> gpurun_out/bench_final_ref.err; echo "ref rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
