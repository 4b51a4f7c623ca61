#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/pytest52.txt; tail -4 gpurun_out/pytest30.txt
timeout 900 python bench.py > gpurun_out/bench52_ours.json 2> gpurun_out/bench52_ours.err; echo "ours rc=$?"; tail -c 300 gpurun_out/bench52_ours.json
timeout 900 python bench.py --impl reference > gpurun_out/bench52_ref.json 2> gpurun_out/bench52_ref.err; echo "ref rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
