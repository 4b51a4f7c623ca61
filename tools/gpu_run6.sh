#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider -s > gpurun_out/pytest6.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest6.txt; tail -4 gpurun_out/pytest6.txt; grep "config 4\|tokens identical\|tok/s" gpurun_out/pytest6.txt | head
python bench.py > gpurun_out/bench6_ours.json 2> gpurun_out/bench6_ours.err; echo "bench ours rc=$?"
python bench.py --impl reference > gpurun_out/bench6_ref.json 2> gpurun_out/bench6_ref.err; echo "bench ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke6.txt 2>&1; echo "smoke rc=$?"
