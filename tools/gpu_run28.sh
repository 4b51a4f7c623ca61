#!/bin/bash
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 600 python bench.py --no-tp1 --no-cpu-baseline "$@" > gpurun_out/b28_$tag.json 2> gpurun_out/b28_$tag.err; echo "$tag rc=$? $(grep -c prefill gpurun_out/b28_$tag.json) $(grep -m1 'Error\|error' gpurun_out/b28_$tag.err | cut -c1-150)"; }
run m1024 --widths 1 --layers 2 --steps 3 --warmup 3 --no-lm-head --prefill-m 1024
run m4096 --widths 1 --layers 2 --steps 3 --warmup 3 --no-lm-head --prefill-m 4096
run m354 --widths 1 --layers 2 --steps 3 --warmup 3 --no-lm-head --prefill-m 354
run all --widths 1 --layers 2 --steps 3 --warmup 3 --no-lm-head
run full
