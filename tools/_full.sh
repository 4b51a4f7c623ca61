set -x
D=gpurun_out/${1:-h2}
mkdir -p $D
timeout 1500 python -m pytest tests -m gpu -q > $D/pytest_gpu.txt 2>&1; tail -n 3 $D/pytest_gpu.txt
timeout 600 python bench.py --impl reference > $D/bench_reference.json 2> $D/bench_reference.err; tail -c 600 $D/bench_reference.json
timeout 600 python bench.py > $D/bench_ours.json 2> $D/bench_ours.err; tail -c 1500 $D/bench_ours.json; tail -3 $D/bench_ours.err
python -c "import __graft_entry__ as g; g.smoke()" > $D/smoke.txt 2>&1; tail -2 $D/smoke.txt
