set -x
mkdir -p gpurun_out/h1
timeout 900 python -m pytest tests/test_gpu_dense.py -x -q > gpurun_out/h1/pytest_dense.txt 2>&1; tail -n 15 gpurun_out/h1/pytest_dense.txt
timeout 300 python tools/bench_dense.py > gpurun_out/h1/bench_dense.json 2> gpurun_out/h1/bench_dense.err; cat gpurun_out/h1/bench_dense.json; tail -3 gpurun_out/h1/bench_dense.err
