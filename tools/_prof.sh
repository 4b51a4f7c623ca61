set -x
mkdir -p gpurun_out/g1
python -m pytest tests -m gpu -q > gpurun_out/g1/pytest_gpu.txt 2>&1; tail -n 2 gpurun_out/g1/pytest_gpu.txt
python bench.py --impl reference > gpurun_out/g1/bench_reference.json 2> gpurun_out/g1/bench_reference.err
python bench.py > gpurun_out/g1/bench_ours.json 2> gpurun_out/g1/bench_ours.err
# launch list of the same command (short form), then one full capture of the dominant kernel
python bench.py --steps 1 --warmup 3 --layers 4 --widths 1,16 --no-detail --no-cpu-baseline > gpurun_out/g1/bench_short.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/g1/launches.csv python bench.py --steps 1 --warmup 3 --layers 4 --widths 1,16 --no-detail --no-cpu-baseline > gpurun_out/g1/ncu_launches.log 2>&1
python tools/prof_one.py --k 4096 --n 32768 --m 1 --copies 4 --pdl && \
ncu --set full --import-source on --clock-control none -k regex:w4a16_decode -c 2 -o gpurun_out/g1/gateup_m1 python tools/prof_one.py --k 4096 --n 32768 --m 1 --copies 4 > gpurun_out/g1/ncu_full.log 2>&1
ncu -i gpurun_out/g1/gateup_m1.ncu-rep --page details > gpurun_out/g1/gateup_m1_details.txt 2>/dev/null
ncu -i gpurun_out/g1/gateup_m1.ncu-rep --page raw --csv > gpurun_out/g1/gateup_m1_raw.csv 2>/dev/null
ncu --set full --clock-control none -k regex:w4a16_decode -c 1 -o gpurun_out/g1/gateup_m16 python tools/prof_one.py --k 4096 --n 32768 --m 16 --copies 4 > gpurun_out/g1/ncu_full16.log 2>&1
ncu -i gpurun_out/g1/gateup_m16.ncu-rep --page raw --csv > gpurun_out/g1/gateup_m16_raw.csv 2>/dev/null
ls -la gpurun_out/g1
