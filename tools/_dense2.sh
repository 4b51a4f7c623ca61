set -x
D=gpurun_out/h3
mkdir -p $D
timeout 600 python bench.py > $D/bench_ours.json 2> $D/bench_ours.err; tail -c 1200 $D/bench_ours.json; tail -3 $D/bench_ours.err
timeout 600 python bench.py --impl reference > $D/bench_reference.json 2> $D/bench_reference.err; tail -c 900 $D/bench_reference.json
python tools/bench_dense.py --once --widths 1,16,64 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $D/dense_launches.csv python tools/bench_dense.py --once --widths 1,16,64 > $D/ncu_launches.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:dense_stream -c 1 -o $D/dense_m1 python tools/bench_dense.py --once --widths 1 > $D/ncu_full.log 2>&1
ncu -i $D/dense_m1.ncu-rep --page details > $D/dense_m1_details.txt 2>/dev/null
ncu -i $D/dense_m1.ncu-rep --page raw --csv > $D/dense_m1_raw.csv 2>/dev/null
ncu --set full --clock-control none -k regex:dense_stream -c 1 -o $D/dense_m64 python tools/bench_dense.py --once --widths 64 > $D/ncu_full64.log 2>&1
ncu -i $D/dense_m64.ncu-rep --page details > $D/dense_m64_details.txt 2>/dev/null
ls -la $D
