D=gpurun_out/${1:-j1}
mkdir -p $D
timeout 1500 python -m pytest tests -m gpu -q > $D/pytest_gpu.txt 2>&1; tail -n 2 $D/pytest_gpu.txt
timeout 600 python bench.py --impl reference > $D/bench_reference.json 2> $D/bench_reference.err
timeout 600 python bench.py > $D/bench_ours.json 2> $D/bench_ours.err; tail -2 $D/bench_ours.err
python -c "import __graft_entry__ as g; g.smoke()" > $D/smoke.txt 2>&1; tail -1 $D/smoke.txt
python tools/bench_dense.py --n 32768 --k 4096 > $D/dense_frspec_head.json 2>&1
python tools/bench_dense.py --n 4096 --k 8192 --widths 1,4,16 > $D/dense_eagle_fc.json 2>&1
timeout 300 python bench.py --steps 1 --warmup 3 --layers 4 --widths 1,16 --no-detail --no-cpu-baseline --no-lm-head > $D/bench_short.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $D/launches.csv python bench.py --steps 1 --warmup 3 --layers 4 --widths 1,16 --no-detail --no-cpu-baseline --no-lm-head > $D/ncu_launches.log 2>&1
python - <<P
import json
for f in ("bench_ours","bench_reference"):
    d=json.load(open("$D/"+f+".json"))
    print(f, d["value"], d["ms_per_step"], {m:v["ms"] for m,v in d["per_width"].items()}, "e2e", d["e2e"]["value"], "roof", d["roofline"]["frac"], {m:v["us"] for m,v in d["lm_head"]["per_width"].items()})
for f in ("dense_frspec_head","dense_eagle_fc"):
    d=json.loads(open("$D/"+f+".json").read().strip().splitlines()[-1])
    print(f, {m:(v["ours"]["us"],v["cublas"]["us"]) for m,v in d["per_width"].items()})
P
