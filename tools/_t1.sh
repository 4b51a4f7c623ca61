mkdir -p gpurun_out/i4
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/i4/pytest_gpu.txt 2>&1; tail -n 3 gpurun_out/i4/pytest_gpu.txt
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --widths 1,4,16,32 --steps 10 --no-detail --no-cpu-baseline --no-lm-head > gpurun_out/i4/$name.json 2> gpurun_out/i4/$name.err
  python - <<P
import json
d=json.load(open("gpurun_out/i4/$name.json"))
print("$name", {m:v["ms"] for m,v in d["per_width"].items()})
P
}
run A X=1
run OCC2 B200_W4A16_V1_OCC=2
run A2 X=1
python tools/chain.py --m 1 --combo "" --combo "gate_up=__half/nt1/tn4x64/kb8x16/w16" --combo "gate_up=__half/nt1/tn4x64/kb8x16/w16,down=__half/nt1/tn4x64/kb8x16/w16" 2>&1 | cut -c1-160 | tail -5
python tools/bench_dense.py --n 4096 --k 4096 --widths 1,16 2>&1 | tail -1 | cut -c1-900
