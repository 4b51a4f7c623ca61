mkdir -p gpurun_out/i5
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/i5/pytest_gpu.txt 2>&1; tail -n 2 gpurun_out/i5/pytest_gpu.txt
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --widths 1,4,16,32 --steps 10 --no-detail --no-cpu-baseline --no-lm-head > gpurun_out/i5/$name.json 2> gpurun_out/i5/$name.err
  python - <<P
import json
d=json.load(open("gpurun_out/i5/$name.json"))
print("$name", {m:v["ms"] for m,v in d["per_width"].items()})
P
}
run A X=1
run A2 X=1
