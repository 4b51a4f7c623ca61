mkdir -p gpurun_out/k2
for mode in 1 2; do
  echo "== pad mode $mode"
  B200_W4A16_TC_PAD=$mode PDL_ALL=1 timeout 300 python tools/stress_m64.py 3000 > gpurun_out/k2/stress_pad$mode.txt 2>&1; cat gpurun_out/k2/stress_pad$mode.txt
  B200_W4A16_TC_PAD=$mode timeout 300 python bench.py --widths 64 --steps 10 --no-cpu-baseline --no-lm-head > gpurun_out/k2/bench64_pad$mode.json 2> gpurun_out/k2/bench64_pad$mode.err
  python - <<P
import json
d=json.load(open("gpurun_out/k2/bench64_pad$mode.json"))
print("pad=$mode", {m:v["ms"] for m,v in d["per_width"].items()}, {k:v["64"]["us"] for k,v in d["detail"].items()})
P
done
