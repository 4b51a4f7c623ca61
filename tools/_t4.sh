mkdir -p gpurun_out/k1
timeout 300 python tools/stress_m64.py 1500 > gpurun_out/k1/stress_after.txt 2>&1; cat gpurun_out/k1/stress_after.txt
B200_W4A16_TC_CORESIDENT=1 timeout 300 python tools/stress_m64.py 1500 > gpurun_out/k1/stress_coresident.txt 2>&1; cat gpurun_out/k1/stress_coresident.txt
for e in 0 1; do
B200_W4A16_TC_CORESIDENT=$e timeout 300 python bench.py --widths 64 --steps 10 --no-cpu-baseline --no-lm-head > gpurun_out/k1/bench64_$e.json 2> gpurun_out/k1/bench64_$e.err
python - <<P
import json
d=json.load(open("gpurun_out/k1/bench64_$e.json"))
print("coresident=$e", {m:v["ms"] for m,v in d["per_width"].items()}, {k:v["64"]["us"] for k,v in d["detail"].items()})
P
done
