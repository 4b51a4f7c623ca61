mkdir -p gpurun_out/i1
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --widths 1,4,16,32 --steps 10 --no-detail --no-cpu-baseline --no-lm-head > gpurun_out/i1/$name.json 2> gpurun_out/i1/$name.err
  python - <<P
import json
d=json.load(open("gpurun_out/i1/$name.json"))
print("$name", {m:v["ms"] for m,v in d["per_width"].items()})
P
}
run A X=1
run B B200_W4A16_MIX_BYTES_M8=10000000 B200_W4A16_MIX_BYTES_M16=10000000 B200_W4A16_MIX_BYTES_M32=10000000
run C B200_W4A16_MIX_BYTES_M8=10000000 B200_W4A16_MIX_BYTES_M16=40000000 B200_W4A16_MIX_BYTES_M32=40000000
run D B200_W4A16_MIX_BYTES_M8=40000000 B200_W4A16_MIX_BYTES_M16=100000000 B200_W4A16_MIX_BYTES_M32=100000000
run A2 X=1
