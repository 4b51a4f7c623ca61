#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python tools/stress_prefill_chain.py --m 354 --passes 30000 --force 1013
timeout 900 python tools/stress_prefill_chain.py --m 1024 --passes 12000
timeout 900 python tools/stress_prefill_chain.py --m 300 --passes 10000 --force 1013
} > gpurun_out/stress64.txt 2>&1
cut -c1-300 gpurun_out/stress64.txt | tail -12
python tools/chain.py --m 1024 --layers 4 --reps 5 --combo "" | awk '{ printf "%s %s us\n", $1,$2 }'
python tools/chain.py --m 4096 --layers 4 --reps 5 --combo "" | awk '{ printf "%s %s us\n", $1,$2 }'
timeout 600 python bench.py --no-tp1 --no-cpu-baseline --no-lm-head --widths 1 --layers 2 --steps 3 --warmup 3 > gpurun_out/bench64_prefill.json 2>/dev/null; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench64_prefill.json').read().strip().splitlines()[-1])
print({m:(v['us_per_layer'],v['tflops']) for m,v in d['prefill']['per_m'].items()})
PY
