#!/bin/bash
mkdir -p gpurun_out
for v in 17 9; do
  B200_W4A16_TMEM_MIN_M=$v timeout 600 python bench.py --no-cpu-baseline --no-lm-head --no-prefill --widths 16 --tp-widths 16 --steps 10 --warmup 3 > gpurun_out/bench66_$v.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench66_$v.json').read().strip().splitlines()[-1])
print("TMEM_MIN_M=$v", {w:v['ms'] for w,v in d['per_width'].items()}, {w:v['ms'] for w,v in d.get('llama70b_tp1',{}).get('per_width',{}).items()})
PY
done
for m in 16 24 32; do B200_W4A16_TMEM_MIN_M=9 python tools/chain.py --m $m --layers 8 --reps 5 --combo "" | awk '{ printf "tmem>=9 %s %s us\n", $1,$2 }'; B200_W4A16_TMEM_MIN_M=99 python tools/chain.py --m $m --layers 8 --reps 5 --combo "" | awk '{ printf "mma.sync %s %s us\n", $1,$2 }'; done
