#!/usr/bin/env python
"""Per-CTA phase timeline of a PDL chain of streaming-kernel launches (tuning aid).
  python tools/timeline.py --shapes qkv,o,gate_up,down --m 1 --layers 4 [--cfg 0] [--no-pdl]
Prints, per launch: kernel span and the median / max over CTAs of each phase boundary relative to the
launch's first CTA start, plus the gap to the previous launch."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import cpmcu_b200 as B  # noqa: E402
from cpmcu_b200 import capi as L  # noqa: E402

SHAPES = {"qkv": (4096, 4608), "o": (4096, 4096), "gate_up": (4096, 32768), "down": (16384, 4096)}
ap = argparse.ArgumentParser()
ap.add_argument("--shapes", default="qkv,o,gate_up,down")
ap.add_argument("--m", type=int, default=1)
ap.add_argument("--layers", type=int, default=4)
ap.add_argument("--cfg", type=int, default=0)
ap.add_argument("--no-pdl", action="store_true")
ap.add_argument("--cfg-name", default="")
args = ap.parse_args()
dev = "cuda"
if args.cfg_name:
    args.cfg = next(c for c in list(range(1, 400)) if args.cfg_name in (L.config_name(c) or ""))
    print("cfg", args.cfg, L.config_name(args.cfg))
names = args.shapes.split(",")
m = args.m
ws_, lay = {}, []
for li in range(args.layers):
    for nm in names:
        k, n = SHAPES[nm]
        qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev)
        sc = (torch.rand((k // 128, n), device=dev) * 0.018 + 0.002).half()
        lay.append((nm, k, n, qw, sc))
acts = {k: torch.randn((m, k), device=dev).half() for k in (4096, 16384)}
outs = {n: torch.empty((m, n), dtype=torch.float16, device=dev) for n in (4608, 4096, 32768)}
ws = torch.zeros(B.workspace_ints(32768), dtype=torch.int32, device=dev)
ct = torch.empty(64 * 32768, dtype=torch.float32, device=dev)
nl = len(lay)


def run():
    for (nm, k, n, qw, sc) in lay:
        B.gemm(acts[k], qw, sc, ws, ct, out=outs[n], force_config=args.cfg, pdl=not args.no_pdl)


run()
torch.cuda.synchronize()
buf = torch.zeros(nl * 8192, dtype=torch.int64, device=dev)
L.lib().b200_w4a16_debug_timeline(buf.data_ptr(), nl)
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        run()
    L.lib().b200_w4a16_debug_timeline(None, 0)
    for _ in range(3):
        g.replay()
    s.synchronize()
    buf.zero_()
    s.synchronize()
    g.replay()
    s.synchronize()
ts = buf.cpu().numpy().reshape(nl, 1024, 8)
prev_end = None
PH = ["start", "priorDone", "1stFull", "loopEnd", "ticket", "tailWait", "fetched", "exit"]
print("times in us relative to the launch's first CTA start; med/max over CTAs")
print(f"{'launch':14s} {'ctas':>5s} {'gap':>6s} {'span':>6s} | " + " ".join(f"{p:>13s}" for p in PH))
for i, (nm, k, n, _, _) in enumerate(lay):
    t = ts[i]
    live = t[:, 0] > 0
    t = t[live].astype(np.float64)
    t0 = t[:, 0].min()
    end = t[:, 7].max()
    gap = (t0 - prev_end) / 1e3 if prev_end is not None else 0.0
    cols = []
    for j in range(8):
        v = (t[:, j][t[:, j] > 0] - t0) / 1e3
        cols.append(f"{np.median(v):6.2f}/{v.max():6.2f}" if len(v) else "      -      ")
    print(f"{nm + ' L' + str(i // len(names)):14s} {int(live.sum()):5d} {gap:6.2f} {(end - t0) / 1e3:6.2f} | " + " ".join(cols))
    prev_end = end
print(f"chain total: {(ts[..., 7].max() - ts[ts[..., 0] > 0][:, 0].min()) / 1e3:.2f} us for {nl} launches")
