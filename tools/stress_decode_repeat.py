#!/usr/bin/env python
"""Repeatability stress of the streaming kernels (default pick per M, so M = 32 runs the tensor-memory kernel): the
MiniCPM4-8B layer chain under PDL, every output compared bit for bit with the first pass.
  python tools/stress_decode_repeat.py --m 32 --passes 1000"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as b  # noqa: E402
from cpmcu_b200 import capi as L  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--m", type=int, default=32)
ap.add_argument("--passes", type=int, default=1000)
ap.add_argument("--bf16", action="store_true")
ap.add_argument("--gated", action="store_true", help="gate_up with the fused SiLU(gate) * up epilogue (compared with the first pass)")
args = ap.parse_args()
DT = torch.bfloat16 if args.bf16 else torch.float16
DEV = "cuda"
m = args.m
shapes = [(4096, 4608), (4096, 4096), (4096, 32768), (16384, 4096)]
print("picked:", [L.config_name(L.pick_config(m, n, k, 128, L.FP16)) for k, n in shapes], flush=True)
g = torch.Generator(device=DEV).manual_seed(5)
w = []
for k, n in shapes:
    for _ in range(2):
        qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
        sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).to(DT)
        a = torch.randn((m, k), device=DEV, generator=g).to(DT)
        ws = torch.zeros(b.workspace_ints(n), dtype=torch.int32, device=DEV)
        ct = torch.empty(max(b.reduce_bytes_fast(m, n, k) // 4, 64), dtype=torch.float32, device=DEV)
        w.append((a, qw, sc, ws, ct))
refs = [b.gemm(a, qw, sc, ws, ct) for a, qw, sc, ws, ct in w]
outs = [torch.empty_like(r) for r in refs]
torch.cuda.synchronize()
bad = 0
for rep in range(args.passes):
    for (a, qw, sc, ws, ct), o in zip(w, outs):
        b.gemm(a, qw, sc, ws, ct, out=o, pdl=True)
    torch.cuda.synchronize()
    for i, (o, r) in enumerate(zip(outs, refs)):
        if not torch.equal(o, r):
            bad += 1
            d = o != r
            print(f"pass {rep} GEMM {i}: {int(d.sum())} elements differ", flush=True)
        o.fill_(float("nan"))
print("m", m, "passes", args.passes, "-> mismatching outputs:", bad, "; workspace clean:", all(int(x[3].abs().max()) == 0 for x in w), flush=True)
