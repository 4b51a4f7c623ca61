#!/usr/bin/env python
"""Back-to-back launches of one tcgen05 configuration under PDL, compared with the single-CTA tile (config 1002):
which output tiles differ, and in which half of the 256 columns.
  python tools/repro_prefill.py [--cfg 1013] [--k 16384 --n 4096 --m 1024] [--launches 3] [--no-pdl]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as B  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cfg", type=int, default=1013)
ap.add_argument("--k", type=int, default=16384)
ap.add_argument("--n", type=int, default=4096)
ap.add_argument("--m", type=int, default=1024)
ap.add_argument("--launches", type=int, default=3)
ap.add_argument("--no-pdl", action="store_true")
ap.add_argument("--trials", type=int, default=5)
args = ap.parse_args()
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(4321)
k, n, m = args.k, args.n, args.m
q = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
s = (torch.rand((k // 128, n), device=dev, generator=g) * 0.018 + 0.002).half()
a = torch.randn((m, k), device=dev, generator=g).half()
ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=dev)
ct = torch.empty(B.reduce_max_m(m) * n, dtype=torch.float32, device=dev)
ref = B.gemm(a, q, s, ws, ct, force_config=1002)
torch.cuda.synchronize()
bad_total = 0
for t in range(args.trials):
    outs = [torch.full((m, n), float("nan"), dtype=torch.float16, device=dev) for _ in range(args.launches)]
    torch.cuda.synchronize()
    for o in outs:
        B.gemm(a, q, s, ws, ct, out=o, force_config=args.cfg, pdl=not args.no_pdl)
    torch.cuda.synchronize()
    for i, o in enumerate(outs):
        diff = (o != ref) | torch.isnan(o)
        nbad = int(diff.sum())
        if nbad == 0:
            continue
        bad_total += nbad
        tiles = diff.view(m // 256 if m % 256 == 0 else 1, -1, n // 256, 256) if m % 256 == 0 else None
        msg = f"trial {t} launch {i}: {nbad} elements differ"
        if tiles is not None:
            per = tiles.sum(dim=(1, 3))   # [mt][nt256]
            idx = per.nonzero().tolist()
            halves = diff.view(m, n // 128, 128).sum(dim=(0, 2)).nonzero().flatten().tolist()
            toks = diff.sum(dim=1).nonzero().flatten()
            msg += f"; (mt, nt) tiles {idx[:12]} counts {[int(per[i0, i1]) for i0, i1 in idx[:12]]}; 128-column blocks {halves[:16]}; " \
                   f"tokens {int(toks.min())}..{int(toks.max())} ({toks.numel()} rows); nan {int(torch.isnan(o).sum())}"
            r, c = diff.nonzero()[0].tolist()
            msg += f"; first ({r},{c}): got {float(o[r, c]):.4f} want {float(ref[r, c]):.4f}"
        print(msg, flush=True)
print("cfg", args.cfg, "pdl", not args.no_pdl, "k n m", k, n, m, "launches", args.launches, "-> differing elements:", bad_total, flush=True)
