#!/usr/bin/env python
"""Stress the chunked-prefill chain as tests/test_gpu_parity.py::test_tcgen05_pair_tile_repeatable_under_pdl_chains
runs it (b.gemm, auto-picked tiles, PDL, default stream) and report WHERE an output differs from the single-CTA tile:
  python tools/stress_prefill_chain.py --m 354 --passes 200 [--no-pdl] [--force CFG]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as b  # noqa: E402
from cpmcu_b200 import capi as L  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--m", type=int, default=354)
ap.add_argument("--passes", type=int, default=200)
ap.add_argument("--layers", type=int, default=2)
ap.add_argument("--no-pdl", action="store_true")
ap.add_argument("--force", type=int, default=0, help="force this configuration for every GEMM instead of the picked one")
ap.add_argument("--ref", type=int, default=1002)
ap.add_argument("--bf16", action="store_true")
ap.add_argument("--gated", action="store_true", help="gate_up with the fused SiLU(gate) * up epilogue (compared with the first pass)")
args = ap.parse_args()
DT = torch.bfloat16 if args.bf16 else torch.float16
DEV = "cuda"
m = args.m
shapes = [(4096, 4608), (4096, 4096), (4096, 32768), (16384, 4096)]
picked = [L.config_name(L.pick_config(m, n, k, 128, L.BF16 if args.bf16 else L.FP16)) for k, n in shapes]
print("picked:", picked, flush=True)
g = torch.Generator(device=DEV).manual_seed(77)
w = []
for k, n in shapes:
    for _ in range(args.layers):
        qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
        sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).to(DT)
        a = torch.randn((m, k), device=DEV, generator=g).to(DT)
        w.append((a, qw, sc, torch.zeros(b.workspace_ints(n), dtype=torch.int32, device=DEV)))
c_tmp = torch.empty(b.reduce_max_m(m) * 32768, dtype=torch.float32, device=DEV)
gated = [args.gated and qw.shape[1] // 2 == 32768 for a, qw, sc, ws in w]
refs = [b.gemm(a, qw, sc, ws, c_tmp, force_config=0 if gt else args.ref, gated_silu=gt) for (a, qw, sc, ws), gt in zip(w, gated)]
outs = [torch.empty_like(r) for r in refs]
torch.cuda.synchronize()
nbad = 0
for rep in range(args.passes):
    for (a, qw, sc, ws), o, gt in zip(w, outs, gated):
        b.gemm(a, qw, sc, ws, c_tmp, out=o, pdl=not args.no_pdl, force_config=0 if gt else args.force, gated_silu=gt)
    torch.cuda.synchronize()
    for i, (o, r) in enumerate(zip(outs, refs)):
        if not torch.equal(o, r):
            nbad += 1
            diff = (o != r) | torch.isnan(o)
            rows = diff.sum(dim=1).nonzero().flatten()
            cols = diff.sum(dim=0).nonzero().flatten()
            rr, cc = diff.nonzero()[0].tolist()
            print(f"pass {rep} GEMM {i} ({picked[i // args.layers]}): {int(diff.sum())} elements differ; tokens {int(rows.min())}..{int(rows.max())} "
                  f"({rows.numel()}), columns {int(cols.min())}..{int(cols.max())} ({cols.numel()}); first ({rr},{cc}) got {float(o[rr, cc]):.4f} "
                  f"want {float(r[rr, cc]):.4f}; nan {int(torch.isnan(o).sum())}", flush=True)
        o.fill_(float("nan"))
print("m", m, "passes", args.passes, "pdl", not args.no_pdl, "-> mismatching outputs:", nbad, flush=True)
