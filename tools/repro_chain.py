#!/usr/bin/env python
"""bench.py's prefill chain (operator API, auto-picked tiles, PDL) for one chunk size, eagerly, then every output
checked against the single-CTA tile (config 1002).   python tools/repro_chain.py --m 1024 [--layers 4] [--no-pdl]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as B  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--m", type=int, default=1024)
ap.add_argument("--layers", type=int, default=4)
ap.add_argument("--no-pdl", action="store_true")
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
dev = "cuda"
SHAPES = [("qkv", 4096, 4608), ("o", 4096, 4096), ("gate_up", 4096, 32768), ("down", 16384, 4096)]
g = torch.Generator(device=dev).manual_seed(4321)
B.set_pdl(not args.no_pdl)
m, layers = args.m, args.layers
stream = torch.cuda.Stream()
lins, qws, scs = [], [], []
for name, k, n in SHAPES:
    q = torch.randint(-2**31, 2**31 - 1, (layers, k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
    s = (torch.rand((layers, k // 128, n), device=dev, generator=g) * 0.018 + 0.002).half()
    row = []
    for l in range(layers):
        lin = B.W4A16GPTQMarlinLinear(k, n, 128, dtype=torch.float16, device=dev)
        lin.weight, lin.scales = q[l], s[l]
        lin.workspace = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=dev)
        row.append(lin)
    lins.append(row)
    qws.append(q)
    scs.append(s)
acts = [torch.randn((m, k), device=dev, generator=g).half() for _, k, n in SHAPES]
outs = [[torch.full((m, n), float("nan"), dtype=torch.float16, device=dev) for l in range(layers)] for _, k, n in SHAPES]
c_tmp = torch.empty(max(B.reduce_max_m(m) * n for _, k, n in SHAPES) + 64, dtype=torch.float32, device=dev)
print("tiles:", {name: B.config_name(B.pick_config(m, n, k, 128, torch.float16)) if hasattr(B, "pick_config") else "?" for name, k, n in SHAPES}, flush=True)
refs = [[None] * layers for _ in SHAPES]
for i, (name, k, n) in enumerate(SHAPES):
    for l in range(layers):
        refs[i][l] = B.gemm(acts[i], qws[i][l], scs[i][l], lins[i][l].workspace, c_tmp, force_config=1002)
torch.cuda.synchronize()
bad = 0
for rep in range(args.reps):
    for l in range(layers):
        for i in range(4):
            lins[i][l].prefill(stream, m, acts[i], None, c_tmp, tgt=outs[i][l])
    stream.synchronize()
    for i, (name, k, n) in enumerate(SHAPES):
        for l in range(layers):
            d = int(((outs[i][l] != refs[i][l]) | torch.isnan(outs[i][l])).sum())
            if d:
                bad += d
                diff = outs[i][l] != refs[i][l]
                rows = diff.sum(dim=1).nonzero().flatten()
                cols = diff.sum(dim=0).nonzero().flatten()
                print("rep", rep, name, "layer", l, ":", d, "elements differ; tokens", int(rows.min()), "..", int(rows.max()),
                      "columns", int(cols.min()), "..", int(cols.max()), flush=True)
            outs[i][l].fill_(float("nan"))
    torch.cuda.synchronize()
print("m", m, "pdl", not args.no_pdl, "reps", args.reps, "-> differing elements:", bad, flush=True)
