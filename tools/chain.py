#!/usr/bin/env python
"""Time the bench chain (qkv -> o -> gate_up -> down) x layers in one CUDA graph with per-shape forced
configurations (tuning aid).  Each --combo is "qkv=<cfg>,o=<cfg>,gate_up=<cfg>,down=<cfg>" where <cfg>
is a configuration id or a (sub)name; missing shapes use the library default (0).
  python tools/chain.py --m 1 --layers 32 --combo "gate_up=__half/nt1/tn4x64/kb8x16/w8" --combo ""
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as B  # noqa: E402
from cpmcu_b200 import capi as L  # noqa: E402

SHAPES = [("qkv", 4096, 4608), ("o", 4096, 4096), ("gate_up", 4096, 32768), ("down", 16384, 4096)]
ap = argparse.ArgumentParser()
ap.add_argument("--m", type=int, default=1)
ap.add_argument("--layers", type=int, default=32)
ap.add_argument("--combo", action="append", default=[])
ap.add_argument("--no-pdl", action="store_true")
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--only", default="", help="chain only this shape (qkv|o|gate_up|down)")
args = ap.parse_args()
dev = "cuda"
m = args.m
ALL = list(range(1, 400)) + list(range(1001, 1040))


def resolve(x):
    if not x or x == "0":
        return 0
    if x.isdigit():
        return int(x)
    return next(c for c in ALL if x in (L.config_name(c) or ""))


lay = []
for li in range(args.layers):
    for nm, k, n in SHAPES:
        if args.only and nm != args.only:
            continue
        qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev)
        sc = (torch.rand((k // 128, n), device=dev) * 0.018 + 0.002).half()
        lay.append((nm, k, n, qw, sc))
acts = {k: torch.randn((m, k), device=dev).half() for k in (4096, 16384)}
outs = {n: torch.empty((m, n), dtype=torch.float16, device=dev) for n in (4608, 4096, 32768)}
ws = torch.zeros(B.workspace_ints(32768), dtype=torch.int32, device=dev)
ct = torch.empty(max(B.reduce_bytes_fast(m, n, k) for _, k, n in SHAPES) // 4 + 64, dtype=torch.float32, device=dev)
total_bytes = sum(k * n // 2 + (k // 128) * n * 2 for _, k, n, _, _ in lay)

for combo in (args.combo or [""]):
    cfg = {nm: 0 for nm, _, _ in SHAPES}
    for part in combo.split(","):
        if "=" in part:
            a, b = part.split("=", 1)
            cfg[a.strip()] = resolve(b.strip())

    def run():
        for (nm, k, n, qw, sc) in lay:
            B.gemm(acts[k], qw, sc, ws, ct, out=outs[n], force_config=cfg[nm], pdl=not args.no_pdl)

    try:
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            run()
            s.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                run()
            for _ in range(3):
                g.replay()
            s.synchronize()
            best = 1e9
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for _ in range(args.reps):
                e0.record(s)
                g.replay()
                e1.record(s)
                s.synchronize()
                best = min(best, e0.elapsed_time(e1))
        assert int(ws.abs().max()) == 0, "workspace not left zeroed"
        names = {nm: (L.config_name(c) if c else "default:" + L.config_name(L.pick_config(m, n, k))) for (nm, k, n), c in zip(SHAPES, cfg.values())}
        print(f"m={m} {best * 1e3 / args.layers:7.2f} us/layer  {total_bytes / best / 1e6:7.1f} GB/s  {combo or 'default'}   {names}", flush=True)
    except Exception as e:  # noqa: BLE001
        print(f"m={m} FAILED {combo}: {e}", flush=True)
