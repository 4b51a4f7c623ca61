#!/usr/bin/env python
"""Launch one (shape, M, config) a few times over rotating weight copies — the command ncu wraps.
  python tools/prof_one.py --k 4096 --n 32768 --m 1 --cfg 2010 [--copies 6] [--pdl]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as B  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--k", type=int, default=4096)
ap.add_argument("--n", type=int, default=32768)
ap.add_argument("--m", type=int, default=1)
ap.add_argument("--cfg", default="0", help="configuration id, or several separated by commas (each launched in turn)")
ap.add_argument("--copies", type=int, default=6)
ap.add_argument("--pdl", action="store_true")
ap.add_argument("--cfg-name", default="", help="pick the configuration id by (sub)name, e.g. __half/nt1/tn4x64/kb8x16/w8")
a_ = ap.parse_args()
dev = "cuda"
if a_.cfg_name:
    from cpmcu_b200 import capi as _L
    a_.cfg = "%d" % next(c for c in list(range(1, 400)) if a_.cfg_name in (_L.config_name(c) or ""))
    print("cfg", a_.cfg, _L.config_name(a_.cfg))
k, n, m = a_.k, a_.n, a_.m
qw = torch.randint(-2**31, 2**31 - 1, (a_.copies, k // 16, 2 * n), dtype=torch.int32, device=dev)
sc = (torch.rand((a_.copies, k // 128, n), device=dev) * 0.018 + 0.002).half()
a = torch.randn((m, k), device=dev).half()
out = torch.empty((m, n), dtype=torch.float16, device=dev)
ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=dev)
ct = torch.empty(max(B.reduce_bytes_fast(m, n, k) // 4, 64), dtype=torch.float32, device=dev)
for cfg in [int(c) for c in str(a_.cfg).split(",")]:
    for rep in range(2):
        for i in range(a_.copies):
            B.gemm(a, qw[i], sc[i], ws, ct, out=out, force_config=cfg, pdl=a_.pdl)
torch.cuda.synchronize()
print("done", B.launch_count())
