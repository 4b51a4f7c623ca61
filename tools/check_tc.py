#!/usr/bin/env python
"""The tensor-memory streaming kernel (geometry 7, '/tmem') against the default streaming kernel and fp32 torch:
  python tools/check_tc.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as B  # noqa: E402
from cpmcu_b200 import capi as L  # noqa: E402

dev = "cuda"
g = torch.Generator(device=dev).manual_seed(5)


TILE = os.environ.get("TC_TILE", "tn4x64/kb8x16")


def cfg_id(dtype, nt, grouped=True):
    want = f"{'__half' if dtype == torch.float16 else '__nv_bfloat16'}/nt{nt}/{TILE}/w8/s5/{'true' if grouped else 'false'}/tmem"
    for c in range(1, 200):
        if (L.config_name(c) or "") == want:
            return c
    raise RuntimeError(want)


bad = 0
for dtype in (torch.float16, torch.bfloat16):
    for (m, k, n) in [(1, 1024, 512), (4, 4096, 4608), (16, 4096, 4096), (16, 4096, 32768), (9, 16384, 4096), (32, 4096, 4608),
                      (20, 2048, 768), (1, 4096, 32768)]:
        qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
        sc = (torch.rand((k // 128, n), device=dev, generator=g) * 0.018 + 0.002).to(dtype)
        a = torch.randn((m, k), device=dev, generator=g).to(dtype)
        ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=dev)
        ct = torch.empty(max(B.reduce_bytes_fast(m, n, k) // 4, 64), dtype=torch.float32, device=dev)
        ref = B.gemm(a, qw, sc, ws, ct)
        nt = 1 if m <= 8 else (2 if m <= 16 else 4)
        cid = cfg_id(dtype, nt)
        out = torch.full((m, n), float("nan"), dtype=dtype, device=dev)
        B.gemm(a, qw, sc, ws, ct, out=out, force_config=cid)
        torch.cuda.synchronize()
        out2 = torch.full((m, n), float("nan"), dtype=dtype, device=dev)
        B.gemm(a, qw, sc, ws, ct, out=out2, force_config=cid, pdl=True)
        torch.cuda.synchronize()
        exact = a.float() @ B.dequant(qw, sc, k, n).float()
        err = float((out.float() - exact).abs().max() / exact.abs().mean())
        err_ref = float((ref.float() - exact).abs().max() / exact.abs().mean())
        same = bool(torch.equal(out, out2))
        ok = err <= err_ref * 1.05 + 1e-6 and same and int(ws.abs().max()) == 0 and bool(torch.isfinite(out.float()).all())
        bad += 0 if ok else 1
        print(f"{str(dtype)[6:]:9s} m={m:3d} k={k:5d} n={n:5d} cfg {cid:3d}: max rel err {err:.2e} (default kernel {err_ref:.2e}) repeatable {same} "
              f"{'ok' if ok else 'FAIL'}", flush=True)
print("failures:", bad)
