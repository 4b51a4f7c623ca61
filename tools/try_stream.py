#!/usr/bin/env python
"""Quick GPU check of the cluster split-K streaming kernel: every geometry x cluster size on a few
shapes against fp64 matmul with the library's bit-exact dequant.  Run under `timeout`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as B  # noqa: E402
from cpmcu_b200 import capi  # noqa: E402

DEV = "cuda"
worst = 0.0
for (m, k, n) in [(1, 512, 256), (5, 1024, 320), (16, 2048, 1088), (27, 4096, 4608), (64, 1024, 512)]:
    for dtype in (torch.float16, torch.bfloat16):
        g = torch.Generator(device=DEV).manual_seed(m + n)
        qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
        sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).to(dtype)
        a = torch.randn((m, k), device=DEV, generator=g).to(dtype)
        ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=DEV)
        ct = torch.empty(max(B.reduce_bytes(m, n, k) // 4, 1), dtype=torch.float32, device=DEV)
        ref = a.double() @ B.dequant(qw, sc, k, n).double()
        first = None
        for cfg in range(2001, 2001 + 80):
            out = B.gemm(a, qw, sc, ws, ct, force_config=cfg)
            out2 = B.gemm(a, qw, sc, ws, ct, force_config=cfg, pdl=True)
            torch.cuda.synchronize()
            assert torch.equal(out, out2), (cfg, "not deterministic")
            err = (out.double() - ref).abs()
            rel = (err / ref.abs().clamp_min(ref.abs().mean().item() * 1e-2)).max().item()
            worst = max(worst, rel)
            tol = 1e-2 if dtype == torch.float16 else 2e-2
            flag = "" if rel <= tol else "  <-- FAIL"
            print(f"m={m:3d} k={k:5d} n={n:5d} {str(dtype)[6:]:8s} cfg={cfg} {capi.config_name(cfg)}/S{1 + ((cfg - 2001) & 7)} "
                  f"max_rel={rel:.2e}{flag}", flush=True)
print("worst", worst)
print("OK" if worst <= 2e-2 else "FAIL")
