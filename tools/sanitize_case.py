#!/usr/bin/env python
"""One small launch per kernel family, for compute-sanitizer (racecheck / synccheck / memcheck):
  compute-sanitizer --tool racecheck python tools/sanitize_case.py
Families: streaming kernel with split columns (M = 4, 20, 32), whole-column plan (contract-sized scratch), gated SiLU,
tcgen05 split-K (M = 64), whole-K (M = 300) and the CTA-pair tile, the un-quantised streaming linear, dequant dump, repack."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as B  # noqa: E402

dev = "cuda"
g = torch.Generator(device=dev).manual_seed(1)


def case(m, k, n, contract=False, **kw):
    qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
    sc = (torch.rand((k // 128, n), device=dev, generator=g) * 0.018 + 0.002).half()
    a = torch.randn((m, k), device=dev, generator=g).half()
    ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=dev)
    nbytes = B.reduce_max_m(m) * n * 4 if contract else max(B.reduce_bytes_fast(m, n, k), 64)
    ct = torch.empty(nbytes // 4, dtype=torch.float32, device=dev)
    out = B.gemm(a, qw, sc, ws, ct, **kw)
    torch.cuda.synchronize()
    if not kw.get("gated_silu"):
        ref = a.float() @ B.dequant(qw, sc, k, n).float()
        err = float((out.float() - ref).abs().max() / ref.abs().mean())
        assert err < 2e-2, (m, k, n, err)
    assert int(ws.abs().max()) == 0
    print("ok", m, k, n, contract, kw, flush=True)


case(4, 1024, 512)
case(20, 2048, 512)
case(32, 1024, 768)
case(16, 1024, 512, contract=True)
case(8, 512, 1024, gated_silu=True)
case(64, 1024, 256)
case(300, 512, 384)
case(300, 512, 512, force_config=1013)   # cta_group::2 pair tile
w = torch.randn((1024, 512), device=dev, generator=g).half()
x = torch.randn((4, 512), device=dev, generator=g).half()
B.dense_linear(x, w)
torch.cuda.synchronize()
print("sanitize cases done", flush=True)
