"""Debug: tcgen05 split-K at M=64 across K / cluster sizes; prints where the wrong outputs sit."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cpmcu_b200 as B
from cpmcu_b200 import capi as L
DEV = "cuda"
for (m, k, n) in [(64, 4096, 4096), (64, 8192, 4096), (64, 16384, 4096), (64, 16384, 1024), (128, 16384, 4096)]:
    g = torch.Generator(device=DEV).manual_seed(m + n)
    qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
    sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).half()
    a = torch.randn((m, k), device=DEV, generator=g).half()
    ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=DEV)
    ct = torch.empty(max(B.reduce_bytes(m, n, k) // 4, 64), dtype=torch.float32, device=DEV)
    w = B.dequant(qw, sc, k, n)
    exact = a.double() @ w.double()
    print(f"== m={m} k={k} n={n} default cfg {L.pick_config(m, n, k)} {L.config_name(L.pick_config(m, n, k))}")
    for cfg in [0] + list(range(1005, 1013)):
        name = L.config_name(cfg) if cfg else "default"
        if cfg and (("bm64" in name and m > 64)):
            continue
        for rep in range(3):
            out = B.gemm(a, qw, sc, ws, ct, force_config=cfg)
            torch.cuda.synchronize()
            err = (out.double() - exact).abs()
            bad = err > 0.05 + 0.01 * exact.abs()
            nb = int(bad.sum())
            if nb:
                rows = bad.any(dim=1).nonzero().flatten().tolist()
                cols = bad.any(dim=0).nonzero().flatten()
                print(f"  cfg {cfg} {name} rep {rep}: {nb} bad, max {float(err.max()):.3f}; rows {rows[:8]}..({len(rows)}) "
                      f"cols {cols[:6].tolist()}..{cols[-3:].tolist()} ({len(cols)}) coltiles {sorted(set((cols // 128).tolist()))[:20]}")
            elif rep == 0:
                print(f"  cfg {cfg} {name}: ok (max err {float(err.max()):.4f})")
