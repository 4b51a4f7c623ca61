import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cpmcu_b200 as B
dev = "cuda"
iters = 600
g = torch.Generator(device=dev).manual_seed(1)
m, k, n = 64, 16384, 4096
qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
sc = (torch.rand((k // 128, n), device=dev, generator=g) * 0.018 + 0.002).half()
a = torch.randn((m, k), device=dev, generator=g).half()
a2 = torch.randn((7, k), device=dev, generator=g).half()
ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=dev)
ct = torch.empty(max(B.reduce_bytes(m, n, k), B.reduce_bytes(7, n, k), 4) // 4 + 64, dtype=torch.float32, device=dev)
ref = B.gemm(a, qw, sc, ws, ct).clone()
for name, pdl_mode, inter, sync, cfgs in [("plain", 0, False, False, 0), ("sync-each", 0, False, True, 0), ("pdl", 1, False, False, 0),
                                    ("interleave", 0, True, False, 0), ("pdl+interleave", 1, True, False, 0)]:
    outs = [torch.empty_like(ref) for _ in range(8)]
    bad = 0
    rows = set()
    for i in range(iters):
        if inter and i % 3 == 0:
            B.gemm(a2, qw, sc, ws, ct)
        out = outs[i % 8]
        B.gemm(a, qw, sc, ws, ct, out=out, pdl=bool(pdl_mode))
        if sync:
            torch.cuda.synchronize()
        if not torch.equal(out, ref):
            bad += 1
            d = (out.float() - ref.float()).abs()
            idx = torch.nonzero(d > 0)
            rows.add((int(idx[:, 0].min()), int(idx[:, 0].max()), int(idx[:, 1].min()), int(idx[:, 1].max()), int(idx.shape[0])))
    torch.cuda.synchronize()
    print(f"{name}: {bad}/{iters} bad; (tok_min, tok_max, col_min, col_max, count) samples: {sorted(rows)[:6]}", flush=True)
for S in (1, 2, 4, 8):
    # forced cluster sizes of the split-K bm64 tile
    cfg = None
    for c in range(1001, 1040):
        nm = B.capi.config_name(c)
        if nm == "tc/bn128/bm64/splitk":
            cfg = c
            break
    import math
    c = cfg + int(math.log2(S))
    out = torch.empty_like(ref)
    bad = 0
    try:
        r0 = B.gemm(a, qw, sc, ws, ct, force_config=c).clone()
        for i in range(iters):
            B.gemm(a, qw, sc, ws, ct, out=out, force_config=c)
            if not torch.equal(out, r0):
                bad += 1
        print(f"S={S} cfg={c}: {bad}/{iters} bad, max diff vs default ref {float((r0.float()-ref.float()).abs().max()):.3g}", flush=True)
    except Exception as e:
        print(f"S={S}: {e}")
