import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cpmcu_b200 as B
dev = "cuda"
iters = 300
m = int(sys.argv[1]) if len(sys.argv) > 1 else 64
g = torch.Generator(device=dev).manual_seed(1)
names = {c: B.capi.config_name(c) for c in range(1001, 1030)}
for (k, n) in [(4096, 4608), (4096, 4096), (4096, 32768), (16384, 4096)]:
    copies = 3
    qws = [torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g) for _ in range(copies)]
    sc = (torch.rand((k // 128, n), device=dev, generator=g) * 0.018 + 0.002).half()
    a = torch.randn((m, k), device=dev, generator=g).half()
    ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=dev)
    ct = torch.empty(max(B.reduce_bytes(32, n, k), 4) // 4 + 64, dtype=torch.float32, device=dev)
    dflt = B.capi.pick_config(m, n, k)
    for cfg in [0] + list(range(1001, 1013)):
        nm = "default=" + names.get(dflt, str(dflt)) if cfg == 0 else names.get(cfg, "")
        if cfg and not nm:
            continue
        try:
            ref = B.gemm(a, qws[0], sc, ws, ct, force_config=cfg).clone()
        except Exception as e:
            continue
        out = torch.empty_like(ref)
        bad = 0
        for i in range(iters):
            B.gemm(a, qws[0], sc, ws, ct, out=out, force_config=cfg)
            if not torch.equal(out, ref):
                bad += 1
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                for i in range(12):
                    B.gemm(a, qws[i % copies], sc, ws, ct, out=out, force_config=cfg, pdl=True, stream=s)
            gr.replay(); s.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            best = 1e9
            for _ in range(5):
                e0.record(s); gr.replay(); e1.record(s); s.synchronize()
                best = min(best, e0.elapsed_time(e1) / 12 * 1e3)
        print(f"m={m} k={k} n={n} cfg={cfg} {nm}: bad {bad}/{iters}  {best:.2f} us", flush=True)
