"""Stress the tcgen05 split-K path (32 < M <= 256): repeated launches must be bit-identical (fixed reduction
order); prints the number of mismatching launches per shape.  Other GEMMs run in between so that shared-memory /
TMEM state left by different kernels varies."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cpmcu_b200 as B

dev = "cuda"
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 300
g = torch.Generator(device=dev).manual_seed(1)
SHAPES = [(64, 16384, 4096), (64, 4096, 4096), (64, 4096, 32768), (48, 16384, 4096), (128, 16384, 4096), (256, 4096, 4608)]
if len(sys.argv) > 2 and sys.argv[2] == "decode":   # the streaming regime (persistent stream-K kernel, PDL co-residency)
    SHAPES = [(1, 4096, 4608), (1, 4096, 32768), (1, 16384, 4096), (8, 16384, 4096), (16, 4096, 32768), (16, 16384, 4096),
              (32, 4096, 4096), (32, 4096, 32768)]
for (m, k, n) in SHAPES:
    qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
    sc = (torch.rand((k // 128, n), device=dev, generator=g) * 0.018 + 0.002).half()
    a = torch.randn((m, k), device=dev, generator=g).half()
    a2 = torch.randn((7, k), device=dev, generator=g).half()
    ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=dev)
    ct = torch.empty(max(B.reduce_bytes_fast(m, n, k), B.reduce_bytes_fast(7, n, k), 4) // 4 + 64, dtype=torch.float32, device=dev)
    ref = B.gemm(a, qw, sc, ws, ct).clone()
    out = torch.empty_like(ref)
    bad = 0
    worst = 0.0
    for i in range(iters):
        if i % 3 == 0:
            B.gemm(a2, qw, sc, ws, ct)          # a streaming-kernel launch in between
        B.gemm(a, qw, sc, ws, ct, out=out, pdl=(os.environ.get("PDL_ALL") == "1" or i % 2 == 1))
        if not torch.equal(out, ref):
            bad += 1
            worst = max(worst, float((out.float() - ref.float()).abs().max()))
    torch.cuda.synchronize()
    print(f"m={m} k={k} n={n} cfg={B.capi.config_name(B.capi.pick_config(m, n, k))}: {bad}/{iters} mismatching launches, worst abs diff {worst:.3g}", flush=True)
