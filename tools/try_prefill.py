#!/usr/bin/env python
"""Quick GPU check of the tcgen05 kernel: every tile shape, a few (m, k, n), against a torch fp32
matmul with the library's own bit-exact dequant, plus timing at BASELINE's prefill shapes.
Run under `timeout` on the GPU box:  timeout 300 python tools/try_prefill.py [--time]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cpmcu_b200 as B  # noqa: E402
from cpmcu_b200 import capi  # noqa: E402

DEV = "cuda"


def problem(m, k, n, dtype, seed=0):
    g = torch.Generator(device=DEV).manual_seed(seed)
    qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
    sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).to(dtype)
    a = torch.randn((m, k), device=DEV, generator=g).to(dtype)
    return a, qw, sc


def check(m, k, n, dtype, cfg):
    a, qw, sc = problem(m, k, n, dtype, seed=m + n)
    ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=DEV)
    ct = torch.empty(max(B.reduce_bytes_fast(min(m, 64), n, k) // 4, 1), dtype=torch.float32, device=DEV)
    out = B.gemm(a, qw, sc, ws, ct, force_config=cfg)
    torch.cuda.synchronize()
    w = B.dequant(qw, sc, k, n)
    ref = a.double() @ w.double()
    err = (out.double() - ref).abs()
    denom = ref.abs().clamp_min(ref.abs().max().item() * 1e-3)
    rel = (err / denom).max().item()
    print(f"  m={m:5d} k={k:5d} n={n:5d} {str(dtype)[6:]:8s} cfg={cfg} {capi.config_name(cfg) if cfg else 'auto':16s} "
          f"max_rel={rel:.3e} mean_abs={err.mean().item():.3e}", flush=True)
    return rel


def timeit(m, k, n, cfg, iters=20):
    a, qw, sc = problem(m, k, n, torch.float16)
    ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=DEV)
    ct = torch.empty(1024, dtype=torch.float32, device=DEV)
    out = torch.empty((m, n), dtype=torch.float16, device=DEV)
    for _ in range(3):
        B.gemm(a, qw, sc, ws, ct, out=out, force_config=cfg)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        B.gemm(a, qw, sc, ws, ct, out=out, force_config=cfg)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    tf = 2.0 * m * k * n / ms / 1e9
    print(f"  m={m:5d} k={k:5d} n={n:5d} cfg={cfg} {capi.config_name(cfg) if cfg else capi.config_name(capi.pick_config(m, n, k)):16s} "
          f"{ms * 1e3:9.1f} us  {tf:7.1f} TFLOP/s", flush=True)


def main():
    torch.cuda.set_device(0)
    worst = 0.0
    print("correctness:")
    for cfg in (1001, 1002, 1003, 1004):
        for (m, k, n) in [(128, 128, 128), (256, 256, 256), (100, 256, 256), (354, 512, 320), (513, 1024, 1152)]:
            worst = max(worst, check(m, k, n, torch.float16, cfg))
        worst = max(worst, check(300, 384, 448, torch.bfloat16, cfg))
    for cfg in range(1005, 1013):
        for (m, k, n) in [(16, 256, 128), (40, 1024, 320), (64, 2048, 1152), (100, 512, 256), (128, 4096, 4608)]:
            if m > 64 and cfg < 1009:
                continue
            worst = max(worst, check(m, k, n, torch.float16, cfg))
        worst = max(worst, check(61, 1024, 448, torch.bfloat16, cfg))
    worst = max(worst, check(1024, 4096, 4608, torch.float16, 0))
    print("worst max_rel:", worst)
    if worst > 1e-2:
        print("FAIL")
        sys.exit(1)
    if "--time" in sys.argv:
        print("timing (hot L2, back-to-back launches):")
        for m in (1024, 4096):
            for (k, n) in [(4096, 4608), (4096, 4096), (4096, 32768), (16384, 4096)]:
                for cfg in (0, 1001, 1002, 1003, 1004):
                    timeit(m, k, n, cfg)
    print("OK")


if __name__ == "__main__":
    main()
