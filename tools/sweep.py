"""Tuning sweep (run on the B200 box): every decode configuration x MiniCPM4-8B layer shape x M,
weights rotated over > 2x L2 so every launch streams from HBM; CUDA-graph replays timed with events.
Also times the reference Marlin (oracle/_ref) the same way.  Prints one line per measurement and
writes gpurun_out/sweep.json."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cpmcu_b200 as B  # noqa: E402
from cpmcu_b200 import capi as L  # noqa: E402
from oracle import ref_marlin  # noqa: E402

SHAPES = {"qkv": (4096, 4608), "o": (4096, 4096), "gate_up": (4096, 32768), "down": (16384, 4096)}
HBM = 6541.8


def alg_bytes(k, n):
    return k * n // 2 + (k // 128) * n * 2


def time_graph(fn_launch_all, reps=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn_launch_all()  # warm-up (module load, attribute set, descriptor cache)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fn_launch_all()
        g.replay()
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(reps):
            e0.record(s)
            g.replay()
            e1.record(s)
            s.synchronize()
            best = min(best, e0.elapsed_time(e1))
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ms", default="1,4,16")
    ap.add_argument("--shapes", default="qkv,o,gate_up,down")
    ap.add_argument("--geoms", default="all")
    ap.add_argument("--family", default="both", choices=["v1", "v2", "both", "tc", "all"])
    ap.add_argument("--pdl-only", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.json"))
    ap.add_argument("--ss", default="", help="cluster sizes to try, e.g. 1,2,3,4,6,8 (default all)")
    args = ap.parse_args()
    args.ss = [int(x) for x in args.ss.split(",")] if args.ss else []
    dev = "cuda"
    results = []
    for name in args.shapes.split(","):
        k, n = SHAPES[name]
        nbytes = alg_bytes(k, n)
        copies = max(4, int(300e6 // nbytes) + 1)
        qw = torch.randint(-2**31, 2**31 - 1, (copies, k // 16, 2 * n), dtype=torch.int32, device=dev)
        sc = (torch.rand((copies, k // 128, n), device=dev) * 0.018 + 0.002).half()
        for m in [int(x) for x in args.ms.split(",")]:
            a = torch.randn((m, k), device=dev).half()
            out = torch.empty((m, n), dtype=torch.float16, device=dev)
            ws = torch.zeros(B.workspace_ints(n), dtype=torch.int32, device=dev)
            ct = torch.empty(max(B.reduce_bytes(m, n, k) // 4, ref_marlin.reduce_max_m(m) * n if ref_marlin.available() else 0, 64 * n),
                             dtype=torch.float32, device=dev)
            # reference Marlin
            ref_out = None
            if ref_marlin.available():
                def run_ref():
                    for i in range(copies):
                        ref_marlin.gemm(a, qw[i], sc[i], ws, ct, out=out)
                t = time_graph(run_ref) / copies * 1e3
                ref_out = ref_marlin.gemm(a, qw[0], sc[0], ws, ct).clone()
                r = dict(shape=name, m=m, impl="reference_marlin", us=t, gbs=nbytes / t / 1e3, frac=nbytes / t / 1e3 / HBM)
                results.append(r)
                print(json.dumps(r), flush=True)
            default_cfg = L.pick_config(m, n, k)
            cands = []
            nt = 1 if m <= 8 else (2 if m <= 16 else 4)
            if args.family in ("v1", "both") and m <= 64:
                for cfg in range(1, 400):
                    cname = L.config_name(cfg)
                    if not cname:
                        break
                    if cname.startswith(f"__half/nt{nt}/") and cname.endswith("true"):
                        cands.append((cfg, cname))
            if args.family in ("v2", "both", "all") and m <= 64:
                for cfg in range(2001, 2001 + 128):
                    cname = L.config_name(cfg)
                    S = 1 + ((cfg - 2001) & 7)
                    if cname and (not args.ss or S in args.ss) and (args.geoms == "all" or str((cfg - 2001) >> 3) in args.geoms.split(",")):
                        cands.append((cfg, f"{cname}/S{S}"))
            if args.family in ("tc", "all"):
                for cfg in range(1001, 1013):
                    cname = L.config_name(cfg)
                    if not cname:
                        continue
                    if "splitk" in cname:
                        if ("bm64" in cname and m > 64) or m > 128:
                            continue
                        cname = f"{cname}/S{1 << ((cfg - 1005) & 3)}"
                    cands.append((cfg, cname))
            for cfg, cname in cands:
                for pdl in ((True,) if args.pdl_only else (False, True)):
                    def run():
                        for i in range(copies):
                            B.gemm(a, qw[i], sc[i], ws, ct, out=out, force_config=cfg, pdl=pdl)
                    try:
                        t = time_graph(run) / copies * 1e3
                    except Exception as e:  # noqa: BLE001
                        print(f"# {name} m={m} {cname} pdl={pdl}: {e}", flush=True)
                        continue
                    err = None
                    if ref_out is not None:
                        mine = B.gemm(a, qw[0], sc[0], ws, ct, force_config=cfg)
                        err = float(((mine.float() - ref_out.float()).abs().max() /
                                     ref_out.float().abs().mean()).item())
                    r = dict(shape=name, m=m, impl=cname, cfg=cfg, pdl=pdl, default=(cfg == default_cfg), us=t,
                             gbs=nbytes / t / 1e3, frac=nbytes / t / 1e3 / HBM, err_vs_ref=err)
                    results.append(r)
                    print(json.dumps(r), flush=True)
        del qw, sc
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
