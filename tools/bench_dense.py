"""Un-quantised lm_head (73448 x 4096 fp16) at tree-verify widths: the streaming
kernel of this library (b200_w4a16_dense_linear_forward) against cuBLAS (torch.nn.functional.linear, what the
reference's linear<T>() calls).  CUDA events around graph replays; the 602 MB weight matrix is larger than L2
(126 MB), so every launch streams from HBM.  Prints one JSON line."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=73448)
    ap.add_argument("--k", type=int, default=4096)
    ap.add_argument("--widths", default="1,4,16,64")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--once", action="store_true", help="one plain launch per width (for ncu)")
    args = ap.parse_args()
    import torch
    import cpmcu_b200 as B
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    try:
        peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
        peak_src = "MEASURED_PEAKS.json"
    except Exception:
        peak, peak_src = 6500.0, "fallback"
    n, k = args.n, args.k
    g = torch.Generator(device=dev).manual_seed(0)
    w = (torch.randn((n, k), generator=g, device=dev, dtype=torch.float32) * 0.02).half()
    bytes_w = n * k * 2
    res = {}
    for m in [int(x) for x in args.widths.split(",")]:
        a = torch.randn((m, k), generator=g, device=dev, dtype=torch.float32).half()
        out = torch.empty((m, n), dtype=torch.float16, device=dev)
        if args.once:
            B.dense_linear(a, w, out)
            torch.cuda.synchronize()
            continue
        row = {}
        for name in ("ours", "cublas"):
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())

            def chain():
                for _ in range(args.reps):
                    if name == "ours":
                        B.dense_linear(a, w, out, pdl=True, stream=s)
                    else:
                        torch.matmul(a, w.t(), out=out)
            with torch.cuda.stream(s):
                chain()  # warm-up outside the graph (descriptor cache, cuBLAS workspace / heuristics)
                s.synchronize()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=s):
                    chain()
                for _ in range(3):
                    graph.replay()
                s.synchronize()
                best = None
                for _ in range(args.iters):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(s)
                    graph.replay()
                    e1.record(s)
                    s.synchronize()
                    t = e0.elapsed_time(e1) / args.reps
                    best = t if best is None or t < best else best
            row[name] = {"us": round(best * 1e3, 2), "gbs": round(bytes_w / best / 1e6, 1),
                         "hbm_frac": round(bytes_w / best / 1e6 / peak, 4)}
        ref = torch.matmul(a.float(), w.float().t())
        B.dense_linear(a, w, out)
        torch.cuda.synchronize()
        d = (out.float() - ref).abs()
        row["max_rel"] = float((d / ref.abs().clamp_min(1e-2 * float(ref.abs().mean()))).max())
        row["mean_abs"] = float(d.mean())
        res[str(m)] = row
    print(json.dumps({"op": "dense lm_head stream", "n": n, "k": k, "weight_bytes": bytes_w, "peak_gbs": peak,
                      "peak_source": peak_src, "timing": f"best of {args.iters} graph replays of {args.reps} back-to-back launches",
                      "per_width": res}))


if __name__ == "__main__":
    main()
