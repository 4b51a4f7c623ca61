#!/usr/bin/env python
"""bench.py — the W4A16 GPTQ-Marlin linear path on B200 (BASELINE.json: "W4A16 GEMM GB/s & TFLOPS vs
B200 roofline; MiniCPM4-8B decode tok/s").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (N = 1; BASELINE.json configs[1]): the quantised linears of a 32-layer MiniCPM4-8B
(qkv 4096->4608, o 4096->4096, gate_up 4096->32768, down 16384->4096; group 128, fp16) at the
speculative tree-verify widths M = 1, 4, 16, 64.  One STEP = one pass over all 32 layers x 4 linears
at each of the four widths = 512 GEMMs streaming 4 x 3.91 GB of weights + scales.  The 3.64 GiB of
weights are far larger than L2 (126 MB), so every launch streams from HBM (no flush needed).
Synthetic data: uniformly random int4 nibbles, scales U(0.002, 0.02), activations N(0, 1).

`value`  = algorithmic weight+scale bytes per second (K*N/2 + (K/128)*N*2 per GEMM),
           device-timed with CUDA events, inputs resident in HBM, CUDA-graph replays (the reference
           also decodes from a captured graph, src/entry.cu:540-562).
`e2e`    = the same metric through the operator API (W4A16GPTQMarlinLinear.prefill -> C ABI) with the
           step's activations copied from pinned host memory and the step's results copied back
           inside the timed region.  Weights stay resident: loading them is the reference's
           load_model step (w4a16_gptq_marlin_linear.cuh:90-102), not part of the path.
`--impl reference` runs the UNMODIFIED reference Marlin kernel (oracle/_ref/libmarlin_ref.so: the
reference's src/qgemm/gptq_marlin compiled in place for sm_100) through the identical harness on the
same GPU; if that library is absent it times the CPU oracle port instead.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SHAPES_8B = [("qkv", 4096, 4608), ("o", 4096, 4096), ("gate_up", 4096, 32768), ("down", 16384, 4096)]
# Llama-3-70B (BASELINE.json configs[4]); parallel style per linear: col = split N, row = split K + allreduce
SHAPES_70B = [("qkv", 8192, 10240, "col"), ("o", 8192, 8192, "row"), ("gate_up", 8192, 57344, "col"),
              ("down", 28672, 8192, "row")]
WIDTHS = [1, 4, 16, 64]
GROUP = 128


def alg_bytes(k, n):
    return k * n // 2 + (k // GROUP) * n * 2


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d["bf16_tflops_sustained"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001
            self.proc = None

    def stop(self):
        out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        if self.proc is None:
            return out
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------------
# arms
# ------------------------------------------------------------------------------------------------
class OursArm:
    """cpmcu_b200 through its operator API (W4A16GPTQMarlinLinear.prefill -> C ABI)."""
    name = "cpmcu_b200 (sm_100a)"

    def __init__(self, torch, dev, pdl=True):
        import cpmcu_b200 as B
        self.B, self.torch, self.dev = B, torch, dev
        B.lib()
        B.set_pdl(pdl)
        self.pdl = pdl

    def make_linear(self, k, n, qweight, scales):
        lin = self.B.W4A16GPTQMarlinLinear(k, n, GROUP, dtype=self.torch.float16, device=self.dev)
        lin.weight, lin.scales = qweight, scales
        lin.workspace = self.torch.zeros(self.B.workspace_ints(n), dtype=self.torch.int32, device=self.dev)
        return lin

    def scratch_floats(self, m, n, k):
        return self.B.reduce_bytes_fast(m, n, k) // 4

    def run(self, lin, m, a, out, c_tmp, stream):
        lin.prefill(stream, m, a, None, c_tmp, tgt=out)

    def launches(self):
        return self.B.launch_count()


class RefMarlinArm:
    """the unmodified reference Marlin kernel (sm_100 recompile) through oracle/ref_shim.cu."""
    name = "reference Marlin (CPM.cu src/qgemm/gptq_marlin, sm_100 recompile, unmodified)"

    def __init__(self, torch, dev):
        from oracle import ref_marlin
        self.R, self.torch, self.dev = ref_marlin, torch, dev
        ref_marlin.lib()
        self.n_launch = 0

    def make_linear(self, k, n, qweight, scales):
        ws = self.torch.zeros((n // 64) * 16, dtype=self.torch.int32, device=self.dev)
        return dict(k=k, n=n, w=qweight, s=scales, ws=ws)

    def scratch_floats(self, m, n, k):
        return self.R.reduce_max_m(m) * n

    def run(self, lin, m, a, out, c_tmp, stream):
        self.R.gemm(a, lin["w"], lin["s"], lin["ws"], c_tmp, out=out, stream=stream)
        self.n_launch += 1

    def launches(self):
        return self.n_launch


# ------------------------------------------------------------------------------------------------
def cpu_baseline(torch, qw0, sc0, shapes, m, budget_s=25.0):
    """the CPU oracle (plain-C OpenMP port, else numpy + torch matmul) on a bounded sample: the four
    linears of ONE layer at width m.  Returns the cpu_baseline object."""
    import numpy as np
    a_host = {name: np.random.default_rng(0).standard_normal((m, k)).astype(np.float16) for name, k, n in shapes}
    host = [(name, k, n, qw0[i].cpu().numpy(), sc0[i].cpu().numpy()) for i, (name, k, n) in enumerate(shapes)]
    total_bytes = sum(alg_bytes(k, n) for _, k, n in shapes)
    t_used, reps = 0.0, 0
    try:
        from oracle import oracle_c
        if not oracle_c.available():
            raise RuntimeError("liboracle_c.so not built")
        oracle_c.lib()
        cores, kind_note = oracle_c.threads(), "oracle/marlin_oracle.c (OpenMP)"

        def one_pass():
            for name, k, n, q, s in host:
                oracle_c.gemm(a_host[name], q, s, k, n, GROUP, 0)
    except Exception:  # noqa: BLE001
        from oracle import marlin_oracle as mo
        torch.set_num_threads(os.cpu_count() or 1)
        cores, kind_note = torch.get_num_threads(), "oracle/marlin_oracle.py dequant + torch fp32 matmul"

        def one_pass():
            for name, k, n, q, s in host:
                w = mo.dequant(q, s, k, n, GROUP, mo.FP16)
                torch.from_numpy(a_host[name]).float() @ torch.from_numpy(w).float()
    while reps < 1 or (t_used < budget_s / 2 and reps < 5):
        t0 = time.perf_counter()
        one_pass()
        t_used += time.perf_counter() - t0
        reps += 1
    sec = t_used / reps
    return dict(value=total_bytes / sec / 1e9, unit="GB/s", cores=cores, kind="port",
                sample=f"{kind_note}: 1 of 32 layers (qkv, o, gate_up, down) at M={m}, {reps} pass(es), "
                       f"{sec:.2f} s per pass", seconds_per_layer_pass=sec)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--widths", default=",".join(map(str, WIDTHS)))
    ap.add_argument("--no-pdl", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-detail", action="store_true")
    ap.add_argument("--no-lm-head", action="store_true", help="skip the un-quantised lm_head leg (SURVEY 8f-2)")
    ap.add_argument("--no-prefill", action="store_true", help="skip the chunked-prefill leg (BASELINE.json configs[2])")
    ap.add_argument("--prefill-m", default="1024,4096,354", help="chunk sizes of the prefill leg (354 = 106850 %% 2048, the ragged last chunk)")
    ap.add_argument("--prefill-layers", type=int, default=4)
    ap.add_argument("--no-tp1", action="store_true", help="skip the unsharded Llama-3-70B record (the tp = 1 base of the scaling curve)")
    ap.add_argument("--workload", default="auto", choices=["auto", "minicpm4", "llama70b"],
                    help="auto: MiniCPM4-8B linears at N=1, tensor-parallel Llama-3-70B linears at N>1")
    ap.add_argument("--collective", default="fused", choices=["fused", "oneshot", "nccl"],
                    help="all-reduce after the row-parallel linears (N > 1): this library's one-shot NVLink kernel, or NCCL")
    ap.add_argument("--tp-layers", type=int, default=16)
    ap.add_argument("--tp-widths", default="1,16,64")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    widths = [int(x) for x in args.widths.split(",")]

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: cpmcu_b200 has no CPU path")

    if world > 1 or args.workload == "llama70b":
        from bench_tp import main_tp  # tensor-parallel Llama-3-70B shapes (configs[4])
        return main_tp(args, rank, world, local_rank, use_reference=(args.impl == "reference"))

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    pk = peaks()

    use_ref_kernel = False
    if args.impl == "reference":
        from oracle import ref_marlin
        use_ref_kernel = ref_marlin.available()
        if not use_ref_kernel:
            return reference_cpu_arm(args, torch, dev, widths)
    arm = RefMarlinArm(torch, dev) if use_ref_kernel else OursArm(torch, dev, pdl=not args.no_pdl)

    # ---- weights (resident in HBM, as after the reference's load_model) ----
    g = torch.Generator(device=dev).manual_seed(1234)
    L = args.layers
    qw, sc, lins = [], [], []
    for name, k, n in SHAPES_8B:
        q = torch.randint(-2**31, 2**31 - 1, (L, k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
        s = (torch.rand((L, k // GROUP, n), device=dev, generator=g) * 0.018 + 0.002).half()
        qw.append(q)
        sc.append(s)
        lins.append([arm.make_linear(k, n, q[l], s[l]) for l in range(L)])
    layer_bytes = sum(alg_bytes(k, n) for _, k, n in SHAPES_8B)
    pass_bytes = layer_bytes * L                      # one width
    step_bytes = pass_bytes * len(widths)
    step_flops = sum(2 * m * k * n for m in widths for _, k, n in SHAPES_8B) * L

    # ---- activations / outputs / scratch ----
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    maxm = max(widths)
    acts_host = {m: [torch.randn((m, k), generator=torch.Generator().manual_seed(m + i)).half().pin_memory()
                     for i, (_, k, n) in enumerate(SHAPES_8B)] for m in widths}
    acts = {m: [t.to(dev) for t in acts_host[m]] for m in widths}
    outs = {m: [torch.empty((m, n), dtype=torch.float16, device=dev) for _, k, n in SHAPES_8B] for m in widths}
    res_host = {m: torch.empty((m, SHAPES_8B[3][2]), dtype=torch.float16).pin_memory() for m in widths}
    c_tmp = torch.empty(max(arm.scratch_floats(m, n, k) for m in widths for _, k, n in SHAPES_8B) + 64,
                        dtype=torch.float32, device=dev)

    def one_pass(m):
        for l in range(L):
            for i in range(4):
                arm.run(lins[i][l], m, acts[m][i], outs[m][i], c_tmp, stream)

    # ---- capture one graph per width (warm-up launch first: module load, attributes, descriptors) ----
    graphs = {}
    launches_per_step = 0
    for m in widths:
        one_pass(m)
        stream.synchronize()
        n0 = arm.launches()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=stream):
            one_pass(m)
        launches_per_step += arm.launches() - n0
        graphs[m] = gr
    stream.synchronize()

    def kernel_step(evs=None):
        for j, m in enumerate(widths):
            graphs[m].replay()
            if evs is not None:
                evs[j + 1].record(stream)

    def e2e_step():
        for m in widths:
            for i in range(4):
                acts[m][i].copy_(acts_host[m][i], non_blocking=True)
            graphs[m].replay()
            res_host[m].copy_(outs[m][3], non_blocking=True)

    # ---- kernel-only: W warm-up steps, then exactly K timed steps ----
    for _ in range(args.warmup):
        kernel_step()
    stream.synchronize()
    K = args.steps
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(widths) + 1)] for _ in range(K)]
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t_begin.record(stream)
    for s_i in range(K):
        ev[s_i][0].record(stream)
        kernel_step(ev[s_i])
    t_end.record(stream)
    torch.cuda.synchronize()
    total_ms = t_begin.elapsed_time(t_end)
    ms_per_step = total_ms / K
    per_width_ms = {m: sum(ev[s_i][j].elapsed_time(ev[s_i][j + 1]) for s_i in range(K)) / K
                    for j, m in enumerate(widths)}

    # ---- end to end: host activations in, host results out, every step ----
    for _ in range(args.warmup):
        e2e_step()
    torch.cuda.synchronize()
    t_begin.record(stream)
    for _ in range(K):
        e2e_step()
    t_end.record(stream)
    torch.cuda.synchronize()
    e2e_ms = t_begin.elapsed_time(t_end) / K
    clocks = sampler.stop()
    h2d = sum(2 * m * k for m in widths for _, k, n in SHAPES_8B)
    d2h = sum(2 * m * SHAPES_8B[3][2] for m in widths)

    # ---- per (shape, width) detail, outside the timed region: one graph of L same-shape GEMMs ----
    detail = {}
    if not args.no_detail:
        for i, (name, k, n) in enumerate(SHAPES_8B):
            detail[name] = {}
            for m in widths:
                def chain():
                    for l in range(L):
                        arm.run(lins[i][l], m, acts[m][i], outs[m][i], c_tmp, stream)
                chain()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=stream):
                    chain()
                for _ in range(3):
                    gr.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                best = 1e30
                for _ in range(5):
                    e0.record(stream)
                    gr.replay()
                    e1.record(stream)
                    stream.synchronize()
                    best = min(best, e0.elapsed_time(e1))
                us = best / L * 1e3
                gbs = alg_bytes(k, n) / us / 1e3
                tf = 2.0 * m * k * n / us / 1e6
                detail[name][str(m)] = dict(us=round(us, 3), gbs=round(gbs, 1), hbm_frac=round(gbs / pk["hbm"], 4),
                                            tflops=round(tf, 2), tensor_frac=round(tf / pk["tf_burst"], 4))
                del gr

    # ---- roofline of the dominant kernel: the HBM-bound streaming kernel (widths <= 32) ----
    low = [m for m in widths if m <= 32] or widths
    low_ms = sum(per_width_ms[m] for m in low)
    achieved = pass_bytes * len(low) / (low_ms * 1e-3) / 1e9
    # traffic: DRAM bytes need ncu counters, which bench.py does not collect (a number taken under a profiler is not
    # a bench value); the per-round capture is committed under profiles/ and quoted in DESIGN.md
    traffic, traffic_note = None, "; traffic: null here (ncu capture of the current kernel: see profiles/r02)"
    roofline = dict(bound="hbm", achieved=round(achieved, 1), peak=pk["hbm"], unit="GB/s",
                    frac=round(achieved / pk["hbm"], 4), traffic=traffic,
                    kernel="w4a16_decode_kernel" if not use_ref_kernel else "marlin::Marlin",
                    note=f"all launches at widths {low}: algorithmic weight+scale bytes / summed device time of "
                         f"those passes (back-to-back launches, gaps charged to the kernel); peak = {pk['source']}; "
                         f"frac of nominal 8000 GB/s = {achieved / 8000.0:.4f}" + traffic_note)

    value = step_bytes / (ms_per_step * 1e-3) / 1e9
    e2e_value = step_bytes / (e2e_ms * 1e-3) / 1e9
    line = {
        "metric": "w4a16_gemm_weight_stream_GBps",
        "value": round(value, 1), "unit": "GB/s", "n_gpus": 1, "steps": K, "warmup": args.warmup,
        "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16", "data": "synthetic",
        "config": {"workload": f"MiniCPM4-8B W4A16 linears (qkv 4096x4608, o 4096x4096, gate_up 4096x32768, "
                               f"down 16384x4096; group 128) x {L} layers at tree-verify widths M={widths}; "
                               f"one step = {4 * L * len(widths)} GEMMs",
                   "l2": "inputs larger than L2: 3.64 GiB of weights streamed per width, no flush needed",
                   "launch": "one CUDA graph per width" + ("" if use_ref_kernel else f", PDL={'on' if arm.pdl else 'off'}"),
                   "impl": arm.name},
        "tflops": round(step_flops / (ms_per_step * 1e-3) / 1e12, 2),
        "per_width": {str(m): dict(ms=round(per_width_ms[m], 4),
                                   gbs=round(pass_bytes / per_width_ms[m] / 1e6, 1),
                                   hbm_frac=round(pass_bytes / per_width_ms[m] / 1e6 / pk["hbm"], 4),
                                   gemm_only_steps_per_s=round(1e3 / per_width_ms[m], 1)) for m in widths},
        "e2e": {"value": round(e2e_value, 1), "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(e2e_ms, 4)},
        "gpu_launches": launches_per_step * K,
        "roofline": roofline,
        "clocks": clocks,
        "detail": detail,
    }
    if use_ref_kernel:
        line["impl"] = "reference"
        line["roofline"]["note"] += "; this is the reference's kernel, reported for comparison"
    if not args.no_prefill:
        line["prefill"] = prefill_leg(torch, dev, arm, stream, pk, local_rank,
                                      [int(x) for x in args.prefill_m.split(",")], args.prefill_layers)
    if not args.no_lm_head:
        line["lm_head"] = lm_head_leg(torch, dev, stream, widths, pk, ours=not use_ref_kernel)
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(torch, [q[0] for q in qw], [s[0] for s in sc], SHAPES_8B, 16)
    if not args.no_tp1:
        # the unsharded Llama-3-70B linears (configs[4]) on ONE GPU: the tp = 1 base of the 1 -> 8 scaling curve
        del qw, sc, lins, graphs, acts, outs
        torch.cuda.empty_cache()
        from bench_tp import measure_tp
        t1 = measure_tp(args, torch, dev, None, 0, 1, local_rank, use_reference=use_ref_kernel, steps=max(3, K // 4))
        line["llama70b_tp1"] = {"value": t1["value"], "unit": "GB/s", "ms_per_step": t1["ms_per_step"],
                                "per_width": t1["per_width"], "workload": t1["config"]["workload"]}
    print(json.dumps(line), flush=True)


def prefill_leg(torch, dev, arm, stream, pk, gpu_index, ms, layers, seconds=1.5):
    """BASELINE.json configs[2]: chunked prefill.  The four MiniCPM4-8B linears x `layers` layers (distinct weights:
    4 layers = 489 MB > L2) at chunk sizes M, one CUDA graph per M, replayed back to back for >= `seconds` seconds so
    that the number is a SUSTAINED one under the power cap and the clock sampler sees the region.  Same harness for
    both arms: ours = one persistent tcgen05 launch per linear; reference = marlin_mm's `par` M loop
    (gptq_marlin_mm.cu:161-206).  TFLOP/s = 2*M*N*K / device time, against the measured dense bf16 peaks of
    MEASURED_PEAKS.json (burst: a kernel timed alone; sustained: inside a long loop) and the nominal 2250."""
    g = torch.Generator(device=dev).manual_seed(4321)
    qw, sc, lins = [], [], []
    for name, k, n in SHAPES_8B:
        q = torch.randint(-2**31, 2**31 - 1, (layers, k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
        s = (torch.rand((layers, k // GROUP, n), device=dev, generator=g) * 0.018 + 0.002).half()
        qw.append(q)
        sc.append(s)
        lins.append([arm.make_linear(k, n, q[l], s[l]) for l in range(layers)])
    out = {}
    for m in ms:
        acts = [torch.randn((m, k), device=dev, generator=g).half() for _, k, n in SHAPES_8B]
        outs = [torch.empty((m, n), dtype=torch.float16, device=dev) for _, k, n in SHAPES_8B]
        c_tmp = torch.empty(max(arm.scratch_floats(m, n, k) for _, k, n in SHAPES_8B) + 64, dtype=torch.float32, device=dev)

        def chain():
            for l in range(layers):
                for i in range(4):
                    arm.run(lins[i][l], m, acts[i], outs[i], c_tmp, stream)
        chain()
        stream.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=stream):
            chain()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e30
        for _ in range(3):   # warm-up, also the burst figure (a replay timed alone)
            e0.record(stream)
            gr.replay()
            e1.record(stream)
            stream.synchronize()
            best = min(best, e0.elapsed_time(e1))
        reps = max(5, int(seconds * 1e3 / best))
        sampler = ClockSampler(gpu_index)
        sampler.start()
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(reps):
            gr.replay()
        e1.record(stream)
        torch.cuda.synchronize()
        clocks = sampler.stop()
        ms_layer = e0.elapsed_time(e1) / reps / layers
        flops = sum(2.0 * m * k * n for _, k, n in SHAPES_8B)
        tf = flops / (ms_layer * 1e-3) / 1e12
        tf_burst = flops / (best / layers * 1e-3) / 1e12
        out[str(m)] = dict(us_per_layer=round(ms_layer * 1e3, 2), tflops=round(tf, 1),
                           frac_burst=round(tf / pk["tf_burst"], 4), frac_sustained=round(tf / pk["tf_sustained"], 4),
                           frac_nominal_2250=round(tf / 2250.0, 4), replays=reps,
                           tflops_first_replays=round(tf_burst, 1), clocks=clocks)
        del gr, acts, outs, c_tmp
    return {"workload": f"chunked prefill: MiniCPM4-8B W4A16 linears (qkv, o, gate_up, down) x {layers} layers at chunk sizes "
                        f"M={ms}; one CUDA graph per M replayed back to back for >= {seconds} s (sustained, power cap included)",
            "kernel": arm.name, "peaks": {"burst": pk["tf_burst"], "sustained": pk["tf_sustained"], "nominal": 2250.0,
                                          "source": pk["source"]},
            "per_m": out}


def lm_head_leg(torch, dev, stream, widths, pk, ours, reps=16, iters=5):
    """SURVEY.md section 8(f)-2, reported beside the headline metric (not part of `value`): the un-quantised
    lm_head (73448 x 4096 fp16 = 602 MB > L2, streamed from HBM every launch) at the same widths — this
    library's streaming kernel (ours) or cuBLAS via torch.matmul, which is what the reference's linear<T>()
    calls (src/model/linear.cuh:8-37)."""
    import cpmcu_b200 as B
    n, k = 73448, 4096
    g = torch.Generator(device=dev).manual_seed(7)
    w = (torch.randn((n, k), generator=g, device=dev, dtype=torch.float32) * 0.02).half()
    res = {}
    for m in widths:
        if m > 64:
            continue
        a = torch.randn((m, k), generator=g, device=dev, dtype=torch.float32).half()
        out = torch.empty((m, n), dtype=torch.float16, device=dev)

        def chain():
            for _ in range(reps):
                if ours:
                    B.dense_linear(a, w, out, pdl=True, stream=stream)
                else:
                    torch.matmul(a, w.t(), out=out)
        chain()
        stream.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=stream):
            chain()
        for _ in range(3):
            gr.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e30
        for _ in range(iters):
            e0.record(stream)
            gr.replay()
            e1.record(stream)
            stream.synchronize()
            best = min(best, e0.elapsed_time(e1))
        us = best / reps * 1e3
        gbs = n * k * 2 / us / 1e3
        res[str(m)] = dict(us=round(us, 2), gbs=round(gbs, 1), hbm_frac=round(gbs / pk["hbm"], 4),
                           hbm_frac_nominal_8000=round(gbs / 8000.0, 4))
        del gr
    return {"workload": "un-quantised lm_head 73448x4096 fp16, weight bytes / device time, back-to-back launches "
                        "in one CUDA graph" + (", PDL on" if ours else ""),
            "kernel": "dense_stream_kernel (b200_w4a16_dense_linear_forward)" if ours else "cuBLAS via torch.matmul",
            "note": "peak = the copy (read+write) bandwidth of MEASURED_PEAKS.json; a read-only stream can exceed it",
            "per_width": res}


def reference_cpu_arm(args, torch, dev, widths):
    """fallback of --impl reference when oracle/_ref/libmarlin_ref.so is absent: the CPU oracle port,
    each step a bounded sample (one layer at M=16)."""
    g = torch.Generator().manual_seed(1234)
    qw, sc = [], []
    for name, k, n in SHAPES_8B:
        qw.append(torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, generator=g))
        sc.append((torch.rand((k // GROUP, n), generator=g) * 0.018 + 0.002).half())
    vals = []
    for _ in range(max(1, min(args.steps, 3))):
        vals.append(cpu_baseline(torch, qw, sc, SHAPES_8B, 16, budget_s=1.0))
    cb = vals[-1]
    v = statistics.median(x["value"] for x in vals)
    cb["value"] = v
    print(json.dumps({
        "impl": "reference", "metric": "w4a16_gemm_weight_stream_GBps", "value": round(v, 3), "unit": "GB/s",
        "n_gpus": 1, "steps": len(vals), "warmup": 0, "ms_per_step": round(cb["seconds_per_layer_pass"] * 1e3, 2),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": "MiniCPM4-8B W4A16 linears, CPU oracle port, bounded sample: 1 layer at M=16 per step"},
        "cpu_baseline": cb,
        "e2e": {"value": round(v, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


if __name__ == "__main__":
    main()
