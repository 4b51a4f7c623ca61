"""Multi-GPU leg of bench.py: Llama-3-70B-shaped W4A16 linears, tensor-parallel over N B200s
(BASELINE.json configs[4]).  One process per GPU (torchrun), NCCL over NVLink.

Per layer:  qkv 8192->10240 (column-parallel, fused q|k|v sliced per part)  ->  o 8192->8192 (row-parallel +
all-reduce)  ->  gate_up 8192->57344 (column-parallel, gate|up sliced per part)  ->  down 28672->8192
(row-parallel + all-reduce).  One STEP = `layers` such layers at every width M in `widths`, replayed from one
CUDA graph per width (GEMMs and NCCL all-reduces captured together).  Strong scaling: the model is fixed, each
GPU streams 1/N of the weights; `value` = algorithmic weight+scale bytes of the WHOLE model per second.
Weights are synthetic and generated directly in their sharded shape (the slicing itself is tested on CPU
against the oracle: tests/test_sharding_cpu.py).
"""
import json
import os
import sys
import time

SHAPES_70B = [  # name, K, N, style, fused parts along N
    ("qkv", 8192, 10240, "col", [8192, 1024, 1024]),
    ("o", 8192, 8192, "row", None),
    ("gate_up", 8192, 57344, "col", [28672, 28672]),
    ("down", 28672, 8192, "row", None),
]
GROUP = 128


def alg_bytes(k, n):
    return k * n // 2 + (k // GROUP) * n * 2


def local_dims(k, n, style, parts, rank, world):
    from cpmcu_b200 import sharding as sh
    if style == "col":
        n_local = sum(e - b for b, e in (sh.even_split(p, world, rank, 64) for p in (parts or [n])))
        return k, n_local
    k0, k1 = sh.even_split(k, world, rank, 128)
    return k1 - k0, n


def check_tp_result(torch, dist, dev, world, rank, widths, ar, stream):
    """Outside the timed region, once per width: a tensor-parallel layer computed from REAL shards of one full
    weight (cpmcu_b200/sharding.py) must equal the unsharded GEMM of this library — column-parallel: the rank's
    column slices; row-parallel + all-reduce: the full output, identical on every rank."""
    import cpmcu_b200 as B
    from cpmcu_b200 import sharding as sh
    report = {}
    for name, k, n, style, parts in (("qkv", 8192, 10240, "col", [8192, 1024, 1024]), ("o", 8192, 8192, "row", None)):
        g = torch.Generator(device=dev).manual_seed(99)           # same seed on every rank: identical full tensors
        q = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
        s = (torch.rand((k // GROUP, n), device=dev, generator=g) * 0.018 + 0.002).half()
        ws = torch.zeros((n // 64) * 16, dtype=torch.int32, device=dev)
        for m in widths:
            a = torch.randn((m, k), device=dev, generator=g).half()
            ct = torch.empty(B.reduce_bytes_fast(m, n, k) // 4 + 64, dtype=torch.float32, device=dev)
            full = B.gemm(a, q, s, ws, ct, stream=stream)
            if style == "col":
                ql, sl = sh.shard_fused_columns(q, s, parts, rank, world)
                loc = B.gemm(a, ql, sl, ws, ct, stream=stream)
                cols, off = [], 0
                for p in parts:
                    b, e = sh.even_split(p, world, rank, 64)
                    cols.append(full[:, off + b:off + e])
                    off += p
                want = torch.cat(cols, dim=1)
                stream.synchronize()
                err = float((loc.float() - want.float()).abs().max())
                ok = err <= 1e-2 * float(want.float().abs().mean()) + 1e-3   # split-K order differs between the two launches
            else:
                k0, k1 = sh.even_split(k, world, rank, 128)
                ql, sl = sh.shard_rows(q, s, k0, k1)
                al = a[:, k0:k1].contiguous()
                out = torch.empty((m, n), dtype=torch.float16, device=dev)
                if ar is not None:
                    buf = ar.next_input(m, n)
                    B.gemm(al, ql, sl, ws, ct, out=buf, stream=stream)
                    ar.reduce(buf, out, stream=stream)
                else:
                    B.gemm(al, ql, sl, ws, ct, out=out, stream=stream)
                    stream.synchronize()
                    dist.all_reduce(out)
                stream.synchronize()
                torch.cuda.synchronize()
                err = float((out.float() - full.float()).abs().max())
                # partials are rounded to fp16 before the fp32 sum: W roundings instead of one
                ok = err <= 1e-2 * float(full.float().abs().mean()) * max(1, world) ** 0.5 + 1e-3
                gathered = [torch.empty_like(out) for _ in range(world)]
                dist.all_gather(gathered, out)
                ok = ok and all(torch.equal(gathered[0], t) for t in gathered)   # bit-identical on every rank
            report[f"{name}/m{m}"] = round(err, 5)
            if not ok:
                raise RuntimeError(f"[rank {rank}] tensor-parallel {name} at M={m} differs from the unsharded GEMM: max abs err {err}")
    return report


def measure_tp(args, torch, dev, dist, rank, world, local_rank, use_reference=False, steps=None):
    """one measurement of the Llama-3-70B linears over `world` ranks (world = 1: unsharded on this GPU); returns the
    JSON line as a dict on every rank (values are max-over-ranks)."""
    from bench import ClockSampler, peaks
    pk = peaks()
    widths = [int(x) for x in args.tp_widths.split(",")]
    L = args.tp_layers
    eff_world, eff_rank = world, rank
    if use_reference:
        from oracle import ref_marlin
        ref_marlin.lib()
    else:
        import cpmcu_b200 as B
        B.lib()
        B.set_pdl(not args.no_pdl)

    g = torch.Generator(device=dev).manual_seed(1234 + eff_rank)
    lins = []   # per shape: dict(k_local, n_local, style, q[L], s[L], ws)
    for name, k, n, style, parts in SHAPES_70B:
        kl, nl = local_dims(k, n, style, parts, eff_rank, eff_world)
        q = torch.randint(-2**31, 2**31 - 1, (L, kl // 16, 2 * nl), dtype=torch.int32, device=dev, generator=g)
        s = (torch.rand((L, kl // GROUP, nl), device=dev, generator=g) * 0.018 + 0.002).half()
        ws = torch.zeros((nl // 64) * 16, dtype=torch.int32, device=dev)
        lins.append(dict(name=name, k=kl, n=nl, style=style, q=q, s=s, ws=ws))
    model_bytes = sum(alg_bytes(k, n) for _, k, n, _, _ in SHAPES_70B) * L   # whole model, one width
    local_bytes = sum(alg_bytes(d["k"], d["n"]) for d in lins) * L

    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    acts_host = {m: [torch.randn((m, d["k"]), generator=torch.Generator().manual_seed(m + i)).half().pin_memory()
                     for i, d in enumerate(lins)] for m in widths}
    acts = {m: [t.to(dev) for t in acts_host[m]] for m in widths}
    outs = {m: [torch.empty((m, d["n"]), dtype=torch.float16, device=dev) for d in lins] for m in widths}
    res_host = {m: torch.empty((m, lins[3]["n"]), dtype=torch.float16).pin_memory() for m in widths}
    if use_reference:
        scratch = max(ref_marlin.reduce_max_m(m) * d["n"] for m in widths for d in lins)
    else:
        scratch = max(B.reduce_bytes_fast(m, d["n"], d["k"]) // 4 for m in widths for d in lins)
    c_tmp = torch.empty(scratch + 64, dtype=torch.float32, device=dev)
    n_launch = [0]

    # the collective after the row-parallel linears: this library's one-shot NVLink all-reduce over
    # symmetric (peer-mapped) buffers; NCCL all-reduce if symmetric memory is unavailable or --collective nccl
    ar = None
    collective = "none"
    if eff_world > 1:
        collective = "NCCL all-reduce (fp16, sum) inside the CUDA graph"
        if getattr(args, "collective", "fused") in ("oneshot", "fused"):
            try:
                from cpmcu_b200.collective import OneShotAllReduce
                ar = OneShotAllReduce(max(widths) * max(d["n"] for d in lins if d["style"] == "row"),
                                      torch.float16, dev)
                collective = ("one-shot all-reduce kernel of this library over NVLink peer memory "
                              "(b200_w4a16_allreduce_oneshot; fp32 sum in rank order), inside the CUDA graph")
            except Exception as e:  # noqa: BLE001
                sys.stderr.write(f"[rank {rank}] symmetric memory unavailable ({e!r}); using NCCL all-reduce\n")
                ar = None
        flag = torch.tensor([1 if ar is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)   # every rank must take the same path
        if int(flag.item()) == 0:
            ar, collective = None, "NCCL all-reduce (fp16, sum) inside the CUDA graph"
    # the all-reduce fused into the row-parallel GEMM launch (M <= 32); wider steps keep the separate collective
    fused = None
    if eff_world > 1 and not use_reference and getattr(args, "collective", "fused") == "fused":
        from cpmcu_b200.collective import FusedRowParallel
        fused = FusedRowParallel(32 * max(d["n"] for d in lins if d["style"] == "row"), torch.float16, dev)
        collective = ("M <= 32: all-reduce fused into the row-parallel GEMM launch (b200_w4a16_gemm_ar: tiles pushed to "
                      "every peer over NVLink, reduced locally in rank order); M > 32: " + collective)
    tp_check = None
    if eff_world > 1 and not use_reference:
        tp_check = check_tp_result(torch, dist, dev, eff_world, eff_rank, widths, ar, stream)

    def gemm(i, l, m, out):
        d = lins[i]
        if use_reference:
            ref_marlin.gemm(acts[m][i], d["q"][l], d["s"][l], d["ws"], c_tmp, out=out, stream=stream)
        else:
            B.gemm(acts[m][i], d["q"][l], d["s"][l], d["ws"], c_tmp, out=out, pdl=not args.no_pdl,
                   stream=stream)
        n_launch[0] += 1

    def one_pass(m, comm=True, compute=True):
        for l in range(L):
            for i, d in enumerate(lins):
                reduce_here = comm and d["style"] == "row" and eff_world > 1
                if reduce_here and fused is not None and m <= 32 and compute:
                    fused.gemm(acts[m][i], d["q"][l], d["s"][l], d["ws"], c_tmp, outs[m][i], pdl=not args.no_pdl, stream=stream)
                    n_launch[0] += 1
                    continue
                if reduce_here and ar is not None:
                    buf = ar.next_input(m, d["n"])      # the GEMM writes its partial into the symmetric buffer
                    if compute:
                        gemm(i, l, m, buf)
                    ar.reduce(buf, outs[m][i], stream=stream)
                    n_launch[0] += 1
                    continue
                if compute:
                    gemm(i, l, m, outs[m][i])
                if reduce_here:
                    dist.all_reduce(outs[m][i])

    def capture(m, comm=True, compute=True):
        one_pass(m, comm, compute)
        stream.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=stream):
            one_pass(m, comm, compute)
        return gr

    graphs, graphs_nocomm, graphs_comm = {}, {}, {}
    launches_per_step = 0
    for m in widths:
        n0 = n_launch[0]
        graphs[m] = capture(m)
        launches_per_step += (n_launch[0] - n0) // 2
        if eff_world > 1:
            graphs_nocomm[m] = capture(m, comm=False)
            graphs_comm[m] = capture(m, compute=False)
    stream.synchronize()

    def barrier():
        if eff_world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, nsteps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(nsteps):
            fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / nsteps], device=dev)
        if eff_world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def kernel_step():
        for m in widths:
            graphs[m].replay()

    def e2e_step():
        for m in widths:
            for i in range(len(lins)):
                acts[m][i].copy_(acts_host[m][i], non_blocking=True)
            graphs[m].replay()
            res_host[m].copy_(outs[m][3], non_blocking=True)

    for _ in range(max(args.warmup, 3)):
        kernel_step()
    K = steps if steps else args.steps
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_per_step = timed(kernel_step, K)
    per_width = {}
    for m in widths:
        t_all = timed(graphs[m].replay, max(3, K // 2))
        ent = dict(ms=round(t_all, 4), gbs=round(model_bytes / t_all / 1e6, 1))
        if eff_world > 1:
            t_g = timed(graphs_nocomm[m].replay, max(3, K // 2))
            t_c = timed(graphs_comm[m].replay, max(3, K // 2))
            ent.update(gemm_only_ms=round(t_g, 4), allreduce_only_ms=round(t_c, 4),
                       allreduce_share=round(max(0.0, t_all - t_g) / t_all, 4))
        per_width[str(m)] = ent
    for _ in range(3):
        e2e_step()
    e2e_ms = timed(e2e_step, K)
    clocks = sampler.stop()
    if fused is not None:
        fused.check()
    if ar is not None:
        ar.check()   # a peer that never arrived is an error (NaN output + error word), never a silently wrong sum
    if eff_world > 1:
        dist.barrier()

    step_bytes = model_bytes * len(widths)
    value = step_bytes / (ms_per_step * 1e-3) / 1e9
    low = [m for m in widths if m <= 32] or widths
    low_ms = sum(per_width[str(m)]["ms"] for m in low)
    achieved = model_bytes * len(low) / (low_ms * 1e-3) / 1e9
    peak = pk["hbm"] * eff_world
    h2d = sum(2 * m * d["k"] for m in widths for d in lins)
    d2h = sum(2 * m * lins[3]["n"] for m in widths)
    line = {
        "metric": "w4a16_gemm_weight_stream_GBps", "value": round(value, 1), "unit": "GB/s", "n_gpus": world,
        "steps": K, "warmup": max(args.warmup, 3), "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": f"Llama-3-70B W4A16 linears (qkv 8192x10240 col, o 8192x8192 row+allreduce, gate_up "
                               f"8192x57344 col, down 28672x8192 row+allreduce; group 128) x {L} layers at widths "
                               f"M={widths}, tensor-parallel over {eff_world} GPU(s)",
                   "parallelism": f"tp{eff_world}", "collective": collective,
                   "l2": f"inputs larger than L2: {local_bytes / 2**30:.2f} GiB of weights per GPU per width",
                   "impl": "reference Marlin, unsharded on one GPU (the reference has no multi-GPU path)"
                   if use_reference else "cpmcu_b200 (sm_100a)"},
        "per_width": per_width,
        "e2e": {"value": round(step_bytes / (e2e_ms * 1e-3) / 1e9, 1), "unit": "GB/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": round(e2e_ms, 4)},
        "gpu_launches": launches_per_step * K * eff_world,
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": round(peak, 1), "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": None,
                     "kernel": "marlin::Marlin" if use_reference else "w4a16_decode_kernel",
                     "note": f"widths {low}: whole-model algorithmic bytes / max-over-ranks device time; peak = "
                             f"{eff_world} x {pk['hbm']} GB/s ({pk['source']})"},
        "clocks": clocks,
    }
    if tp_check is not None:
        line["tp_result_check"] = {"max_abs_err_vs_unsharded": tp_check,
                                   "note": "sharded (sharding.py) vs unsharded GEMM of one full 70B qkv / o weight per width, "
                                           "outside the timed region; row-parallel result bit-identical on every rank"}
    del graphs, graphs_nocomm, graphs_comm, lins, acts, outs, c_tmp
    torch.cuda.empty_cache()
    return line


def main_tp(args, rank, world, local_rank, use_reference=False):
    import torch
    import torch.distributed as dist

    if use_reference and rank != 0:
        return  # the reference arm is one process on one GPU; the other ranks exit without work
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1 and not use_reference:
        dist.init_process_group("nccl", device_id=dev)
    if use_reference:
        # the reference has no multi-GPU path: rank 0 runs the UNSHARDED model on one GPU through the
        # unmodified reference Marlin kernel; the other ranks exit without work
        line = measure_tp(args, torch, dev, None, 0, 1, local_rank, use_reference=True)
        line["n_gpus"] = world
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": line["value"], "unit": "GB/s", "cores": 0, "kind": "reference",
                                "sample": "GPU kernel of the reference (it has no CPU implementation of this path): "
                                          "oracle/_ref/libmarlin_ref.so, unsharded model on rank 0"}
        print(json.dumps(line), flush=True)
        return
    # the tp = 1 base of the scaling curve: THIS library, unsharded, same layers and widths, on rank 0's GPU
    tp1 = None
    if rank == 0 and world > 1:
        tp1 = measure_tp(args, torch, dev, None, 0, 1, local_rank, steps=max(3, args.steps // 4))
    if world > 1:
        dist.barrier()
    line = measure_tp(args, torch, dev, dist if world > 1 else None, rank, world, local_rank)
    if rank != 0:
        # leave without tearing NCCL down: destroying a communicator whose collectives were captured in
        # live CUDA graphs can hang (seen on the first 2-GPU run); the process is done anyway
        sys.stdout.flush()
        os._exit(0)
    if tp1 is not None:
        line["llama70b_tp1"] = {"value": tp1["value"], "unit": "GB/s", "ms_per_step": tp1["ms_per_step"],
                                "per_width": tp1["per_width"]}
        line["efficiency_vs_own_tp1"] = round(line["value"] / (world * tp1["value"]), 4)
    print(json.dumps(line), flush=True)
    if world > 1:
        sys.stdout.flush()
        time.sleep(0.2)
        os._exit(0)
