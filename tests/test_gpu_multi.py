"""Multi-GPU parity (-m gpu; skipped below 2 visible GPUs): tensor-parallel W4A16 linears on REAL shards.

One process per GPU (torch.multiprocessing), NCCL for the plumbing, this library's NVLink all-reduce kernel for the
collective.  Llama-3-70B shapes of BASELINE.json configs[4]: qkv 8192->10240 (column-parallel, fused parts sliced
per part), o 8192->8192 and down 28672->8192 (row-parallel + all-reduce), fp16 and bf16, M = 1 / 16 / 64, one-shot
and two-shot reduction, launched from a CUDA graph and replayed.  Checks, per case:
  * column-parallel: the rank's output equals the matching columns of the UNSHARDED b200_w4a16_gemm (tolerance: the
    split-K order of the two launches differs);
  * row-parallel: all-reduced output within tolerance of the unsharded GEMM and of the exact product (fp64 matmul of
    the library's bit-exact dequant) — max floored relative error <= 1e-2 for fp16, <= 3e-2 for bf16 (the partials
    are rounded to T before the fp32 sum: W + 1 roundings of an 8-bit significand) — BIT-IDENTICAL on every rank and
    from replay to replay;
  * a peer that never calls the collective produces an ERROR on the waiting rank (error word + NaN output), not a sum.
"""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = {torch.float16: 1e-2, torch.bfloat16: 3e-2}   # bf16 carries 8 significand bits: W rounded partials + one more rounding


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rel(x, ref):
    """max |x - ref| / max(|ref|, mean |ref|): the floored relative error of oracle.marlin_oracle.error_stats"""
    ref = ref.double()
    d = (x.double() - ref).abs()
    return float((d / torch.maximum(ref.abs(), ref.abs().mean())).max())


def _worker(rank, world, port, results):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist
    import cpmcu_b200 as B
    from cpmcu_b200 import sharding as sh
    from cpmcu_b200.collective import FusedRowParallel, OneShotAllReduce
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    log = []
    try:
        stream = torch.cuda.Stream(device=dev)
        torch.cuda.set_stream(stream)
        for tdt in (torch.float16, torch.bfloat16):
            ar = OneShotAllReduce(64 * 8192, tdt, dev)
            fr = FusedRowParallel(32 * 8192, tdt, dev)
            for name, k, n, style, parts in (("qkv", 8192, 10240, "col", [8192, 1024, 1024]),
                                             ("o", 8192, 8192, "row", None), ("down", 28672, 8192, "row", None)):
                g = torch.Generator(device=dev).manual_seed(k + n)      # same seed everywhere: identical full tensors
                q = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=dev, generator=g)
                s = (torch.rand((k // 128, n), device=dev, generator=g) * 0.018 + 0.002).to(tdt)
                ws = torch.zeros((n // 64) * 16, dtype=torch.int32, device=dev)
                w64 = B.dequant(q, s, k, n).double()
                if style == "col":
                    ql, sl = sh.shard_fused_columns(q, s, parts, rank, world)
                else:
                    k0, k1 = sh.even_split(k, world, rank, 128)
                    ql, sl = sh.shard_rows(q, s, k0, k1)
                for m in (1, 16, 64):
                    a = torch.randn((m, k), device=dev, generator=g).to(tdt)
                    ct = torch.empty(B.reduce_bytes_fast(m, n, k) // 4 + 64, dtype=torch.float32, device=dev)
                    full = B.gemm(a, q, s, ws, ct, stream=stream)
                    exact = a.double() @ w64
                    assert _rel(full, exact) <= TOL[tdt]
                    if style == "col":
                        loc = B.gemm(a, ql, sl, ws, ct, stream=stream)
                        cols, off = [], 0
                        for p in parts:
                            b, e = sh.even_split(p, world, rank, 64)
                            cols.append(full[:, off + b:off + e])
                            off += p
                        want = torch.cat(cols, dim=1)
                        stream.synchronize()
                        r = _rel(loc, want)
                        assert r <= TOL[tdt], (name, m, tdt, r)
                        log.append((name, str(tdt), m, "col", r))
                        continue
                    al = a[:, k0:k1].contiguous()
                    if m <= 32:
                        # the all-reduce fused into the GEMM launch (b200_w4a16_gemm_ar): same checks, plus post-ops
                        # applied AFTER the sum: RN(RN(res + sum) * T(scale)) on the fused result == on a plain one
                        out = torch.empty((m, n), dtype=tdt, device=dev)

                        def fstep():
                            for _ in range(3):
                                fr.gemm(al, ql, sl, ws, ct, out, pdl=True, stream=stream)
                        fstep()
                        stream.synchronize()
                        gr = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(gr, stream=stream):
                            fstep()
                        gr.replay()
                        stream.synchronize()
                        first = out.clone()
                        for _ in range(5):
                            gr.replay()
                        stream.synchronize()
                        fr.check()
                        assert torch.equal(first, out), (name, m, tdt, "fused: not repeatable")
                        gathered = [torch.empty_like(out) for _ in range(world)]
                        dist.all_gather(gathered, out)
                        assert all(torch.equal(gathered[0], t) for t in gathered), (name, m, tdt, "fused: ranks differ")
                        r_full, r_exact = _rel(out, full), _rel(out, exact)
                        assert r_exact <= TOL[tdt] and r_full <= TOL[tdt], (name, m, tdt, "fused", r_full, r_exact)
                        res = torch.randn((m, n), device=dev, generator=g).to(tdt)
                        out2 = torch.empty_like(out)
                        fr.gemm(al, ql, sl, ws, ct, out2, residual=res, out_scale=0.25, stream=stream)
                        stream.synchronize()
                        assert torch.equal(out2, (res + out) * torch.tensor(0.25, device=dev).to(tdt)), (name, m, tdt, "fused post-ops")
                        assert int(ws.abs().max()) == 0
                        log.append((name, str(tdt), m, "fused", r_exact))
                        del gr
                    for two_shot in (False, True):
                        out = torch.empty((m, n), dtype=tdt, device=dev)

                        def step():
                            for _ in range(2):      # an even number of reduces per graph keeps the buffer alternation
                                buf = ar.next_input(m, n)
                                B.gemm(al, ql, sl, ws, ct, out=buf, pdl=True, stream=stream)
                                ar.reduce(buf, out, stream=stream, two_shot=two_shot)
                        step()
                        stream.synchronize()
                        gr = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(gr, stream=stream):
                            step()
                        gr.replay()
                        stream.synchronize()
                        first = out.clone()
                        for _ in range(5):
                            gr.replay()
                        stream.synchronize()
                        ar.check()
                        assert torch.equal(first, out), (name, m, tdt, two_shot, "not repeatable")
                        gathered = [torch.empty_like(out) for _ in range(world)]
                        dist.all_gather(gathered, out)
                        assert all(torch.equal(gathered[0], t) for t in gathered), (name, m, tdt, two_shot, "ranks differ")
                        r_full, r_exact = _rel(out, full), _rel(out, exact)
                        # W fp16/bf16-rounded partials are summed in fp32: W roundings instead of one
                        assert r_exact <= TOL[tdt] and r_full <= TOL[tdt], (name, m, tdt, two_shot, r_full, r_exact)
                        assert int(ws.abs().max()) == 0
                        log.append((name, str(tdt), m, "two-shot" if two_shot else "one-shot", r_exact))
                        del gr
            del ar, fr
        # ---- a peer that never arrives: error, not a wrong sum ----
        ar = OneShotAllReduce(8192, torch.float16, dev)
        dist.barrier()
        if rank == 0:
            buf = ar.next_input(1, 8192)
            buf.fill_(1.0)
            out = torch.zeros((1, 8192), dtype=torch.float16, device=dev)
            ar.reduce(buf, out, stream=stream)
            with pytest.raises(RuntimeError, match="did not arrive"):
                ar.check()
            assert bool(torch.isnan(out).all()), "a timed-out all-reduce must not return a plausible sum"
            log.append(("dead-peer", "float16", 1, "error raised", 0.0))
        dist.barrier()
        results[rank] = ("ok", log)
    except BaseException as e:  # noqa: BLE001
        import traceback
        results[rank] = ("fail", traceback.format_exc() + repr(e))
    finally:
        os._exit(0)   # no NCCL teardown: destroying communicators with live captured graphs can hang


def test_tensor_parallel_linears_and_nvlink_allreduce():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    world = 2 if n < 4 else 4
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    results = mgr.dict()
    port = _free_port()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, world, port, results)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=900)
    for p in procs:
        if p.is_alive():
            p.kill()
    for r in range(world):
        assert r in results, f"rank {r} died without a result"
        status, payload = results[r]
        assert status == "ok", f"rank {r}:\n{payload}"
    cases = results[0][1]
    assert len(cases) >= 2 * (3 + 2 * 3 * 2 + 2 * 2) + 1
    print("\n" + "\n".join(f"{c}" for c in cases))
