"""Tensor-parallel sharding of the Marlin format (cpmcu_b200/sharding.py) on CPU: slicing identities against
the oracle, and the world_size-2 row-parallel path (partial GEMMs + all-reduce) over gloo."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cpmcu_b200 import sharding as sh
from oracle import marlin_oracle as mo


def _exact(a, q, s, k, n):
    return mo.gemm(a, q, s, k, n, 128, mo.FP16)[1]


def test_even_split_alignment():
    assert sh.even_split(10240, 8, 0, 64) == (0, 1280)
    assert sh.even_split(10240, 8, 7, 64) == (8960, 10240)
    spans = [sh.even_split(4608, 4, r, 64) for r in range(4)]
    assert spans[0][0] == 0 and spans[-1][1] == 4608
    assert all(spans[i][1] == spans[i + 1][0] for i in range(3))
    assert all(b % 64 == 0 and e % 64 == 0 for b, e in spans)
    with pytest.raises(ValueError):
        sh.even_split(100, 2, 0, 64)


@pytest.mark.parametrize("world", [2, 4])
def test_column_shards_are_valid_marlin_tensors(world):
    """a chunk-aligned column slice of (qweight, scales) is the Marlin packing of those columns"""
    m, k, n = 3, 256, 512
    p = mo.make_problem(m, k, n, 128, mo.FP16, seed=1)
    full = _exact(p["a"], p["qweight"], p["scales_marlin"], k, n)
    outs = []
    for r in range(world):
        n0, n1 = sh.even_split(n, world, r, 64)
        q, s = sh.shard_columns(p["qweight"], p["scales_marlin"], n0, n1)
        assert q.shape == (k // 16, 2 * (n1 - n0)) and s.shape == (k // 128, n1 - n0)
        # the shard unpacks to exactly those columns of the integer weights
        assert np.array_equal(mo.marlin_unpack(q), mo.marlin_unpack(p["qweight"])[:, n0:n1])
        outs.append(_exact(p["a"], q, s, k, n1 - n0))
    assert np.array_equal(np.concatenate(outs, axis=1), full)
    with pytest.raises(ValueError):
        sh.shard_columns(p["qweight"], p["scales_marlin"], 32, 128)


def test_fused_column_shards_keep_parts_together():
    """q|k|v fused along N: every rank gets its share of q, of k and of v (gptq2marlin.py:162-230 fuses before packing)"""
    m, k = 2, 256
    parts = [256, 64, 64]  # q, k, v widths
    n = sum(parts)
    p = mo.make_problem(m, k, n, 128, mo.FP16, seed=2)
    full = _exact(p["a"], p["qweight"], p["scales_marlin"], k, n)
    world = 2
    # only q can be split two ways on 64-column chunks; k and v (one chunk each) go to rank 0 / nobody:
    # use chunk-granular parts that do divide
    parts2 = [256, 128, 128]
    n2 = sum(parts2)
    p2 = mo.make_problem(m, k, n2, 128, mo.FP16, seed=3)
    full2 = _exact(p2["a"], p2["qweight"], p2["scales_marlin"], k, n2)
    cols = {0: [], 1: []}
    off = 0
    for w in parts2:
        for r in range(world):
            b, e = sh.even_split(w, world, r, 64)
            cols[r].append((off + b, off + e))
        off += w
    for r in range(world):
        q, s = sh.shard_fused_columns(p2["qweight"], p2["scales_marlin"], parts2, r, world)
        want = np.concatenate([full2[:, b:e] for b, e in cols[r]], axis=1)
        assert np.array_equal(_exact(p2["a"], q, s, k, q.shape[1] // 2), want)
    with pytest.raises(ValueError):
        sh.shard_fused_columns(p["qweight"], p["scales_marlin"], [100, 100], 0, 2)
    assert full.shape == (m, n)


def test_row_shards_sum_to_full():
    m, k, n = 4, 512, 128
    p = mo.make_problem(m, k, n, 128, mo.FP16, seed=4)
    full = _exact(p["a"], p["qweight"], p["scales_marlin"], k, n)
    acc = np.zeros_like(full)
    for r in range(4):
        k0, k1 = sh.even_split(k, 4, r, 128)
        q, s = sh.shard_rows(p["qweight"], p["scales_marlin"], k0, k1)
        acc += _exact(p["a"][:, k0:k1], q, s, k1 - k0, n)
    assert np.allclose(acc, full, rtol=1e-12, atol=1e-9)
    with pytest.raises(ValueError):
        sh.shard_rows(p["qweight"], p["scales_marlin"], 64, 256)


def _tp_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m, k, n = 5, 512, 256
        p = mo.make_problem(m, k, n, 128, mo.FP16, seed=9)  # same seed -> same full tensors on every rank

        def gemm_fn(a, q, s, out):
            kk, nn = a.shape[1], s.shape[1]
            out.copy_(torch.from_numpy(_exact(a.numpy(), q.numpy(), s.numpy(), kk, nn)))

        def all_reduce(t):
            dist.all_reduce(t, op=dist.ReduceOp.SUM)

        qw, sc, a = torch.from_numpy(p["qweight"]), torch.from_numpy(p["scales_marlin"]), torch.from_numpy(p["a"])
        # layer = column-parallel (k -> n) followed by row-parallel (n -> k2), like gate_up -> down
        col = sh.TPW4A16Linear(qw, sc, k, n, "col", rank, world, gemm_fn)
        h_local = torch.empty((m, col.n_local), dtype=torch.float64)
        col.forward(col.local_input(a), h_local)
        full = torch.from_numpy(_exact(p["a"], p["qweight"], p["scales_marlin"], k, n))
        n0, n1 = sh.even_split(n, world, rank, 64)
        ok_col = bool(torch.equal(h_local, full[:, n0:n1]))

        p2 = mo.make_problem(m, n, 128, 128, mo.FP16, seed=10)
        qw2, sc2 = torch.from_numpy(p2["qweight"]), torch.from_numpy(p2["scales_marlin"])
        row = sh.TPW4A16Linear(qw2, sc2, n, 128, "row", rank, world, gemm_fn, all_reduce)
        a2 = torch.from_numpy(p2["a"])
        out = torch.empty((m, 128), dtype=torch.float64)
        row.forward(row.local_input(a2), out)
        full2 = torch.from_numpy(_exact(p2["a"], p2["qweight"], p2["scales_marlin"], n, 128))
        ok_row = bool(torch.allclose(out, full2, rtol=1e-12, atol=1e-9))
        ret[rank] = (ok_col, ok_row, row.k_local, col.n_local)
    finally:
        dist.destroy_process_group()


def test_tensor_parallel_world2_gloo():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_tp_worker, args=(world, port, ret), nprocs=world, join=True)
    assert len(ret) == world
    for r in range(world):
        ok_col, ok_row, k_local, n_local = ret[r]
        assert ok_col and ok_row
        assert k_local == 128 and n_local == 128
