"""Tensor-parallel sharding of Marlin-packed W4A16 linears (BASELINE.json configs[4]).

The reference has no multi-GPU code; this is the natural sharding of its weight format:

  column-parallel (split N; qkv, gate_up):   qweight int32[K/16][2N] -> columns [2*n0, 2*n1)
                                             scales  T[K/g][N]       -> columns [n0, n1)
      n0, n1 multiples of 64: a 64-column chunk is 128 consecutive int32 of a row, and the scale
      permutation (gptq2marlin.py:55-57,99-107) stays inside a 64-column chunk, so chunk-aligned
      slices are valid Marlin tensors.  Fused tensors (q|k|v, gate|up: gptq2marlin.py:162-230) are
      sliced PER PART and re-concatenated so every rank holds matching heads / gate-up pairs.
      No collective: the outputs feed the rank-local next op.
  row-parallel (split K; o, down):           qweight rows [k0/16, k1/16), scale rows [k0/128, k1/128)
      k0, k1 multiples of 128 (the group).  Every rank produces a full [M][N] partial; one
      all-reduce(sum) over NVLink finishes the layer.

All functions work on numpy arrays and torch tensors alike (pure slicing / concatenation), so the
host logic is testable on CPU; the GEMM itself always goes through the C ABI on the GPU.
"""
from __future__ import annotations

from typing import Callable, Optional, Sequence


def _cat(parts, axis):
    first = parts[0]
    if hasattr(first, "numpy") or type(first).__module__.startswith("torch"):
        import torch
        return torch.cat(list(parts), dim=axis).contiguous()
    import numpy as np
    return np.ascontiguousarray(np.concatenate(list(parts), axis=axis))


def _contig(x):
    return x.contiguous() if hasattr(x, "contiguous") else __import__("numpy").ascontiguousarray(x)


def even_split(total: int, world: int, rank: int, align: int) -> tuple[int, int]:
    """[begin, end) of `rank`'s share of `total`, in units of `align`, as even as possible."""
    if total % align:
        raise ValueError(f"size {total} is not a multiple of {align}")
    units = total // align
    b = (units * rank) // world
    e = (units * (rank + 1)) // world
    return b * align, e * align


def shard_columns(qweight, scales, n0: int, n1: int):
    """columns [n0, n1) of a Marlin linear; n0, n1 multiples of 64"""
    if n0 % 64 or n1 % 64 or not (0 <= n0 < n1):
        raise ValueError(f"column shard [{n0}, {n1}) must be 64-aligned and non-empty")
    return _contig(qweight[:, 2 * n0:2 * n1]), _contig(scales[:, n0:n1])


def shard_rows(qweight, scales, k0: int, k1: int, group_size: int = 128):
    """rows [k0, k1) of a Marlin linear; k0, k1 multiples of the group (128), or of 16 when channelwise"""
    align = 128 if group_size == 128 else 16
    if k0 % align or k1 % align or not (0 <= k0 < k1):
        raise ValueError(f"row shard [{k0}, {k1}) must be {align}-aligned and non-empty")
    q = _contig(qweight[k0 // 16:k1 // 16, :])
    s = _contig(scales[k0 // 128:k1 // 128, :]) if group_size == 128 else _contig(scales)
    return q, s


def shard_fused_columns(qweight, scales, parts: Sequence[int], rank: int, world: int, align: int = 64):
    """column shard of a fused tensor whose N axis is the concatenation of `parts` (e.g. [q, k, v] or
    [gate, up]): rank r gets its even share of EVERY part, concatenated in part order."""
    n_total = scales.shape[1]
    if sum(parts) != n_total:
        raise ValueError(f"parts {list(parts)} do not add up to N = {n_total}")
    qs, ss = [], []
    off = 0
    for p in parts:
        b, e = even_split(p, world, rank, align)
        if e > b:
            q, s = shard_columns(qweight, scales, off + b, off + e)
            qs.append(q)
            ss.append(s)
        off += p
    return _cat(qs, 1), _cat(ss, 1)


class TPW4A16Linear:
    """One tensor-parallel W4A16 linear.  style = "col" (split N, no collective) or "row" (split K,
    all-reduce of the [M][N] partial).  `gemm_fn(a, qweight, scales, out)` computes the local GEMM
    (the C ABI on a GPU; tests inject the oracle), `all_reduce_fn(t)` sums `t` over ranks in place
    (torch.distributed over NCCL / NVLink on the GPUs, gloo in the CPU tests)."""

    def __init__(self, qweight, scales, k: int, n: int, style: str, rank: int, world: int,
                 gemm_fn: Callable, all_reduce_fn: Optional[Callable] = None,
                 parts: Optional[Sequence[int]] = None, group_size: int = 128):
        if style not in ("col", "row"):
            raise ValueError("style must be 'col' or 'row'")
        self.style, self.rank, self.world = style, rank, world
        self.k, self.n = k, n
        self.gemm_fn, self.all_reduce_fn = gemm_fn, all_reduce_fn
        if style == "col":
            self.qweight, self.scales = shard_fused_columns(qweight, scales, parts or [n], rank, world)
            self.k_local, self.n_local = k, self.scales.shape[1]
            self.k0 = 0
        else:
            k0, k1 = even_split(k, world, rank, 128 if group_size == 128 else 64)
            self.qweight, self.scales = shard_rows(qweight, scales, k0, k1, group_size)
            self.k_local, self.n_local = k1 - k0, n
            self.k0 = k0
        if style == "row" and world > 1 and all_reduce_fn is None:
            raise ValueError("row-parallel linear needs all_reduce_fn")

    def local_input(self, a_full):
        """the slice of a full [M][K] activation this rank consumes (row-parallel: its K range)"""
        return a_full if self.style == "col" else _contig(a_full[:, self.k0:self.k0 + self.k_local])

    def forward(self, a_local, out):
        self.gemm_fn(a_local, self.qweight, self.scales, out)
        if self.style == "row" and self.world > 1:
            self.all_reduce_fn(out)
        return out
