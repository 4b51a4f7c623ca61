"""Host side of the one-shot NVLink all-reduce that follows the row-parallel linears (o, down) of a
tensor-parallel model.  torch.distributed + its symmetric-memory allocator are the
plumbing (peer-mapped buffers and their addresses); the collective itself is this library's kernel
(`b200_w4a16_allreduce_oneshot`, csrc/aux_kernels.cuh), not NCCL."""
from __future__ import annotations

import ctypes

from . import capi as _l

MAX_BLOCKS, MAX_WORLD = 128, 8


class OneShotAllReduce:
    """Two alternating symmetric input buffers of `max_elems` elements + signal pads.

        ar = OneShotAllReduce(max_elems, torch.float16, device)
        buf = ar.next_input(m, n)          # the row-parallel GEMM writes its [m, n] partial here
        gemm(..., out=buf)
        ar.reduce(buf, out)                # out[m, n] = sum over ranks, identical on every rank

    Every rank must issue the same sequence of calls.  Capturable in a CUDA graph (an even number of
    reduce() calls per captured graph keeps the buffer alternation consistent across replays)."""

    def __init__(self, max_elems: int, dtype, device, group=None):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        if self.world > MAX_WORLD:
            raise ValueError(f"OneShotAllReduce supports up to {MAX_WORLD} ranks")
        gname = self.group.group_name
        if hasattr(symm, "enable_symm_mem_for_group"):  # needed by older torch, a deprecated no-op in newer
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                try:
                    symm.enable_symm_mem_for_group(gname)
                except Exception:  # noqa: BLE001
                    pass
        max_elems = (max_elems + 7) // 8 * 8
        self.max_elems, self.dtype, self.device = max_elems, dtype, device
        self.data = symm.empty((2, max_elems), dtype=dtype, device=device)
        self.res = symm.empty((2, max_elems), dtype=dtype, device=device)   # two-shot result slices
        self.sig = symm.empty((2, 2 * MAX_BLOCKS * MAX_WORLD), dtype=torch.int32, device=device)
        self.data.zero_()
        self.res.zero_()
        self.sig.zero_()
        torch.cuda.synchronize(device)
        hd = symm.rendezvous(self.data, gname)
        hr = symm.rendezvous(self.res, gname)
        hs = symm.rendezvous(self.sig, gname)
        dist.barrier(self.group)
        esz = self.data.element_size()
        PtrArr = ctypes.c_void_p * self.world
        self._in = [PtrArr(*[int(p) + i * max_elems * esz for p in hd.buffer_ptrs]) for i in range(2)]
        self._res = [PtrArr(*[int(p) + i * max_elems * esz for p in hr.buffer_ptrs]) for i in range(2)]
        self._sig = [PtrArr(*[int(p) + i * 2 * MAX_BLOCKS * MAX_WORLD * 4 for p in hs.buffer_ptrs]) for i in range(2)]
        # per buffer set: int32[2 * MAX_BLOCKS] barrier epochs + the sticky error word (+ padding)
        self.epoch = torch.zeros((2, 2 * MAX_BLOCKS + 8), dtype=torch.int32, device=device)
        self._handles = (hd, hr, hs)
        self._turn = 0

    def check(self):
        """raises if any reduce() so far saw a peer that never arrived (its output was filled with NaN);
        synchronises the device — call it outside timed / captured regions"""
        import torch
        torch.cuda.synchronize(self.device)
        if int(self.epoch[:, 2 * MAX_BLOCKS].max().item()) != 0:
            raise RuntimeError(f"one-shot all-reduce: a peer of rank {self.rank} did not arrive within 2 s "
                               "(dead or desynchronised rank); the affected outputs are NaN")

    def next_input(self, m: int, n: int):
        """the symmetric buffer the NEXT reduce() will read, viewed as [m, n]"""
        assert m * n <= self.max_elems
        return self.data[self._turn, : m * n].view(m, n)

    def reduce(self, buf, out, stream=None, two_shot=None):
        i = self._turn
        assert buf.data_ptr() == self.data[i].data_ptr(), "reduce() must be given the buffer of next_input()"
        elems = buf.numel()
        if two_shot is None:
            # measured on 4 B200s (profiles/r01/f_allreduce_4gpu.txt): the second cross-GPU barrier of the
            # two-shot form costs more than the bytes it saves up to [64, 8192] fp16 -> one-shot by default
            two_shot = False
        rc = _l.lib().b200_w4a16_allreduce(
            ctypes.cast(self._in[i], ctypes.c_void_p), ctypes.cast(self._res[i], ctypes.c_void_p),
            ctypes.cast(self._sig[i], ctypes.c_void_p), _l._ptr(self.epoch[i]), _l._ptr(out), elems,
            _l.dtype_code(self.dtype), self.rank, self.world, 1 if two_shot else 0, _l._stream(stream))
        if rc:
            _l._raise(rc)
        self._turn ^= 1
        return out
