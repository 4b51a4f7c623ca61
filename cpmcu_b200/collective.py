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


AR_TILES = 1024


class FusedRowParallel:
    """Row-parallel W4A16 linear with the all-reduce fused into the GEMM launch (`b200_w4a16_gemm_ar`, M <= 32):

        fr = FusedRowParallel(max_elems=32 * 8192, dtype=torch.float16, device=dev)
        fr.gemm(a_local, qweight_rows, scales_rows, workspace, c_tmp, out)   # out[m, n] = sum over ranks, same bits on all

    torch.distributed's symmetric-memory allocator provides the peer-mapped receive buffers and counters; the data
    movement and the reduction are this library's kernel.  Every rank must issue the same sequence of calls."""

    def __init__(self, max_elems: int, dtype, device, group=None):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        if not 2 <= self.world <= MAX_WORLD:
            raise ValueError(f"FusedRowParallel needs 2..{MAX_WORLD} ranks")
        gname = self.group.group_name
        if hasattr(symm, "enable_symm_mem_for_group"):
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                try:
                    symm.enable_symm_mem_for_group(gname)
                except Exception:  # noqa: BLE001
                    pass
        self.slot_elems = (max_elems + 7) // 8 * 8
        self.dtype, self.device = dtype, device
        self.recv = symm.empty((2 * self.world * self.slot_elems,), dtype=dtype, device=device)
        self.flag = symm.empty((self.world * AR_TILES,), dtype=torch.int32, device=device)
        self.recv.zero_()
        self.flag.zero_()
        torch.cuda.synchronize(device)
        hr = symm.rendezvous(self.recv, gname)
        hf = symm.rendezvous(self.flag, gname)
        dist.barrier(self.group)
        PtrArr = ctypes.c_void_p * self.world
        self._recv = PtrArr(*[int(p) for p in hr.buffer_ptrs])
        self._flag = PtrArr(*[int(p) for p in hf.buffer_ptrs])
        self.seen = torch.zeros(AR_TILES + 8, dtype=torch.int32, device=device)
        self._handles = (hr, hf)
        self._ar = _l.FusedAR()
        self._ar.recv = ctypes.cast(self._recv, ctypes.c_void_p)
        self._ar.flag = ctypes.cast(self._flag, ctypes.c_void_p)
        self._ar.seen = self.seen.data_ptr()
        self._ar.slot_elems = self.slot_elems
        self._ar.rank, self._ar.world = self.rank, self.world

    def gemm(self, a, qweight, scales, workspace, c_tmp, out, *, group_size=128, bias=None, residual=None,
             out_scale=0.0, pdl=False, stream=None):
        m, k = a.shape
        n = qweight.shape[1] // 2
        assert m * n <= self.slot_elems and out.shape == (m, n)
        o = _l.Opts()
        o.bias = None if bias is None else bias.data_ptr()
        o.residual = None if residual is None else residual.data_ptr()
        o.c_tmp_bytes = 0 if c_tmp is None else c_tmp.numel() * c_tmp.element_size()
        o.out_scale = float(out_scale)
        o.pdl = 1 if pdl else 0
        rc = _l.lib().b200_w4a16_gemm_ar(_l._ptr(a), _l._ptr(qweight), _l._ptr(scales), _l._ptr(workspace), _l._ptr(out),
                                         m, n, k, group_size, _l.dtype_code(a.dtype), _l._ptr(c_tmp), _l._stream(stream),
                                         ctypes.byref(o), ctypes.byref(self._ar))
        if rc:
            _l._raise(rc)
        return out

    def check(self):
        """raises if a source rank failed to deliver a tile within 2 s in any call so far (device sync)"""
        import torch
        torch.cuda.synchronize(self.device)
        if int(self.seen[AR_TILES].item()) != 0:
            raise RuntimeError(f"fused all-reduce: a source rank of rank {self.rank} did not deliver within 2 s")
