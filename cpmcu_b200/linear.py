"""Host-side mirror of the reference operator `W4A16GPTQMarlinLinear<T>`
(reference: src/model/w4a16_gptq_marlin/w4a16_gptq_marlin_linear.cuh:11-147): same fields, same
method names and argument meaning, same error behaviour, same tensor-name routing for loading —
with the GEMM and its post-ops going through the C ABI (one fused launch instead of up to three).
"""
from __future__ import annotations

import ctypes

from . import capi as _l


class W4A16GPTQMarlinLinear:
    """dim_in -> dim_out quantised linear.  weight: int32[K/16][2N] Marlin uint4b8; scales: T[K/g][N]."""

    def __init__(self, dim_in: int, dim_out: int, group_size: int, transposed: bool = True,
                 has_bias: bool = False, dtype=None, device="cuda"):
        import torch
        self.dim_in, self.dim_out = dim_in, dim_out
        self.transposed, self.has_bias = transposed, has_bias
        self.is_k_full, self.use_fp32_reduce, self.has_act_order = True, True, False
        self.group_size = group_size
        if group_size == 128:                      # w4a16_gptq_marlin_linear.cuh:58-64
            self.num_groups = dim_in // group_size
        elif group_size == -1:
            self.num_groups = 1
        else:
            raise ValueError("Unsupported group size")
        self.dtype = torch.float16 if dtype is None else dtype
        self.device = torch.device(device)
        self.weight = self.scales = self.bias = self.workspace = self.output = None
        self.qzeros = self.g_idx = self.perm = None  # placeholders, as in the reference

    # -- storage -------------------------------------------------------------------------------
    def init_weight_ptr(self, memory=None):
        """sizes as w4a16_gptq_marlin_linear.cuh:70-82 (memory: ignored, torch owns device memory)"""
        import torch
        k, n = self.dim_in, self.dim_out
        self.weight = torch.empty((k // 16, 2 * n), dtype=torch.int32, device=self.device)
        self.scales = torch.empty((self.num_groups, n), dtype=self.dtype, device=self.device)
        self.workspace = torch.zeros(((n // 64) * 16,), dtype=torch.int32, device=self.device)
        if self.has_bias:
            self.bias = torch.empty((n,), dtype=self.dtype, device=self.device)

    def init_output_ptr(self, memory, num_tokens: int, offset: int = 0) -> int:
        import torch
        self.output = torch.empty((num_tokens, self.dim_out), dtype=self.dtype, device=self.device)
        return offset + num_tokens * self.dim_out * self.output.element_size()

    def load_to_storage(self, name: str, tensor):
        """name routing of w4a16_gptq_marlin_linear.cuh:90-102 ("scales" / "qweight" / "bias")."""
        import torch
        if self.weight is None:
            self.init_weight_ptr()
        t = tensor if isinstance(tensor, torch.Tensor) else torch.as_tensor(tensor)
        if "scales" in name:
            self.scales.copy_(t.reshape(self.scales.shape).to(self.dtype), non_blocking=True)
        elif "qweight" in name:
            self.weight.copy_(t.reshape(self.weight.shape), non_blocking=True)
        elif "bias" in name:
            if self.bias is None:
                raise ValueError("Linear Unsupported name " + name)
            self.bias.copy_(t.reshape(self.bias.shape).to(self.dtype), non_blocking=True)
        else:
            raise ValueError("Linear Unsupported name " + name)

    # -- compute -------------------------------------------------------------------------------
    def prefill(self, stream, num_tokens: int, input, a_tmp, c_tmp, tgt=None, inplace: bool = False):
        """Same contract as the reference's prefill(): result in self.output (tgt None), in tgt
        (not inplace), or tgt += result (inplace); bias added last."""
        if tgt is None and self.output is None:
            self.init_output_ptr(None, num_tokens)
        out = self.output[:num_tokens] if self.output is not None else None
        # the mirror is handed a tensor, so it can state the scratch size (a raw-pointer caller states it once
        # with set_scratch_capacity, or gets the reference's contract)
        c_tmp_bytes = 0 if c_tmp is None else c_tmp.numel() * c_tmp.element_size()
        rc = _l.lib().b200_w4a16_linear_forward_ex(
            _l._ptr(input), _l._ptr(self.weight), _l._ptr(self.scales),
            _l._ptr(self.bias) if self.has_bias else None, _l._ptr(self.workspace),
            _l._ptr(out), _l._ptr(tgt), 1 if inplace else 0, num_tokens, self.dim_out, self.dim_in,
            self.group_size, _l.dtype_code(self.dtype), _l._ptr(a_tmp), _l._ptr(c_tmp), c_tmp_bytes,
            _l._stream(stream))
        if rc:
            _l._raise(rc)
        return tgt if tgt is not None else out

    decode = prefill  # the reference's decode paths call prefill() (w4a16_gptq_marlin_ffn.cuh:67-78)


class Linear:
    """Host-side mirror of the reference's un-quantised `Linear<T>` (src/model/linear.cuh:39-85) for the
    decode / tree-verify widths (num_tokens <= 64): weight T[dim_out][dim_in] (`transposed=True`, the only
    layout CPM.cu instantiates), optional bias, same method names and argument meaning.  The GEMM, the
    inplace add and the bias add are ONE weight-streaming launch through the C ABI."""

    def __init__(self, dim_in: int, dim_out: int, transposed: bool = True, has_bias: bool = False,
                 dtype=None, device="cuda"):
        import torch
        if not transposed:
            raise ValueError("Linear: only transposed=True (weight [dim_out][dim_in]) is supported")
        self.dim_in, self.dim_out = dim_in, dim_out
        self.transposed, self.has_bias = transposed, has_bias
        self.dtype = torch.float16 if dtype is None else dtype
        self.device = torch.device(device)
        self.weight = self.bias = self.output = None
        self.input_scale = 1.0

    def init_weight_ptr(self, memory=None):
        import torch
        self.weight = torch.empty((self.dim_out, self.dim_in), dtype=self.dtype, device=self.device)
        if self.has_bias:
            self.bias = torch.empty((self.dim_out,), dtype=self.dtype, device=self.device)

    def init_output_ptr(self, memory, num_tokens: int, offset: int = 0) -> int:
        import torch
        self.output = torch.empty((num_tokens, self.dim_out), dtype=self.dtype, device=self.device)
        return offset + num_tokens * self.dim_out * self.output.element_size()

    def load_to_storage(self, name: str, tensor):
        """name routing of src/model/linear.cuh:65-73 ("weight" / "bias")."""
        import torch
        if self.weight is None:
            self.init_weight_ptr()
        t = tensor if isinstance(tensor, torch.Tensor) else torch.as_tensor(tensor)
        if "weight" in name:
            self.weight.copy_(t.reshape(self.weight.shape).to(self.dtype), non_blocking=True)
        elif "bias" in name:
            if self.bias is None:
                raise ValueError("Unsupported name " + name)
            self.bias.copy_(t.reshape(self.bias.shape).to(self.dtype), non_blocking=True)
        else:
            raise ValueError("Unsupported name " + name)

    def prefill(self, stream, num_tokens: int, input, tgt=None, inplace: bool = False):
        """src/model/linear.cuh:76-82: result in self.output (tgt None) or tgt; inplace: tgt += result."""
        if tgt is None:
            if self.output is None:
                self.init_output_ptr(None, num_tokens)
            tgt = self.output[:num_tokens]
        rc = _l.lib().b200_w4a16_dense_linear_forward(
            _l._ptr(input), _l._ptr(self.weight), _l._ptr(self.bias) if self.has_bias else None, _l._ptr(tgt),
            1 if inplace else 0, num_tokens, self.dim_out, self.dim_in, float(self.input_scale),
            _l.dtype_code(self.dtype), -1, _l._stream(stream))
        if rc:
            _l._raise(rc)
        return tgt

    decode = prefill


class LMHead(Linear):
    """`LMHead<T>` (src/model/linear.cuh:86-105): the input is scaled by T(head_scale) (rounded to T, as
    elementwise_scale does) before the linear — fused into the same launch, no tmp_hidden_size buffer."""

    def __init__(self, dim_in: int, dim_out: int, head_scale: float = 1.0, transposed: bool = True,
                 has_bias: bool = False, dtype=None, device="cuda"):
        super().__init__(dim_in, dim_out, transposed, has_bias, dtype, device)
        self.head_scale = head_scale
        self.input_scale = head_scale
