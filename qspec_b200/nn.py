"""Host-side mirror of the reference's dual-mode layer modules
(vllm/model_executor/layers/quarot_nn/{linear,normalization,quantization,hadamard}.py and
quarot.PackedQuantizedTensor, third-party/QuaRot/quarot/__init__.py:154-170).

Same class names, constructor arguments, buffers, `forward(..., **kwargs)` dispatch on
`kwargs["w4a4"]`, scratch-buffer kwargs and error behaviour, so `quarot_llama.py`-style
callers run unchanged. Every forward ends in a C-ABI call (qspec_b200.ops); there is no
torch fallback.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import ops


class PackedQuantizedTensor:
    """quarot/__init__.py:154-170."""

    def __init__(self, quantized_x: torch.Tensor, scales_x: torch.Tensor):
        self.quantized_x = quantized_x
        self.scales_x = scales_x

    def size(self):
        return self.quantized_x.size()

    @property
    def device(self):
        return self.quantized_x.device

    @property
    def dtype(self):
        return self.quantized_x.dtype


# ----------------------------------------------------------------------------------------------
# packing helpers (quarot/functional/quantization.py:41-74), torch versions for weight loading
# ----------------------------------------------------------------------------------------------

def pack_i4(q: torch.Tensor) -> torch.Tensor:
    if not torch.is_signed(q):
        raise AssertionError("The tensor to be packed should be signed int")
    if q.numel() and (int(q.min()) < -8 or int(q.max()) > 7):
        raise AssertionError("int4 values must be in [-8, 7]")
    u = (q.to(torch.int16) & 0xF).to(torch.uint8)
    return u[..., 0::2] | (u[..., 1::2] << 4)


def unpack_i4(x: torch.Tensor) -> torch.Tensor:
    x = x.view(torch.uint8)
    lo = (x & 0xF).to(torch.int8)
    hi = (x >> 4).to(torch.int8)
    lo = torch.where(lo >= 8, lo - 16, lo)
    hi = torch.where(hi >= 8, hi - 16, hi)
    out = torch.stack((lo, hi), dim=-1)
    return out.reshape(*x.shape[:-1], x.shape[-1] * 2).to(torch.int32)


def get_hadK(n: int):
    """(hadK or None, K) as quarot/functional/hadamard.py:get_hadK (same precedence: 108 before 28 before 40) for the
    sizes this build supports: powers of two (K=1), 28 * 2^k (Llama-3 intermediate sizes), 108 * 2^k and 40 * 2^k
    (Qwen2.5-14B: intermediate 13824, 40 heads)."""
    def pow2(v):
        return v > 0 and v & (v - 1) == 0
    if n % 108 == 0 and pow2(n // 108):
        from ._hadk import hadK
        return hadK(108), 108
    if n % 28 == 0 and pow2(n // 28):
        from ._had28 import had28
        return had28(), 28
    if n % 40 == 0 and pow2(n // 40):
        from ._hadk import hadK
        return hadK(40), 40
    if pow2(n):
        return None, 1
    raise NotImplementedError(f"online Hadamard of size {n}: only 2^k, 28*2^k, 108*2^k and 40*2^k are built")


# ----------------------------------------------------------------------------------------------
# Linear4bit  (quarot_nn/linear.py:28-152)
# ----------------------------------------------------------------------------------------------

class Linear4bit(torch.nn.Module):
    """Symmetric 4-bit linear layer: ONE packed weight serves W4A4 (draft) and W4A16 (verify)."""

    def __init__(self, in_features, out_features, bias=False, dtype=torch.float16, **kwargs):
        super().__init__()
        self.in_features = in_features
        self.out_features = out_features
        dev = kwargs.get("device", None)
        self.register_buffer("weight_scales", torch.zeros((out_features, 1), requires_grad=False, dtype=dtype, device=dev))
        if kwargs.get("empty_init", False):      # loaders that overwrite the weight skip the reference's randint init
            w = torch.empty((out_features, in_features // 2), dtype=torch.uint8, device=dev)
        else:
            w = torch.randint(1, 7, (out_features, in_features // 2), dtype=torch.uint8, requires_grad=False, device=dev)
        self.register_buffer("weight", w)
        if bias:
            self.register_buffer("bias", torch.zeros((out_features), dtype=dtype))
        else:
            self.bias = None
        # kept for call compatibility: the reference passes a BitBLAS op under "<K>-<N>"
        # (linear.py:46-48, model_runner.py:1118-1131); the W4A16 kernel here needs no tuned op.
        self.a16_matmul = kwargs.get(f"{in_features}-{out_features}", None)

    def forward(self, x, C=None, **kwargs):
        if kwargs.get("w4a4", False):
            return self.forward_w4a4(x, C)
        elif kwargs.get("w4a8", False):
            return self.forward_w4a8(x, C)
        else:
            return self.forward_w4a16(x, C)

    def _lazy_views(self):
        if self.weight_scales.dim() == 2:                      # linear.py:75-78
            self.weight_scales = self.weight_scales.view(-1)
        if self.weight.dtype != torch.int8:
            self.weight = self.weight.view(torch.int8)

    def forward_w4a4(self, x, C=None):
        assert type(x) == PackedQuantizedTensor                # linear.py:72
        xq, scales_x = x.quantized_x, x.scales_x
        self._lazy_views()
        if C is None:
            C = torch.empty(xq.shape[0], self.weight.shape[0], dtype=self.weight_scales.dtype, device=xq.device)
        ops.rowwise_scaled_linear_cutlass_s4s4_unified(xq, scales_x.view(-1)[: xq.shape[0]], self.weight,
                                                       self.weight_scales, self.bias, C)
        return C

    def forward_w4a8(self, x, C=None):
        """linear.py:86-99: int8 activations (PackedQuantizedTensor carrying [T, K] int8) x the same packed int4 weight."""
        assert type(x) == PackedQuantizedTensor
        xq, scales_x = x.quantized_x, x.scales_x
        self._lazy_views()
        return ops.rowwise_scaled_linear_cutlass_s8s4(xq, scales_x.view(-1)[: xq.shape[0]], self.weight, self.weight_scales,
                                                      self.bias, C)

    def forward_w4a16(self, x, C=None):
        self._lazy_views()
        if C is None:
            C = torch.empty(x.shape[0], self.weight.shape[0], dtype=x.dtype, device=x.device)
        ops.gemm_w4a16(x, self.weight, self.weight_scales, self.bias, C)
        return C

    def forward_split(self, x, partials, w4a4=False):
        """Deferred reduction (qspec_b200 extension, no reference counterpart): the GEMM leaves its per-k-slice partial
        accumulators in `partials`; ops.split_add_rms_norm_* finishes the epilogue inside the next norm."""
        self._lazy_views()
        if w4a4:
            assert type(x) == PackedQuantizedTensor
            xq = x.quantized_x
            if xq.shape[-1] == 2 * self.weight.shape[1]:       # int8 container: one int4 value per byte
                return ops.gemm_w4a4_c8_split(xq, x.scales_x.view(-1)[: xq.shape[0]], self.weight, self.weight_scales,
                                              self.bias, partials)
            return ops.gemm_w4a4_split(xq, x.scales_x.view(-1)[: xq.shape[0]], self.weight, self.weight_scales, self.bias,
                                       partials)
        return ops.gemm_w4a16_split(x, self.weight, self.weight_scales, self.bias, partials)

    @staticmethod
    def from_float(module: torch.nn.Linear, weight_scales=None, **kwargs):
        """linear.py:130-152: round(W / scale) packed to int4."""
        w = module.weight.data
        int_module = Linear4bit(module.in_features, module.out_features, bias=module.bias is not None,
                                dtype=w.dtype, **kwargs).to(w.dtype)
        if weight_scales is not None:
            assert weight_scales.shape == (module.out_features, 1), "weight_scales should have shape (out_features, 1)"
            int_module.weight_scales.copy_(weight_scales.to(w.dtype))
            q = (w.float() / weight_scales.float().to(w.device)).round().to(torch.int8)
            int_module.weight.copy_(pack_i4(q).cpu())
            if module.bias is not None:
                int_module.bias.copy_(module.bias)
        return int_module


def get_matmul_op_w4a16(in_features, out_features, with_scaling=True, enable_tuning=True,
                        fast_decoding=False, bias=False):
    """quarot_nn/linear.py:156-211 returned a tuned BitBLAS op; here it is a thin callable with
    the same call signature `op(x, W, output=C, scale=sW, bias=b)`. W may be the reference's
    offset-binary copy (`weight ^ 0x88`) ONLY via `offset_binary=True`; the layer itself passes
    the two's-complement weight and never makes that copy."""

    def op(x, W, output=None, scale=None, bias=None, offset_binary=False):
        if offset_binary:
            W = W ^ (0x88 if W.dtype == torch.uint8 else -120)
        if output is None:
            output = torch.empty(x.shape[0], W.shape[0], dtype=x.dtype, device=x.device)
        return ops.gemm_w4a16(x, W, scale.view(-1), bias, output)

    op.in_features, op.out_features = in_features, out_features
    return op


# ----------------------------------------------------------------------------------------------
# RMSNorm  (quarot_nn/normalization.py:5-82) -- mean-subtracting, no gamma
# ----------------------------------------------------------------------------------------------

class RMSNorm(torch.nn.Module):
    def __init__(self, mean_dim: int, eps=1e-5, fuse=False):
        super().__init__()
        self.eps = eps
        self.mean_dim = mean_dim
        self.fuse = fuse

    def forward(self, x, **kwargs):
        if kwargs.get("w4a4", False):
            return self.fuse_forward(x, kwargs.get("quantized_buffer_qkv", None),
                                     kwargs.get("scale_buffer", None), kwargs.get("input_sum_buffer", None))
        if kwargs.get("w4a8", False):
            return self.a8_forward(x)
        return self.unfuse_forward(x)

    def a8_forward(self, x, out=None, **kwargs):
        """normalization.py:84-105: norm + per-token int8 quantisation (scale = amax / 127)."""
        hidden = x.shape[-1]
        rows = x.numel() // hidden
        out_q = torch.empty(rows, hidden, dtype=torch.int8, device=x.device)
        both = torch.empty(2, rows, dtype=torch.float16, device=x.device)
        input_sum, scaling_factor = both[0], both[1]
        ops.rms_norm_general_fuse_sum_i8(out_q, x, input_sum, scaling_factor, self.eps, True)
        return PackedQuantizedTensor(out_q, scaling_factor)

    def fuse_forward(self, x, out_q=None, scaling_factor=None, input_sum=None):
        hidden = x.shape[-1]
        rows = x.numel() // hidden
        if out_q is None or input_sum is None or scaling_factor is None:
            out_q = torch.empty(rows, hidden // 2, dtype=torch.int8, device=x.device)
            both = torch.empty(2, rows, dtype=torch.float16, device=x.device)
            input_sum, scaling_factor = both[0], both[1]
        ops.rms_norm_general_fuse_sum_i4(out_q, x, input_sum, scaling_factor, self.eps, True)
        return PackedQuantizedTensor(out_q, scaling_factor)

    def unfuse_forward(self, x, out=None):
        if out is None:
            out = torch.empty_like(x)
        ops.rms_norm_general_fuse_sum_fp16(out, x, self.eps)
        return out


# ----------------------------------------------------------------------------------------------
# Quantizer  (quarot_nn/quantization.py:4-29)
# ----------------------------------------------------------------------------------------------

class Quantizer(torch.nn.Module):
    def __init__(self, input_clip_ratio=1.0, **kwargs):
        super().__init__()
        self.input_clip_ratio = input_clip_ratio
        self.w4a4 = kwargs.get("w4a4", False)

    def forward(self, x, scale=None, out=None, **kwargs):
        if kwargs.get("w4a4", False):
            if scale is None or out is None:
                rows, hidden = x.shape
                scale = torch.empty(rows, dtype=torch.float16, device=x.device)
                out = torch.empty(rows, hidden // 2, dtype=torch.int8, device=x.device)
            ops.fuse_sym_quant(x, scale, out, self.input_clip_ratio)
            return PackedQuantizedTensor(out, scale)
        elif kwargs.get("w4a8", False):                        # quantization.py:20-26
            if scale is None or out is None:
                rows, hidden = x.shape
                scale = torch.empty(rows, dtype=torch.float16, device=x.device)
                out = torch.empty(rows, hidden, dtype=torch.int8, device=x.device)
            ops.fuse_sym_quant_i8(x, scale, out, self.input_clip_ratio)
            return PackedQuantizedTensor(out, scale)
        return x


# ----------------------------------------------------------------------------------------------
# OnlineHadamard  (quarot_nn/hadamard.py:5-49)
# ----------------------------------------------------------------------------------------------

def reference_hadamard_scale(n: int) -> float:
    """The rotation scale as the reference evaluates it AT RUN TIME: `1.0 / torch.tensor(n).sqrt()` (quarot_nn/hadamard.py:20-21,
    quarot/functional/hadamard.py:109,121) is computed after `torch.set_default_dtype(torch.float16)` (worker/model_runner.py:
    1109,1152), so the integer tensor's sqrt and the division both happen in float16: 0.1767578125 for 32 heads (not the fp32
    0.17677669) and 0.00835418701171875 for 14336 (fp32: 0.0083519137). 1e-4 .. 3e-4 relative -- enough to move f16 roundings
    and int4 quantisation steps, so the decode model and the oracle use THIS value."""
    return float(1.0 / torch.tensor(float(n), dtype=torch.float16).sqrt())


class OnlineHadamard(torch.nn.Module):
    def __init__(self, hadamard_dim, force_fp32=False):
        super().__init__()
        if force_fp32:
            raise NotImplementedError("force_fp32 Hadamard is not on the QSpec path (hadamard.py:8)")
        self.fp32_had = False
        had_rem_dim, self.rem_dim = get_hadK(hadamard_dim)
        self.hadamard_dim = hadamard_dim
        # the reference hard-codes scale_32 / scale_14336 = 1.0 / torch.tensor(n).sqrt() (hadamard.py:20-21): evaluated in torch's
        # DEFAULT dtype -- float16 when the model is built the way the reference builds it (see reference_hadamard_scale), float32
        # otherwise. Same expression here, so the module follows the caller's default dtype exactly as the reference's does.
        self.scale = float(1.0 / torch.tensor(float(hadamard_dim)).sqrt())
        if had_rem_dim is not None:
            self.register_buffer("had_rem_dim", had_rem_dim.to(torch.float16))
        else:
            self.had_rem_dim = None
        self.qspec = False

    def forward(self, x, out=None, **kwargs):
        if kwargs.get("w4a4", False) or kwargs.get("w4a8", False):
            self.qspec = True                                  # sticky, hadamard.py:32-33
        n = x.shape[-1]
        if self.rem_dim == 1:
            x2 = x.reshape(-1, n)
            if not x2.is_contiguous():
                x2 = x2.contiguous()
            if out is None:
                out = torch.empty_like(x2)
            ops.faster_fast_hadamard_transform(x2, self.scale, out.view(-1, n))
            return out
        # K = 28 / 108: FHT over n/K then hadK across the K blocks, one fused launch
        x2 = x.reshape(-1, n)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        if out is None:
            out = torch.empty_like(x2)
        if self.rem_dim == 28:
            ops.hadamard_k28(x2, n // 28, self.scale, out=out.view(-1, n))
        elif self.rem_dim == 108:
            ops.hadamard_k108(x2, n // 108, self.scale, out=out.view(-1, n))
        else:
            raise NotImplementedError(f"OnlineHadamard.forward with K = {self.rem_dim} (the reference's Qwen variant rotates "
                                      "across 40 heads in Qwen2Attention.o_had_trans: use ops.hadamard_heads40)")
        return out.view(-1, self.rem_dim, n // self.rem_dim)   # hadamard.py:112 shape
