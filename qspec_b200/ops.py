"""Tensor-level operators over the C-ABI library.

The function names and argument meaning mirror the reference's native entry points
 so `quarot_nn`-style callers can switch by changing the import:

  layernorm_ops.rms_norm_general_fuse_sum_i4   third-party/kernels/csrc/layernorm.cpp:32-38
  layernorm_ops.rms_norm_general_fuse_sum_fp16 third-party/kernels/csrc/layernorm.cpp:40-42
  quarot._CUDA.fuse_sym_quant                  third-party/QuaRot/quarot/kernels/bindings.cpp:128-146
  faster_fast_hadamard_transform               third-party/fast-hadamard-transform/csrc/fast_hadamard_transform.cpp:69-110
  rowwise_scaled_linear_cutlass_s4s4_unified   third-party/ao/torchao/ops.py:600-636
  bitblas Matmul (W4A16)                       vllm/model_executor/layers/quarot_nn/linear.py:156-211

All outputs are caller-allocated, contiguous, on the same CUDA device; launches go to
torch.cuda.current_stream(). PyTorch is used for memory and streams only.
"""
from __future__ import annotations

from typing import Optional

import os

import torch

from . import _lib

_F16 = torch.float16
_BF16 = torch.bfloat16


def _dt(t: torch.Tensor) -> int:
    if t.dtype == _F16:
        return _lib.DTYPE_F16
    if t.dtype == _BF16:
        return _lib.DTYPE_BF16
    raise ValueError(f"expected float16 or bfloat16 tensor, got {t.dtype}")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _chk(t: Optional[torch.Tensor], name: str, dtypes=None, allow_none=False):
    if t is None:
        if allow_none:
            return
        raise ValueError(f"{name} must not be None")
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (qspec_b200 has no CPU path)")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    if dtypes is not None and t.dtype not in dtypes:
        raise ValueError(f"{name}: dtype {t.dtype} not in {dtypes}")


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


_I4 = (torch.int8, torch.uint8)

# ---------------------------------------------------------------------------------------------
# workspaces: one zero-initialised, self-cleaning buffer per (device, kind); grown on demand.
# A workspace that has been handed out is never freed: a CUDA graph captured earlier keeps its pointer (and relies on
# its counter region being zero), so growth allocates a new block and retires the old one alive. Use is single-stream
# per device: two streams running GEMMs of the same kind concurrently would share the tile counters -- give a second
# stream its own workspace through the C ABI (include/qspec_b200.h) instead.
# ---------------------------------------------------------------------------------------------
_workspaces: dict = {}
_retired: list = []


def _workspace(kind: str, nbytes: int, device: torch.device) -> torch.Tensor:
    key = (kind, device.index if device.index is not None else torch.cuda.current_device())
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        if ws is not None:
            _retired.append(ws)
        ws = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def reserve_workspaces(max_tokens: int, shapes, device) -> None:
    """Pre-size both GEMM workspaces (call before CUDA-graph capture). shapes = [(N, K), ...]."""
    lib = _lib.load()
    device = torch.device(device)
    n4 = max(lib.qspec_gemm_w4a4_workspace_bytes(max_tokens, n, k) for n, k in shapes)
    n16 = max(lib.qspec_gemm_w4a16_workspace_bytes(m, n, k) for n, k in shapes
              for m in sorted({1, 8, 16, 32, 64, max_tokens}) if m <= max(max_tokens, 1))
    _workspace("w4a4", n4, device)
    _workspace("w4a16", n16, device)


# ---------------------------------------------------------------------------------------------
# (A) norm (+quant)
# ---------------------------------------------------------------------------------------------

def rms_norm_general_fuse_sum_i4(out_q: torch.Tensor, x: torch.Tensor, input_sum: Optional[torch.Tensor],
                                 scaling: torch.Tensor, eps: float, use_per_token_quant: bool = True) -> None:
    if not use_per_token_quant:
        raise NotImplementedError("only the per-token branch is on the QSpec path (layernorm_kernels.cu:903)")
    _chk(x, "input", (_F16,)); _chk(out_q, "out", _I4); _chk(scaling, "scaling", (_F16,))
    _chk(input_sum, "input_sum", (_F16,), allow_none=True)
    K = x.shape[-1]
    rows = x.numel() // K if K else 0
    if out_q.numel() != rows * K // 2 or scaling.numel() < rows:
        raise ValueError("rms_norm_general_fuse_sum_i4: output shapes do not match the input")
    if rows == 0:
        return
    lib = _lib.load()
    _lib.check(lib.qspec_norm_quant_i4(_ptr(x), _ptr(out_q), _ptr(scaling), _ptr(input_sum), rows, K,
                                       float(eps), _stream()), "rms_norm_general_fuse_sum_i4")


def add_rms_norm_fuse_sum_i4(out_q: torch.Tensor, x: torch.Tensor, delta: torch.Tensor, input_sum: Optional[torch.Tensor],
                             scaling: torch.Tensor, eps: float) -> None:
    """x += delta (in place), then rms_norm_general_fuse_sum_i4(out_q, x, ...): one launch (quarot_llama.py:380,389)."""
    _chk(x, "input", (_F16,)); _chk(delta, "delta", (_F16,)); _chk(out_q, "out", _I4); _chk(scaling, "scaling", (_F16,))
    _chk(input_sum, "input_sum", (_F16,), allow_none=True)
    K = x.shape[-1]
    rows = x.numel() // K if K else 0
    if delta.numel() != x.numel() or out_q.numel() != rows * K // 2 or scaling.numel() < rows:
        raise ValueError("add_rms_norm_fuse_sum_i4: shapes do not match the input")
    if rows == 0:
        return
    lib = _lib.load()
    _lib.check(lib.qspec_add_norm_quant_i4(_ptr(x), _ptr(delta), _ptr(out_q), _ptr(scaling), _ptr(input_sum), rows, K,
                                           float(eps), _stream()), "add_rms_norm_fuse_sum_i4")


def add_rms_norm_fuse_sum_fp16(out: torch.Tensor, x: torch.Tensor, delta: torch.Tensor, eps: float) -> None:
    """x += delta (in place), then rms_norm_general_fuse_sum_fp16(out, x, eps): one launch."""
    _chk(x, "input", (_F16, _BF16)); _chk(delta, "delta", (x.dtype,)); _chk(out, "out", (x.dtype,))
    K = x.shape[-1]
    rows = x.numel() // K if K else 0
    if out.numel() != x.numel() or delta.numel() != x.numel():
        raise ValueError("add_rms_norm_fuse_sum_fp16: out and delta must have the input's shape")
    if rows == 0:
        return
    lib = _lib.load()
    _lib.check(lib.qspec_add_norm_f16(_ptr(x), _ptr(delta), _ptr(out), rows, K, float(eps), _dt(x), _stream()),
               "add_rms_norm_fuse_sum_fp16")


def tp_two_step(peers, rows: int) -> bool:
    """Which form of the fused all-reduce an exchange of `rows` token rows uses -- a function of (rows, world) only, so every
    rank decides alike. Two-step (row r summed by rank r % world, pushed to the peers) when (world - 1) * rows >=
    QSPEC_TP_SCATTER_MIN. OFF by default: measured on 8 x B200 (profiles/r02s_two_step_ab.md) the one-step form is not bound by
    the (world - 1) x rows x K x 2 bytes every rank reads but by the hand-over latency between GPUs, and the second hand-over of
    the two-step form costs more than its smaller traffic saves (TP = 8: -1 % on the 256-token verify exchanges alone, +10 % with
    the 64-token draft exchanges too; TP = 4: +4 %). Kept as a tested option for larger exchanges (prefill-sized passes)."""
    if not hasattr(peers, "_reduced"):
        return False
    return (peers.world - 1) * rows >= int(os.environ.get("QSPEC_TP_SCATTER_MIN", str(1 << 30))) and rows <= 1024


def tp_add_rms_norm_fuse_sum_i4(out_q: torch.Tensor, x: torch.Tensor, peers, which: int, input_sum: torch.Tensor,
                                scaling: torch.Tensor, eps: float, phases: int = 3, c8: bool = False) -> None:
    """x += sum over ranks of peers.partial(which) (rank order, fp32, one rounding), then the quantising norm:
    the all-reduce of RowParallelLinear.forward (layers/linear.py:1143-1144) fused into the consumer."""
    _chk(x, "input", (_F16,)); _chk(out_q, "out", _I4); _chk(scaling, "scaling", (_F16,)); _chk(input_sum, "input_sum", (_F16,))
    K = x.shape[-1]
    rows = x.numel() // K if K else 0
    if K != peers.hidden or rows > peers.max_tokens or rows == 0 or peers.buf.dtype != x.dtype:
        raise ValueError("tp_add_rms_norm_fuse_sum_i4: input does not match the peer partial buffers")
    lib = _lib.load()
    _check_container(out_q, rows, K, c8, "tp_add_rms_norm_fuse_sum_i4")
    if c8:
        _lib.check(lib.qspec_tp_add_norm_quant_i4_c8(_ptr(x), peers._partials[which], peers._flags, _ptr(peers.state), peers.world,
                                                     peers.rank, _ptr(out_q), _ptr(scaling), _ptr(input_sum), rows, K,
                                                     float(eps), _stream()), "tp_add_rms_norm_fuse_sum_i4 (int8 container)")
        return
    if tp_two_step(peers, rows):
        _lib.check(lib.qspec_tp_rs_add_norm_quant_i4(_ptr(x), peers._partials[which], peers._flags, peers._reduced[which],
                                                     peers._row_flags, _ptr(peers.state), peers.world, peers.rank, phases,
                                                     _ptr(out_q), _ptr(scaling), _ptr(input_sum), rows, K, float(eps),
                                                     _stream()), "tp_add_rms_norm_fuse_sum_i4 (two-step)")
        return
    _lib.check(lib.qspec_tp_add_norm_quant_i4(_ptr(x), peers._partials[which], peers._flags, _ptr(peers.state), peers.world,
                                              peers.rank, _ptr(out_q), _ptr(scaling), _ptr(input_sum), rows, K,
                                              float(eps), _stream()), "tp_add_rms_norm_fuse_sum_i4")


def tp_add_rms_norm_fuse_sum_fp16(out: torch.Tensor, x: torch.Tensor, peers, which: int, eps: float, phases: int = 3) -> None:
    _chk(x, "input", (_F16, _BF16)); _chk(out, "out", (x.dtype,))
    K = x.shape[-1]
    rows = x.numel() // K if K else 0
    if K != peers.hidden or rows > peers.max_tokens or rows == 0 or peers.buf.dtype != x.dtype or out.numel() != x.numel():
        raise ValueError("tp_add_rms_norm_fuse_sum_fp16: input does not match the peer partial buffers")
    lib = _lib.load()
    if tp_two_step(peers, rows):
        _lib.check(lib.qspec_tp_rs_add_norm_f16(_ptr(x), peers._partials[which], peers._flags, peers._reduced[which],
                                                peers._row_flags, _ptr(peers.state), peers.world, peers.rank, phases,
                                                _ptr(out), rows, K, float(eps), _dt(x), _stream()),
                   "tp_add_rms_norm_fuse_sum_fp16 (two-step)")
        return
    _lib.check(lib.qspec_tp_add_norm_f16(_ptr(x), peers._partials[which], peers._flags, _ptr(peers.state), peers.world,
                                         peers.rank, _ptr(out), rows, K, float(eps), _dt(x), _stream()),
               "tp_add_rms_norm_fuse_sum_fp16")


def rms_norm_general_fuse_sum_fp16(out: torch.Tensor, x: torch.Tensor, eps: float) -> None:
    _chk(x, "input", (_F16, _BF16)); _chk(out, "out", (x.dtype,))
    K = x.shape[-1]
    rows = x.numel() // K if K else 0
    if out.numel() != x.numel():
        raise ValueError("rms_norm_general_fuse_sum_fp16: out must have the input's shape")
    if rows == 0:
        return
    lib = _lib.load()
    _lib.check(lib.qspec_norm_f16(_ptr(x), _ptr(out), rows, K, float(eps), _dt(x), _stream()),
               "rms_norm_general_fuse_sum_fp16")


# ---------------------------------------------------------------------------------------------
# quantiser
# ---------------------------------------------------------------------------------------------

def fuse_sym_quant(x: torch.Tensor, scale: torch.Tensor, q: torch.Tensor, input_clip_ratio: float = 1.0) -> None:
    _chk(x, "x", (_F16,)); _chk(scale, "scale", (_F16,)); _chk(q, "q", _I4)
    if x.dim() != 2:
        raise ValueError("fuse_sym_quant: x must be 2-D")
    rows, K = x.shape
    if q.numel() != rows * K // 2 or scale.numel() < rows:
        raise ValueError("fuse_sym_quant: output shapes do not match the input")
    if rows == 0:
        return
    lib = _lib.load()
    _lib.check(lib.qspec_sym_quant_i4(_ptr(x), _ptr(scale), _ptr(q), rows, K, float(input_clip_ratio),
                                      _stream()), "fuse_sym_quant")


# ---------------------------------------------------------------------------------------------
# (B) Hadamard
# ---------------------------------------------------------------------------------------------

def faster_fast_hadamard_transform(x: torch.Tensor, scale: float, out: torch.Tensor) -> torch.Tensor:
    _chk(x, "x", (_F16, _BF16)); _chk(out, "out", (x.dtype,))
    if x.dim() != 2:
        raise ValueError("faster_fast_hadamard_transform: x must be [batch, dim]")
    rows, n = x.shape
    if out.numel() != x.numel():
        raise ValueError("faster_fast_hadamard_transform: out must have x's shape")
    if rows == 0:
        return out
    lib = _lib.load()
    _lib.check(lib.qspec_fht(_ptr(x), _ptr(out), rows, n, float(scale), _dt(x), _stream()),
               "faster_fast_hadamard_transform")
    return out


def _check_container(out_q: Optional[torch.Tensor], rows: int, K: int, c8: bool, who: str) -> None:
    """out_q holds rows x K int4 values: K/2 bytes per row packed, K bytes per row in the int8 container."""
    if out_q is not None and out_q.numel() != rows * (K if c8 else K // 2):
        raise ValueError(f"{who}: out_q has {out_q.numel()} bytes, expected {rows * (K if c8 else K // 2)} "
                         f"({'int8 container' if c8 else 'packed int4'})")


def hadamard_heads(x: torch.Tensor, num_heads: int, head_dim: int, scale: float,
                   out: Optional[torch.Tensor] = None, out_q: Optional[torch.Tensor] = None,
                   scale_out: Optional[torch.Tensor] = None, clip_ratio: float = 1.0, c8: bool = False) -> None:
    """Fused o_proj rotation (+quant): quarot_llama.py:230-238 in one launch. c8: out_q is the int8 container
    ([rows, H*D] int8, one int4 value per byte) instead of the packed format."""
    _chk(x, "x", (_F16,)); _chk(out, "out", (_F16,), True); _chk(out_q, "out_q", _I4, True)
    _chk(scale_out, "scale_out", (_F16,), True)
    rows = x.numel() // (num_heads * head_dim)
    lib = _lib.load()
    _check_container(out_q, rows, num_heads * head_dim, c8, "hadamard_heads")
    _lib.check((lib.qspec_hadamard_heads_c8 if c8 else lib.qspec_hadamard_heads)(_ptr(x), _ptr(out), _ptr(out_q), _ptr(scale_out), rows, num_heads,
                                        head_dim, float(scale), float(clip_ratio), _stream()),
               "hadamard_heads")


def hadamard_k28(x: Optional[torch.Tensor], n: int, scale: float, gate_up: Optional[torch.Tensor] = None,
                 out: Optional[torch.Tensor] = None, out_q: Optional[torch.Tensor] = None,
                 scale_out: Optional[torch.Tensor] = None, clip_ratio: float = 1.0, c8: bool = False) -> None:
    """Fused down_proj rotation had28 (x) H_n (+SiLU*up in front, +quant behind). c8: out_q is the int8 container."""
    _chk(x, "x", (_F16,), True); _chk(gate_up, "gate_up", (_F16,), True)
    _chk(out, "out", (_F16,), True); _chk(out_q, "out_q", _I4, True); _chk(scale_out, "scale_out", (_F16,), True)
    src = gate_up if gate_up is not None else x
    if src is None:
        raise ValueError("hadamard_k28: x or gate_up is required")
    width = 28 * n * (2 if gate_up is not None else 1)
    rows = src.numel() // width
    lib = _lib.load()
    _check_container(out_q, rows, 28 * n, c8, "hadamard_k28")
    _lib.check((lib.qspec_hadamard_k28_c8 if c8 else lib.qspec_hadamard_k28)(_ptr(x) if gate_up is None else None, _ptr(gate_up), _ptr(out), _ptr(out_q),
                                      _ptr(scale_out), rows, n, float(scale), float(clip_ratio), _stream()),
               "hadamard_k28")


def hadamard_k108(x: Optional[torch.Tensor], n: int, scale: float, gate_up: Optional[torch.Tensor] = None,
                  gate_first: bool = True, out: Optional[torch.Tensor] = None, out_q: Optional[torch.Tensor] = None,
                  scale_out: Optional[torch.Tensor] = None, clip_ratio: float = 1.0) -> None:
    """Qwen2.5-14B down_proj rotation had108 (x) H_128 (+SiLU*mul in front, +quant behind): hadamard_utils.matmul_hadU_cuda
    with K = 108 as Qwen2MLP.forward calls it (qwen2_quarot.py:119-127). gate_up: [gate ; up] (vLLM SiluAndMul) when
    gate_first, else [up ; gate]."""
    _chk(x, "x", (_F16,), True); _chk(gate_up, "gate_up", (_F16,), True)
    _chk(out, "out", (_F16,), True); _chk(out_q, "out_q", _I4, True); _chk(scale_out, "scale_out", (_F16,), True)
    src = gate_up if gate_up is not None else x
    if src is None:
        raise ValueError("hadamard_k108: x or gate_up is required")
    rows = src.numel() // (108 * n * (2 if gate_up is not None else 1))
    lib = _lib.load()
    _lib.check(lib.qspec_hadamard_k108(_ptr(x) if gate_up is None else None, _ptr(gate_up), 1 if gate_first else 0, _ptr(out),
                                       _ptr(out_q), _ptr(scale_out), rows, n, float(scale), float(clip_ratio), _stream()),
               "hadamard_k108")


def hadamard_heads40(x: torch.Tensor, num_heads: int, head_dim: int, scale: float, out: Optional[torch.Tensor] = None,
                     out_q: Optional[torch.Tensor] = None, scale_out: Optional[torch.Tensor] = None,
                     clip_ratio: float = 1.0) -> None:
    """Qwen2.5-14B o_proj rotation (Qwen2Attention.o_had_trans with o_K = 40, qwen2_quarot.py:269-277) (+quant)."""
    _chk(x, "x", (_F16,)); _chk(out, "out", (_F16,), True); _chk(out_q, "out_q", _I4, True)
    _chk(scale_out, "scale_out", (_F16,), True)
    rows = x.numel() // (num_heads * head_dim)
    lib = _lib.load()
    _lib.check(lib.qspec_hadamard_heads40(_ptr(x), _ptr(out), _ptr(out_q), _ptr(scale_out), rows, num_heads, head_dim,
                                          float(scale), float(clip_ratio), _stream()), "hadamard_heads40")


# ---------------------------------------------------------------------------------------------
# GEMMs
# ---------------------------------------------------------------------------------------------

def rowwise_scaled_linear_cutlass_s4s4_unified(input: torch.Tensor, input_scale: torch.Tensor,
                                               weight: torch.Tensor, weight_scale: torch.Tensor,
                                               bias: Optional[torch.Tensor], result: torch.Tensor) -> torch.Tensor:
    """W4A4: result = T(((acc * input_scale[m]) * weight_scale[n]) + bias[n]); exact int32 acc."""
    _chk(input, "input", _I4); _chk(weight, "weight", _I4)
    _chk(result, "result", (_F16, _BF16))
    _chk(input_scale, "input_scale", (result.dtype,)); _chk(weight_scale, "weight_scale", (result.dtype,))
    _chk(bias, "bias", (result.dtype,), True)
    if weight.dim() != 2:
        raise ValueError("expected wq argument to be 2D tensor")
    N, K2 = weight.shape
    if input.shape[-1] != K2:
        raise ValueError(f"expected xq argument to have {K2} columns, but got {input.shape[-1]}")
    M = input.numel() // K2 if K2 else 0
    if input_scale.numel() != M or weight_scale.numel() != N or result.numel() != M * N:
        raise ValueError("rowwise_scaled_linear_cutlass_s4s4_unified: scale/result sizes do not match")
    if bias is not None and bias.numel() != N:
        raise ValueError(f"expected bias argument to have {N} elements")
    if M == 0:
        return result
    lib = _lib.load()
    K = 2 * K2
    need = lib.qspec_gemm_w4a4_workspace_bytes(M, N, K)
    ws = _workspace("w4a4", need, input.device)
    _lib.check(lib.qspec_gemm_w4a4(_ptr(input), _ptr(input_scale), _ptr(weight), _ptr(weight_scale), _ptr(bias),
                                   _ptr(result), M, N, K, _dt(result), _ptr(ws), ws.numel(), _stream()),
               "rowwise_scaled_linear_cutlass_s4s4_unified")
    return result


def gemm_w4a4_acc_i32(input: torch.Tensor, weight: torch.Tensor, acc: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Parity/debug entry: raw int32 accumulators of the tcgen05 main loop."""
    _chk(input, "input", _I4); _chk(weight, "weight", _I4)
    N, K2 = weight.shape
    M = input.numel() // K2
    if acc is None:
        acc = torch.empty(M, N, dtype=torch.int32, device=input.device)
    _chk(acc, "acc", (torch.int32,))
    lib = _lib.load()
    need = lib.qspec_gemm_w4a4_workspace_bytes(M, N, 2 * K2)
    ws = _workspace("w4a4", need, input.device)
    _lib.check(lib.qspec_gemm_w4a4_acc_i32(_ptr(input), _ptr(weight), _ptr(acc), M, N, 2 * K2, _ptr(ws),
                                           ws.numel(), _stream()), "gemm_w4a4_acc_i32")
    return acc


def gemm_w4a16(x: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor,
               bias: Optional[torch.Tensor], out: torch.Tensor) -> torch.Tensor:
    """W4A16: out = x @ (int4(weight) * weight_scale[:, None]).T (+ bias); weight is the SAME
    two's-complement packed tensor the W4A4 path reads (no `^ 0x88` copy)."""
    _chk(x, "x", (_F16, _BF16)); _chk(weight, "weight", _I4); _chk(out, "out", (x.dtype,))
    _chk(weight_scale, "weight_scale", (x.dtype,)); _chk(bias, "bias", (x.dtype,), True)
    N, K2 = weight.shape
    K = 2 * K2
    if x.shape[-1] != K:
        raise ValueError(f"gemm_w4a16: x has {x.shape[-1]} columns, weight expects {K}")
    M = x.numel() // K
    if weight_scale.numel() != N or out.numel() != M * N:
        raise ValueError("gemm_w4a16: scale/out sizes do not match")
    if M == 0:
        return out
    lib = _lib.load()
    need = lib.qspec_gemm_w4a16_workspace_bytes(M, N, K)
    ws = _workspace("w4a16", need, x.device)
    _lib.check(lib.qspec_gemm_w4a16(_ptr(x), _ptr(weight), _ptr(weight_scale), _ptr(bias), _ptr(out), M, N, K,
                                    _dt(x), _ptr(ws), ws.numel(), _stream()), "gemm_w4a16")
    return out


# ---------------------------------------------------------------------------------------------
# W4A8 mode (kwargs["w4a8"], off by default in the reference)
#   layernorm_ops.rms_norm_general_fuse_sum_i8   third-party/kernels/csrc/layernorm.cpp:44-50
#   quarot._CUDA.fuse_sym_quant_i8               third-party/QuaRot/quarot/kernels/bindings.cpp:148-164
#   rowwise_scaled_linear_cutlass_s8s4           torchao, as called at quarot_nn/linear.py:98
# ---------------------------------------------------------------------------------------------

def rms_norm_general_fuse_sum_i8(out_q: torch.Tensor, x: torch.Tensor, input_sum: Optional[torch.Tensor],
                                 scaling: torch.Tensor, eps: float, use_per_token_quant: bool = True) -> None:
    if not use_per_token_quant:
        raise NotImplementedError("only the per-token branch is on the QSpec path (normalization.py:96-103)")
    _chk(x, "input", (_F16,)); _chk(out_q, "out", (torch.int8,)); _chk(scaling, "scaling", (_F16,))
    _chk(input_sum, "input_sum", (_F16,), allow_none=True)
    K = x.shape[-1]
    rows = x.numel() // K if K else 0
    if out_q.numel() != rows * K or scaling.numel() < rows:
        raise ValueError("rms_norm_general_fuse_sum_i8: output shapes do not match the input")
    if rows == 0:
        return
    lib = _lib.load()
    _lib.check(lib.qspec_norm_quant_i8(_ptr(x), _ptr(out_q), _ptr(scaling), _ptr(input_sum), rows, K,
                                       float(eps), _stream()), "rms_norm_general_fuse_sum_i8")


def fuse_sym_quant_i8(x: torch.Tensor, scale: torch.Tensor, q: torch.Tensor, input_clip_ratio: float = 1.0) -> None:
    _chk(x, "x", (_F16,)); _chk(scale, "scale", (_F16,)); _chk(q, "q", (torch.int8,))
    if x.dim() != 2:
        raise ValueError("fuse_sym_quant_i8: x must be 2-D")
    rows, K = x.shape
    if q.numel() != rows * K or scale.numel() < rows:
        raise ValueError("fuse_sym_quant_i8: output shapes do not match the input")
    if rows == 0:
        return
    lib = _lib.load()
    _lib.check(lib.qspec_sym_quant_i8(_ptr(x), _ptr(scale), _ptr(q), rows, K, float(input_clip_ratio),
                                      _stream()), "fuse_sym_quant_i8")


def rowwise_scaled_linear_cutlass_s8s4(input: torch.Tensor, input_scale: torch.Tensor, weight: torch.Tensor,
                                       weight_scale: torch.Tensor, bias: Optional[torch.Tensor] = None,
                                       result: Optional[torch.Tensor] = None) -> torch.Tensor:
    """W4A8: result = T(((acc * input_scale[m]) * weight_scale[n]) + bias[n]); input int8 [M, K], weight packed int4 [N, K/2]."""
    _chk(input, "input", (torch.int8,)); _chk(weight, "weight", _I4)
    _chk(input_scale, "input_scale", (_F16, _BF16)); _chk(weight_scale, "weight_scale", (input_scale.dtype,))
    _chk(bias, "bias", (input_scale.dtype,), True)
    if weight.dim() != 2:
        raise ValueError("expected wq argument to be 2D tensor")
    N, K2 = weight.shape
    K = 2 * K2
    if input.shape[-1] != K:
        raise ValueError(f"expected xq argument to have {K} columns, but got {input.shape[-1]}")
    M = input.numel() // K if K else 0
    if result is None:
        result = torch.empty(*input.shape[:-1], N, dtype=input_scale.dtype, device=input.device)
    _chk(result, "result", (input_scale.dtype,))
    if input_scale.numel() != M or weight_scale.numel() != N or result.numel() != M * N:
        raise ValueError("rowwise_scaled_linear_cutlass_s8s4: scale/result sizes do not match")
    if bias is not None and bias.numel() != N:
        raise ValueError(f"expected bias argument to have {N} elements")
    if M == 0:
        return result
    lib = _lib.load()
    need = lib.qspec_gemm_w4a8_workspace_bytes(M, N, K)
    ws = _workspace("w4a4", need, input.device)
    _lib.check(lib.qspec_gemm_w4a8(_ptr(input), _ptr(input_scale), _ptr(weight), _ptr(weight_scale), _ptr(bias),
                                   _ptr(result), M, N, K, _dt(result), _ptr(ws), ws.numel(), _stream()),
               "rowwise_scaled_linear_cutlass_s8s4")
    return result


def gemm_w4a8_acc_i32(input: torch.Tensor, weight: torch.Tensor, acc: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Parity/debug entry: raw int32 accumulators sum_k a_i8[m,k] * w_i4[n,k]."""
    _chk(input, "input", (torch.int8,)); _chk(weight, "weight", _I4)
    N, K2 = weight.shape
    K = 2 * K2
    M = input.numel() // K
    if acc is None:
        acc = torch.empty(M, N, dtype=torch.int32, device=input.device)
    _chk(acc, "acc", (torch.int32,))
    lib = _lib.load()
    need = lib.qspec_gemm_w4a8_workspace_bytes(M, N, K)
    ws = _workspace("w4a4", need, input.device)
    _lib.check(lib.qspec_gemm_w4a8_acc_i32(_ptr(input), _ptr(weight), _ptr(acc), M, N, K, _ptr(ws), ws.numel(),
                                           _stream()), "gemm_w4a8_acc_i32")
    return acc


# ---------------------------------------------------------------------------------------------
# deferred reduction: GEMM leaves per-k-slice partial accumulators, the consuming norm sums them
# ---------------------------------------------------------------------------------------------

class SplitPartials:
    """Raw partial accumulators of one o_proj / down_proj call (qspec_gemm_w4a4_split / qspec_gemm_w4a16_split) plus
    what the consumer needs to finish the GEMM epilogue: [split][M][N] 32-bit words, activation / weight scales, bias."""
    __slots__ = ("buf", "split", "w4a16", "sA", "sW", "bias", "M", "N")

    def __init__(self, buf, split, w4a16, sA, sW, bias, M, N):
        self.buf, self.split, self.w4a16, self.sA, self.sW, self.bias, self.M, self.N = buf, split, w4a16, sA, sW, bias, M, N


def gemm_split_factor(w4a16, M: int, N: int, K: int) -> int:
    """Slices per output element the deferred-reduction GEMM leaves for this shape; 0 = use the plain GEMM.
    w4a16: False / 0 = W4A4 (packed activations), True / 1 = W4A16, 2 = W4A4 from int8-container activations."""
    return int(_lib.load().qspec_gemm_split_factor(int(w4a16), M, N, K))


def gemm_w4a4_c8_split(input: torch.Tensor, input_scale: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor,
                       bias: Optional[torch.Tensor], partials: torch.Tensor) -> SplitPartials:
    """gemm_w4a4_split for int8-container activations ([M, K] int8, one int4 value per byte)."""
    _chk(input, "input", _I4); _chk(weight, "weight", _I4); _chk(partials, "partials", (torch.int32, torch.float32))
    N, K2 = weight.shape
    K = 2 * K2
    M = input.numel() // K
    lib = _lib.load()
    split = lib.qspec_gemm_split_factor(2, M, N, K)
    if split < 1 or partials.numel() < split * M * N:
        raise ValueError(f"gemm_w4a4_c8_split: unsupported shape or partial buffer too small (split={split})")
    ws = _workspace("w4a8", lib.qspec_gemm_w4a8_workspace_bytes(M, N, K), input.device)
    _lib.check(lib.qspec_gemm_w4a8_split(_ptr(input), _ptr(weight), _ptr(partials), partials.numel() * 4, M, N, K, _ptr(ws),
                                         ws.numel(), _stream()), "gemm_w4a4_c8_split")
    return SplitPartials(partials, split, 2, input_scale, weight_scale, bias, M, N)


def gemm_w4a4_split(input: torch.Tensor, input_scale: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor,
                    bias: Optional[torch.Tensor], partials: torch.Tensor) -> SplitPartials:
    """rowwise_scaled_linear_cutlass_s4s4_unified without its cross-CTA reduction and epilogue: raw int32 partial sums
    per k-slice go to `partials` (int32, >= split*M*N elements); the scales travel with the result to the consumer."""
    _chk(input, "input", _I4); _chk(weight, "weight", _I4); _chk(partials, "partials", (torch.int32, torch.float32))
    N, K2 = weight.shape
    M = input.numel() // K2
    K = 2 * K2
    lib = _lib.load()
    split = lib.qspec_gemm_split_factor(0, M, N, K)
    if split < 1 or partials.numel() < split * M * N:
        raise ValueError(f"gemm_w4a4_split: unsupported shape or partial buffer too small (split={split})")
    ws = _workspace("w4a4", lib.qspec_gemm_w4a4_workspace_bytes(M, N, K), input.device)
    _lib.check(lib.qspec_gemm_w4a4_split(_ptr(input), _ptr(weight), _ptr(partials), partials.numel() * 4, M, N, K, _ptr(ws),
                                         ws.numel(), _stream()), "gemm_w4a4_split")
    return SplitPartials(partials, split, False, input_scale, weight_scale, bias, M, N)


def gemm_w4a16_split(x: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor, bias: Optional[torch.Tensor],
                     partials: torch.Tensor) -> SplitPartials:
    _chk(x, "x", (_F16, _BF16)); _chk(weight, "weight", _I4); _chk(partials, "partials", (torch.int32, torch.float32))
    N, K2 = weight.shape
    K = 2 * K2
    M = x.numel() // K
    lib = _lib.load()
    split = lib.qspec_gemm_split_factor(1, M, N, K)
    if split < 1 or partials.numel() < split * M * N:
        raise ValueError(f"gemm_w4a16_split: unsupported shape or partial buffer too small (split={split})")
    ws = _workspace("w4a16", lib.qspec_gemm_w4a16_workspace_bytes(M, N, K), x.device)
    _lib.check(lib.qspec_gemm_w4a16_split(_ptr(x), _ptr(weight), _ptr(partials), partials.numel() * 4, M, N, K, _dt(x),
                                          _ptr(ws), ws.numel(), _stream()), "gemm_w4a16_split")
    return SplitPartials(partials, split, True, None, weight_scale, bias, M, N)


def split_add_rms_norm_fuse_sum_i4(out_q: torch.Tensor, x: torch.Tensor, sp: SplitPartials, input_sum: torch.Tensor,
                                   scaling: torch.Tensor, eps: float, c8: bool = False) -> None:
    """x += epilogue(sum of sp's slices) (in place), then rms_norm_general_fuse_sum_i4: one launch; bit-identical to the
    plain GEMM followed by add_rms_norm_fuse_sum_i4."""
    _chk(x, "input", (_F16,)); _chk(out_q, "out", _I4); _chk(scaling, "scaling", (_F16,)); _chk(input_sum, "input_sum", (_F16,))
    K = x.shape[-1]
    rows = x.numel() // K if K else 0
    if rows != sp.M or K != sp.N:
        raise ValueError("split_add_rms_norm_fuse_sum_i4: input does not match the partial sums")
    lib = _lib.load()
    _check_container(out_q, rows, K, c8, "split_add_rms_norm_fuse_sum_i4")
    _lib.check((lib.qspec_split_add_norm_quant_i4_c8 if c8 else lib.qspec_split_add_norm_quant_i4)(_ptr(x), _ptr(sp.buf), sp.split, int(sp.w4a16), _ptr(sp.sA), _ptr(sp.sW),
                                                 _ptr(sp.bias), _ptr(out_q), _ptr(scaling), _ptr(input_sum), rows, K, float(eps),
                                                 _stream()), "split_add_rms_norm_fuse_sum_i4")


def split_add_rms_norm_fuse_sum_fp16(out: torch.Tensor, x: torch.Tensor, sp: SplitPartials, eps: float) -> None:
    _chk(x, "input", (_F16, _BF16)); _chk(out, "out", (x.dtype,))
    K = x.shape[-1]
    rows = x.numel() // K if K else 0
    if rows != sp.M or K != sp.N or out.numel() != x.numel():
        raise ValueError("split_add_rms_norm_fuse_sum_fp16: input does not match the partial sums")
    lib = _lib.load()
    _lib.check(lib.qspec_split_add_norm_f16(_ptr(x), _ptr(sp.buf), sp.split, int(sp.w4a16), _ptr(sp.sA), _ptr(sp.sW),
                                            _ptr(sp.bias), _ptr(out), rows, K, float(eps), _dt(x), _stream()),
               "split_add_rms_norm_fuse_sum_fp16")


# ---------------------------------------------------------------------------------------------
# neighbours
# ---------------------------------------------------------------------------------------------

def add(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    _chk(a, "a", (_F16, _BF16)); _chk(b, "b", (a.dtype,)); _chk(out, "out", (a.dtype,))
    lib = _lib.load()
    _lib.check(lib.qspec_add(_ptr(a), _ptr(b), _ptr(out), a.numel(), _dt(a), _stream()), "add")
    return out


def silu_mul(gate_up: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """gate_up = [up ; gate] (quarot_llama.py:279); out = silu(gate) * up."""
    _chk(gate_up, "gate_up", (_F16, _BF16)); _chk(out, "out", (gate_up.dtype,))
    inter = gate_up.shape[-1] // 2
    rows = gate_up.numel() // (2 * inter)
    lib = _lib.load()
    _lib.check(lib.qspec_silu_mul(_ptr(gate_up), _ptr(out), rows, inter, _dt(gate_up), _stream()), "silu_mul")
    return out


def rope_kvcache(qkv: torch.Tensor, cos_sin_cache: torch.Tensor, positions: torch.Tensor, seq_ids: torch.Tensor,
                 cache_k: torch.Tensor, cache_v: torch.Tensor, num_heads: int, num_kv_heads: int, head_dim: int) -> None:
    """In-place neox RoPE on q,k of qkv [T, (Hq+2Hkv)*D]; appends k,v to cache [B, S, Hkv, D]."""
    _chk(qkv, "qkv", (_F16,)); _chk(cos_sin_cache, "cos_sin_cache", (_F16,))
    _chk(positions, "positions", (torch.int32,)); _chk(seq_ids, "seq_ids", (torch.int32,))
    _chk(cache_k, "cache_k", (_F16,)); _chk(cache_v, "cache_v", (_F16,))
    T = positions.numel()
    if cos_sin_cache.shape[0] < cache_k.shape[1]:
        raise ValueError(f"rope_kvcache: rotary table has {cos_sin_cache.shape[0]} positions, the KV cache {cache_k.shape[1]}")
    lib = _lib.load()
    _lib.check(lib.qspec_rope_kvcache(_ptr(qkv), _ptr(cos_sin_cache), _ptr(positions), _ptr(seq_ids), _ptr(cache_k),
                                      _ptr(cache_v), T, num_heads, num_kv_heads, head_dim, cache_k.shape[1], _stream()),
               "rope_kvcache")


def rope_qkhad_kvcache(qkv: torch.Tensor, cos_sin_cache: torch.Tensor, positions: torch.Tensor, seq_ids: torch.Tensor,
                       cache_k: torch.Tensor, cache_v: torch.Tensor, num_heads: int, num_kv_heads: int, head_dim: int,
                       had_scale: float) -> None:
    """rope_kvcache + the per-head Hadamard of q and k after the rotary embedding (Qwen2Attention.qk_had_trans,
    qwen2_quarot.py:279-288)."""
    _chk(qkv, "qkv", (_F16,)); _chk(cos_sin_cache, "cos_sin_cache", (_F16,))
    _chk(positions, "positions", (torch.int32,)); _chk(seq_ids, "seq_ids", (torch.int32,))
    _chk(cache_k, "cache_k", (_F16,)); _chk(cache_v, "cache_v", (_F16,))
    if cos_sin_cache.shape[0] < cache_k.shape[1]:
        raise ValueError(f"rope_qkhad_kvcache: rotary table has {cos_sin_cache.shape[0]} positions, the KV cache {cache_k.shape[1]}")
    lib = _lib.load()
    _lib.check(lib.qspec_rope_qkhad_kvcache(_ptr(qkv), _ptr(cos_sin_cache), _ptr(positions), _ptr(seq_ids), _ptr(cache_k),
                                            _ptr(cache_v), positions.numel(), num_heads, num_kv_heads, head_dim,
                                            cache_k.shape[1], float(had_scale), _stream()), "rope_qkhad_kvcache")


def decode_attention(qkv: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor, positions: torch.Tensor,
                     seq_ids: torch.Tensor, out: torch.Tensor, num_heads: int, num_kv_heads: int, head_dim: int,
                     sm_scale: float) -> torch.Tensor:
    _chk(qkv, "qkv", (_F16,)); _chk(out, "out", (_F16,))
    T = positions.numel()
    lib = _lib.load()
    _lib.check(lib.qspec_decode_attention(_ptr(qkv), _ptr(cache_k), _ptr(cache_v), _ptr(positions), _ptr(seq_ids),
                                          _ptr(out), T, num_heads, num_kv_heads, head_dim, cache_k.shape[1],
                                          float(sm_scale), _stream()), "decode_attention")
    return out


def seq_attention_supported(T: int, num_heads: int, num_kv_heads: int, head_dim: int, cache_len: int) -> bool:
    return bool(_lib.load().qspec_seq_attention_supported(T, num_heads, num_kv_heads, head_dim, cache_len))


def seq_attention(qkv: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor, positions: torch.Tensor,
                  seq_ids: torch.Tensor, out: torch.Tensor, num_heads: int, num_kv_heads: int, head_dim: int,
                  sm_scale: float) -> torch.Tensor:
    """decode_attention for passes with many sequences / several tokens per sequence: keys and values of a (sequence, kv head)
    are read once for all its tokens and query heads."""
    _chk(qkv, "qkv", (_F16,)); _chk(out, "out", (_F16,)); _chk(cache_k, "cache_k", (_F16,)); _chk(cache_v, "cache_v", (_F16,))
    _chk(positions, "positions", (torch.int32,)); _chk(seq_ids, "seq_ids", (torch.int32,))
    lib = _lib.load()
    _lib.check(lib.qspec_seq_attention(_ptr(qkv), _ptr(cache_k), _ptr(cache_v), _ptr(positions), _ptr(seq_ids), _ptr(out),
                                       positions.numel(), cache_k.shape[0], num_heads, num_kv_heads, head_dim, cache_k.shape[1],
                                       float(sm_scale), _stream()), "seq_attention")
    return out


def attn_block_supported(T: int, num_heads: int, num_kv_heads: int, head_dim: int, cache_len: int) -> bool:
    return bool(_lib.load().qspec_attn_block_supported(T, num_heads, num_kv_heads, head_dim, cache_len))


def attn_block(qkv: torch.Tensor, cos_sin_cache: torch.Tensor, positions: torch.Tensor, seq_ids: torch.Tensor,
               cache_k: torch.Tensor, cache_v: torch.Tensor, num_heads: int, num_kv_heads: int, head_dim: int,
               sm_scale: float, had_scale: float, out: Optional[torch.Tensor] = None, out_q: Optional[torch.Tensor] = None,
               scale_out: Optional[torch.Tensor] = None, attn_out: Optional[torch.Tensor] = None,
               input_clip_ratio: float = 1.0) -> None:
    """rope_kvcache + decode_attention + hadamard_heads (+ quantiser) in one launch (quarot_llama.py:207-238); qkv is
    left untouched. Bit-identical to the chain of the three calls."""
    _chk(qkv, "qkv", (_F16,)); _chk(out, "out", (_F16,), allow_none=True); _chk(out_q, "out_q", _I4, allow_none=True)
    _chk(scale_out, "scale_out", (_F16,), allow_none=True); _chk(attn_out, "attn_out", (_F16,), allow_none=True)
    if positions.dtype != torch.int32 or seq_ids.dtype != torch.int32:
        raise ValueError("attn_block: positions and seq_ids must be int32")
    T = positions.numel()
    if cos_sin_cache.shape[0] < cache_k.shape[1]:
        raise ValueError(f"attn_block: rotary table has {cos_sin_cache.shape[0]} positions, the KV cache {cache_k.shape[1]}")
    lib = _lib.load()
    _lib.check(lib.qspec_attn_block(_ptr(qkv), _ptr(cos_sin_cache), _ptr(positions), _ptr(seq_ids), _ptr(cache_k),
                                    _ptr(cache_v), _ptr(attn_out), _ptr(out), _ptr(out_q), _ptr(scale_out), T, num_heads,
                                    num_kv_heads, head_dim, cache_k.shape[1], float(sm_scale), float(had_scale),
                                    float(input_clip_ratio), _stream()), "attn_block")


def argmax_rows(logits: torch.Tensor, out_val: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Greedy token per row of the f16 logits [rows, V]: int64 indices (lowest index among equal maxima, like
    torch.argmax on distinct values); out_val (f32 [rows], optional) receives the maxima."""
    if not logits.is_cuda or logits.dtype != _F16:
        raise ValueError("argmax_rows: logits must be a float16 CUDA tensor (qspec_b200 has no CPU path)")
    if logits.dim() != 2 or logits.stride(1) != 1:
        raise ValueError("argmax_rows: logits must be [rows, V] with unit column stride")
    rows, V = logits.shape
    out = torch.empty(rows, dtype=torch.int64, device=logits.device)
    if rows == 0:
        return out
    lib = _lib.load()
    ws = _workspace("argmax", lib.qspec_argmax_rows_workspace_bytes(rows), logits.device)
    _lib.check(lib.qspec_argmax_rows(_ptr(logits), rows, V, logits.stride(0), _ptr(out), _ptr(out_val), _ptr(ws), ws.numel(),
                                     _stream()), "argmax_rows")
    return out


def tp_argmax_combine(local_max: torch.Tensor, local_idx: torch.Tensor, index_offset: int, peers,
                      out: Optional[torch.Tensor] = None, phases: int = 3) -> torch.Tensor:
    """Greedy token over a vocab-sharded lm_head through peer memory (no collective call): every rank pushes its
    (max logit, global index) pairs into every peer's table and takes the best of its own in rank order -- what the
    all-gather of the sharded logits (layers/logits_processor.py:84-107) followed by an arg max gives. int64 [rows],
    identical on every rank. phases 1 / 2: the two halves as separate launches (tests/tp_emulation.py)."""
    if not local_max.is_cuda or local_max.dtype != torch.float32 or local_idx.dtype != torch.int64:
        raise ValueError("tp_argmax_combine: local_max must be float32 and local_idx int64 CUDA tensors")
    rows = local_max.numel()
    if rows == 0 or rows != local_idx.numel() or rows > peers.max_tokens:
        raise ValueError("tp_argmax_combine: rows must be 1..peers.max_tokens on every rank")
    if out is None:
        out = torch.empty(rows, dtype=torch.int64, device=local_max.device)
    lib = _lib.load()
    _lib.check(lib.qspec_tp_argmax_combine(_ptr(local_max), _ptr(local_idx), int(index_offset), rows, peers._amax_slots,
                                           peers._amax_flags, _ptr(peers.amax_state), peers.max_tokens, peers.world,
                                           peers.rank, phases, _ptr(out), _stream()), "tp_argmax_combine")
    return out


def launch_count() -> int:
    return int(_lib.load().qspec_launch_count())
