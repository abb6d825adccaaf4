"""ctypes binding of the C-ABI library (include/qspec_b200.h).

There is NO fallback: if the shared library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_size_t, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("QSPEC_LIB") or os.path.join(_HERE, "_C", "libqspec_b200.so")

QSPEC_OK = 0
DTYPE_F16 = 0
DTYPE_BF16 = 1

# name -> (restype, argtypes); mirrors include/qspec_b200.h one to one
SIGNATURES = {
    "qspec_abi_version": (c_int, []),
    "qspec_last_error": (c_char_p, []),
    "qspec_launch_count": (c_uint64, []),
    "qspec_sm_count": (c_int, []),
    "qspec_debug_set_gemm_trace": (None, [c_void_p]),
    "qspec_norm_quant_i4": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_norm_f16": (c_int, [c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p]),
    "qspec_add_norm_quant_i4": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_add_norm_f16": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p]),
    "qspec_gemm_split_factor": (c_int, [c_int, c_int, c_int, c_int]),
    "qspec_gemm_w4a4_split": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int, c_void_p, c_size_t,
                                      c_void_p]),
    "qspec_gemm_w4a16_split": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int, c_int, c_void_p,
                                       c_size_t, c_void_p]),
    "qspec_split_add_norm_quant_i4": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                              c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_split_add_norm_f16": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_int, c_int, c_float, c_int, c_void_p]),
    "qspec_tp_add_norm_quant_i4": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                           c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_tp_add_norm_f16": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_int,
                                      c_float, c_int, c_void_p]),
    "qspec_tp_rs_add_norm_quant_i4": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                              c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_tp_rs_add_norm_f16": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                         c_void_p, c_int, c_int, c_float, c_int, c_void_p]),
    "qspec_split_add_norm_quant_i4_c8": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                              c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_tp_add_norm_quant_i4_c8": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                           c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_hadamard_heads_c8": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                     c_float, c_float, c_void_p]),
    "qspec_hadamard_k28_c8": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                   c_float, c_float, c_void_p]),
    "qspec_gemm_w4a8_split": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int, c_void_p, c_size_t,
                                      c_void_p]),
    "qspec_debug_gemm_plan": (c_int, [c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "qspec_tp_argmax_combine": (c_int, [c_void_p, c_void_p, ctypes.c_longlong, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                        c_int, c_int, c_void_p, c_void_p]),
    "qspec_sym_quant_i4": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_norm_quant_i8": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_sym_quant_i8": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]),
    "qspec_gemm_w4a8_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "qspec_gemm_w4a8": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                c_int, c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "qspec_gemm_w4a8_acc_i32": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p,
                                        c_size_t, c_void_p]),
    "qspec_fht": (c_int, [c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p]),
    "qspec_hadamard_heads": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                     c_float, c_float, c_void_p]),
    "qspec_hadamard_k28": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                   c_float, c_float, c_void_p]),
    "qspec_gemm_w4a4_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "qspec_gemm_w4a4": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                c_int, c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "qspec_gemm_w4a4_acc_i32": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p,
                                        c_size_t, c_void_p]),
    "qspec_gemm_w4a16_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "qspec_gemm_w4a16": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                 c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "qspec_add": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_void_p]),
    "qspec_silu_mul": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "qspec_rope_kvcache": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                   c_int, c_int, c_int, c_void_p]),
    "qspec_decode_attention": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                       c_int, c_int, c_int, c_int, c_float, c_void_p]),
    "qspec_seq_attention_supported": (c_int, [c_int, c_int, c_int, c_int, c_int]),
    "qspec_seq_attention": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                    c_int, c_int, c_float, c_void_p]),
    "qspec_argmax_rows_workspace_bytes": (ctypes.c_size_t, [c_int]),
    "qspec_argmax_rows_static_max_rows": (c_int, []),
    "qspec_argmax_rows_static": (c_int, [c_void_p, c_int, c_int, ctypes.c_longlong, c_void_p, c_void_p, c_void_p]),
    "qspec_argmax_rows": (c_int, [c_void_p, c_int, c_int, ctypes.c_longlong, c_void_p, c_void_p, c_void_p, ctypes.c_size_t,
                                  c_void_p]),
    "qspec_hadamard_k108": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_float,
                                    c_void_p]),
    "qspec_hadamard_heads40": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_float,
                                       c_void_p]),
    "qspec_rope_qkhad_kvcache": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                         c_int, c_int, c_float, c_void_p]),
    "qspec_attn_block_supported": (c_int, [c_int, c_int, c_int, c_int, c_int]),
    "qspec_attn_block": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_void_p, c_int, c_int, c_int, c_int, c_int, c_float, c_float, c_float, c_void_p]),
}

_lib = None


class QSpecNativeError(RuntimeError):
    """A C-ABI call returned a non-zero status (the reference raises RuntimeError via TORCH_CHECK)."""


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m qspec_b200.build` "
            "(qspec_b200 has no CPU or PyTorch fallback for its kernels)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the header and the library drifted
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    """Status 1/2 (bad argument / unsupported shape) -> ValueError, anything else -> QSpecNativeError."""
    if status == QSPEC_OK:
        return
    msg = load().qspec_last_error()
    text = msg.decode() if msg else "?"
    if status in (1, 2):
        raise ValueError(f"{what}: {text} (status {status})")
    raise QSpecNativeError(f"{what}: {text} (status {status})")
