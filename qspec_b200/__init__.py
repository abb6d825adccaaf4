"""qspec_b200 -- B200-native (sm_100a) implementation of QSpec's dual-mode int4 linear path.

Host side mirrors the reference's module / plugin interfaces; all compute goes through the
C-ABI CUDA library `qspec_b200/_C/libqspec_b200.so` (include/qspec_b200.h). No CPU fallback.
"""
from . import _lib, ops  # noqa: F401
from .nn import (Linear4bit, OnlineHadamard, PackedQuantizedTensor, Quantizer, RMSNorm,  # noqa: F401
                 get_matmul_op_w4a16, pack_i4, unpack_i4)

__version__ = "0.1.0"
