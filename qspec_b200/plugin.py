"""vLLM quantization-plugin surface for the dual-mode layer.

Mirrors the reference fork's API (no dependency on an installed vLLM):
  QuantizeMethodBase / QuantizationConfig   vllm/model_executor/layers/quantization/base_config.py:9-138
  LinearMethodBase.create_weights / apply   vllm/model_executor/layers/linear.py:85-116
  register_quantization_config / lookup     vllm/model_executor/layers/quantization/__init__.py:36-129
and adds the QSpec mode switch as an explicit flag (the reference threads `w4a4` through
**kwargs from ExecuteModelRequest.w4a4, spec_decode_worker.py:797-812).

`bind_to_vllm()` (or QSPEC_REGISTER_VLLM=1 at import) derives the plugin classes from the importable vLLM's own base
classes and registers them with ITS registry, so `--quantization qspec` works there; without a vLLM the local registry
below serves the same decorator / lookup API (tests).
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from contextlib import contextmanager
from typing import Any, Dict, List, Optional, Type

import torch

from . import ops
from .nn import PackedQuantizedTensor

# ----------------------------------------------------------------------------------------------
# mode switch
# ----------------------------------------------------------------------------------------------
_MODE = {"w4a4": False}


def set_qspec_mode(w4a4: bool) -> None:
    """Draft mode (W4A4) on/off; what SpecDecodeWorker does with execute_model_req.w4a4."""
    _MODE["w4a4"] = bool(w4a4)


def get_qspec_mode() -> bool:
    return _MODE["w4a4"]


@contextmanager
def qspec_mode(w4a4: bool):
    prev = _MODE["w4a4"]
    _MODE["w4a4"] = bool(w4a4)
    try:
        yield
    finally:
        _MODE["w4a4"] = prev


# ----------------------------------------------------------------------------------------------
# base classes (same method names / signatures as the reference)
# ----------------------------------------------------------------------------------------------
class QuantizeMethodBase(ABC):
    @abstractmethod
    def create_weights(self, layer: torch.nn.Module, *weight_args, **extra_weight_attrs):
        raise NotImplementedError

    @abstractmethod
    def apply(self, layer: torch.nn.Module, *args, **kwargs) -> torch.Tensor:
        raise NotImplementedError

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        return


class LinearMethodBase(QuantizeMethodBase):
    @abstractmethod
    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int,
                       output_partition_sizes: List[int], input_size: int, output_size: int,
                       params_dtype: torch.dtype, **extra_weight_attrs):
        raise NotImplementedError

    @abstractmethod
    def apply(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError


class QuantizationConfig(ABC):
    @abstractmethod
    def get_name(self) -> str: ...

    @abstractmethod
    def get_supported_act_dtypes(self) -> List[torch.dtype]: ...

    @classmethod
    @abstractmethod
    def get_min_capability(cls) -> int: ...

    @staticmethod
    @abstractmethod
    def get_config_filenames() -> List[str]: ...

    @classmethod
    @abstractmethod
    def from_config(cls, config: Dict[str, Any]) -> "QuantizationConfig": ...

    @abstractmethod
    def get_quant_method(self, layer: torch.nn.Module, prefix: str) -> Optional[QuantizeMethodBase]: ...


_REGISTRY: Dict[str, Type[QuantizationConfig]] = {}


def register_quantization_config(quantization: str):
    """quantization/__init__.py:36-69: decorator registering a custom config class by name."""

    def _wrapper(cls):
        if not issubclass(cls, QuantizationConfig):
            raise ValueError("The quantization config must be a subclass of `QuantizationConfig`.")
        if quantization in _REGISTRY:
            raise ValueError(f"The quantization method `{quantization}` is already exists.")
        _REGISTRY[quantization] = cls
        return cls

    return _wrapper


def get_quantization_config(quantization: str) -> Type[QuantizationConfig]:
    if quantization not in _REGISTRY:
        raise ValueError(f"Invalid quantization method: {quantization}")
    return _REGISTRY[quantization]


# ----------------------------------------------------------------------------------------------
# QSpec plugin
# ----------------------------------------------------------------------------------------------
@register_quantization_config("qspec")
class QSpecConfig(QuantizationConfig):
    """Explicit config replacing the reference's `"qspec" in model_path` string match
    (worker/model_runner.py:1104)."""

    def __init__(self, input_clip_ratio: float = 1.0):
        self.input_clip_ratio = input_clip_ratio

    def get_name(self) -> str:
        return "qspec"

    def get_supported_act_dtypes(self) -> List[torch.dtype]:
        return [torch.float16, torch.bfloat16]

    @classmethod
    def get_min_capability(cls) -> int:
        return 100                                   # sm_100a only

    @staticmethod
    def get_config_filenames() -> List[str]:
        return ["quantize_config.json", "qspec_config.json"]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "QSpecConfig":
        return cls(float(config.get("input_clip_ratio", 1.0)))

    def get_quant_method(self, layer: torch.nn.Module, prefix: str = "") -> Optional[QuantizeMethodBase]:
        return QSpecLinearMethod(self)


class QSpecLinearMethod(LinearMethodBase):
    """create_weights registers the reference checkpoint tensors (`weight` [N, K/2] packed int4,
    `weight_scales` [N, 1]); apply runs W4A4 or W4A16 on the same weight depending on the mode."""

    def __init__(self, quant_config: QSpecConfig):
        self.quant_config = quant_config

    def create_weights(self, layer, input_size_per_partition, output_partition_sizes, input_size, output_size,
                       params_dtype, **extra_weight_attrs):
        if input_size_per_partition % 128 != 0:
            raise ValueError("qspec: input_size_per_partition must be a multiple of 128")
        n = sum(output_partition_sizes)
        if n % 16 != 0:
            raise ValueError("qspec: output partition size must be a multiple of 16")
        weight = torch.nn.Parameter(torch.empty(n, input_size_per_partition // 2, dtype=torch.int8), requires_grad=False)
        scales = torch.nn.Parameter(torch.empty(n, 1, dtype=params_dtype), requires_grad=False)
        # TP sharding metadata used by the stock weight loaders (linear.py:131-132,1076-1112);
        # the packed dim halves K, so a K-shard of width k maps to k/2 bytes.
        for p, attrs in ((weight, {"input_dim": 1, "output_dim": 0, "packed_dim": 1, "pack_factor": 2}),
                         (scales, {"output_dim": 0})):
            for key, val in {**attrs, **extra_weight_attrs}.items():
                setattr(p, key, val)
        layer.register_parameter("weight", weight)
        layer.register_parameter("weight_scales", scales)
        layer.input_size_per_partition = input_size_per_partition
        layer.output_size_per_partition = n

    def process_weights_after_loading(self, layer) -> None:
        # checkpoint stores uint8; both kernels read the bytes as they are (no repack, no ^0x88 copy)
        if layer.weight.dtype == torch.uint8:
            layer.weight = torch.nn.Parameter(layer.weight.data.view(torch.int8), requires_grad=False)

    def apply(self, layer, x, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        w = layer.weight.data
        if w.dtype == torch.uint8:
            w = w.view(torch.int8)
        sw = layer.weight_scales.data.view(-1)
        if get_qspec_mode():
            if isinstance(x, PackedQuantizedTensor):
                q, s = x.quantized_x, x.scales_x.view(-1)
            else:                                     # generic caller handed f16 activations: quantise here
                if x.dtype != torch.float16:
                    # the reference's quantiser kernel is f16-only (quant.cu:102-166, bindings.cpp:128-146 check Half)
                    raise ValueError(f"qspec: draft mode quantises float16 activations, got {x.dtype}; pass a "
                                     "PackedQuantizedTensor or float16")
                x2 = x.reshape(-1, x.shape[-1])
                if not x2.is_contiguous():
                    x2 = x2.contiguous()
                q = torch.empty(x2.shape[0], x2.shape[1] // 2, dtype=torch.int8, device=x.device)
                s = torch.empty(x2.shape[0], dtype=torch.float16, device=x.device)
                ops.fuse_sym_quant(x2, s, q, self.quant_config.input_clip_ratio)
            out = torch.empty(q.shape[0], w.shape[0], dtype=sw.dtype, device=q.device)
            ops.rowwise_scaled_linear_cutlass_s4s4_unified(q, s[: q.shape[0]].to(sw.dtype), w, sw, bias, out)
            lead = x.quantized_x.shape[:-1] if isinstance(x, PackedQuantizedTensor) else x.shape[:-1]
            return out.view(*lead, w.shape[0])
        if isinstance(x, PackedQuantizedTensor):
            raise ValueError("qspec: verify mode (w4a4=False) expects fp16/bf16 activations")
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        out = torch.empty(x2.shape[0], w.shape[0], dtype=x.dtype, device=x.device)
        ops.gemm_w4a16(x2, w, sw.to(x.dtype), bias, out)
        return out.view(*x.shape[:-1], w.shape[0])


_VLLM_BOUND: Dict[str, Any] = {}


def bind_to_vllm(name: str = "qspec"):
    """Registers the plugin with the IMPORTABLE vLLM (the reference fork, or an upstream vLLM with the same plugin API):
    builds `QSpecConfig` / `QSpecLinearMethod` subclasses on top of vLLM's own `QuantizationConfig` / `LinearMethodBase`
    (so `issubclass` checks in `register_quantization_config`, `get_quantization_config` and the parallel linear layers
    hold) and registers the config under `name`, after which `--quantization qspec` / `quantization="qspec"` selects it
    (vllm/model_executor/layers/quantization/__init__.py:36-129; pattern of
    tests/quantization/test_register_quantization_config.py:21-116). Returns (config class, linear-method class).
    Raises ImportError when no vLLM is importable. Idempotent."""
    if name in _VLLM_BOUND:
        return _VLLM_BOUND[name]
    from vllm.model_executor.layers.linear import LinearMethodBase as _VLinearMethodBase  # type: ignore
    from vllm.model_executor.layers.quantization import register_quantization_config as _vreg  # type: ignore
    from vllm.model_executor.layers.quantization.base_config import QuantizationConfig as _VQuantizationConfig  # type: ignore
    try:
        from vllm.model_executor.layers.linear import LinearBase as _VLinearBase  # type: ignore
    except Exception:                                   # noqa: BLE001 - older forks have no common base
        _VLinearBase = None

    class VllmQSpecLinearMethod(QSpecLinearMethod, _VLinearMethodBase):
        pass

    class VllmQSpecConfig(QSpecConfig, _VQuantizationConfig):
        def __init__(self, input_clip_ratio: float = 1.0):
            _VQuantizationConfig.__init__(self)
            QSpecConfig.__init__(self, input_clip_ratio)

        def get_name(self) -> str:
            return name

        def get_quant_method(self, layer: torch.nn.Module, prefix: str = ""):
            if _VLinearBase is not None and not isinstance(layer, _VLinearBase) and type(layer).__module__.startswith("vllm."):
                return None                              # embeddings, attention, MoE ...: not quantised by this method
            return VllmQSpecLinearMethod(self)

    VllmQSpecConfig.__name__ = VllmQSpecConfig.__qualname__ = "QSpecConfig"
    VllmQSpecLinearMethod.__name__ = VllmQSpecLinearMethod.__qualname__ = "QSpecLinearMethod"
    try:
        _vreg(name)(VllmQSpecConfig)
    except ValueError as e:                              # "already exists": a previous process-wide registration
        if "exist" not in str(e):
            raise
    _VLLM_BOUND[name] = (VllmQSpecConfig, VllmQSpecLinearMethod)
    return _VLLM_BOUND[name]


import os as _os

if _os.environ.get("QSPEC_REGISTER_VLLM", "0") == "1":      # opt-in at import: importing vLLM costs tens of seconds
    bind_to_vllm()
