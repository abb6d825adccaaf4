"""Drop-in proof against the reference's OWN caller (INTEGRATION.md option A).

The reference's model file vllm/model_executor/models/quarot_llama.py is executed UNMODIFIED (from /root/reference when
present, else from the copy `oracle/build_ref.py` staged into the git-ignored oracle/_ref/pyref/ -- the GPU box has no
/root/reference) with `vllm.model_executor.layers.quarot_nn` bound to `qspec_b200.nn`. Its QuarotLlamaAttention /
QuarotLlamaMLP / QuarotDecoderLayer (:62-434) then run one Llama-3-8B decoder layer in both modes -- draft with the
scratch-buffer kwargs of vllm/spec_decode/draft_model_runner.py:277-320 -- and the result is compared with the oracle.

What is stubbed: only the vLLM pieces OUTSIDE the hot path that the file imports (config / sampler / loader names, and
rotary embedding + attention, restated in plain torch fp32 below). Everything on the path -- RMSNorm, Linear4bit (both
modes, fused qkv / gate_up built by the reference's own fuse_qkv / fuse_gate_up), OnlineHadamard (K = 1 and K = 28),
Quantizer, PackedQuantizedTensor -- is `qspec_b200.nn` calling the C ABI.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANDIDATES = ("/root/reference/vllm/model_executor/models/quarot_llama.py",
              os.path.join(ROOT, "oracle", "_ref", "pyref", "quarot_llama.py"))


def _reference_file():
    for p in CANDIDATES:
        if os.path.exists(p):
            return p
    return None


# ---- plain-torch stand-ins for the vLLM ops outside the path ----------------------------------------------------------
def _rope_inplace(positions, q, k, head_size, cos_sin_cache, is_neox):
    """vllm ops.rotary_embedding (neox): in place on the [T, H * D] views q and k; fp32 math in the kernel's op order."""
    assert is_neox
    cs = cos_sin_cache[positions.long()].float()
    D = head_size
    c, s = cs[:, None, : D // 2], cs[:, None, D // 2:]
    for t in (q, k):
        x = t.reshape(t.shape[0], -1, D).float()
        x1, x2 = x[..., : D // 2], x[..., D // 2:]
        t.copy_(torch.cat((x1 * c - x2 * s, x2 * c + x1 * s), -1).to(t.dtype).reshape(t.shape))


class _ERotaryEmbedding(torch.nn.Module):
    def __init__(self, head_size, rotary_dim, max_position_embeddings, base, is_neox_style, dtype):
        super().__init__()
        from qspec_b200.model import build_cos_sin_cache
        self.head_size, self.is_neox_style = head_size, is_neox_style
        self.cos_sin_cache = build_cos_sin_cache(head_size, max_position_embeddings, float(base), "cpu")

    def forward_cuda(self, positions, query, key):
        cache = self.cos_sin_cache.to(query.device)
        _rope_inplace(positions, query, key, self.head_size, cache, self.is_neox_style)
        return query, key


class _Attention(torch.nn.Module):
    """Causal attention over the tokens of ONE sequence in the pass (positions 0..T-1). The arithmetic is the oracle's own
    (OracleModel._attention, numpy fp32, on the host): the int4 draft path amplifies even 1-ulp differences of the
    attention output into percent-level differences after two quantisers, so the piece OUTSIDE the path is made
    bit-identical on both sides and the comparison isolates the path itself."""
    oracle = None                    # set by the test: an OracleModel whose KV cache this stub fills

    def __init__(self, num_heads, head_size, scale, num_kv_heads=None, **kw):
        super().__init__()
        self.h, self.d, self.scale, self.hkv = num_heads, head_size, scale, num_kv_heads or num_heads

    def forward(self, q, k, v, kv_cache, attn_metadata, out=None):
        om = type(self).oracle
        T = q.shape[0]
        om.ck[0][0, 0, :T] = k.detach().cpu().numpy().reshape(T, self.hkv, self.d)
        om.cv[0][0, 0, :T] = v.detach().cpu().numpy().reshape(T, self.hkv, self.d)
        o = om._attention(0, q.detach().cpu().numpy().reshape(T, self.h, self.d), np.arange(T), np.zeros(T, dtype=np.int64))
        o = torch.from_numpy(o).to(q.device)
        if out is not None:
            out.view(T, -1).copy_(o)
            return out.view(T, -1)
        return o


def _load_reference_module(path):
    """Executes the reference file as `vllm.model_executor.models.quarot_llama` inside a sandbox of stub modules."""
    import qspec_b200.nn as qnn
    saved = {k: v for k, v in sys.modules.items() if k == "quarot" or k.startswith("quarot.") or k == "vllm" or k.startswith("vllm.")}
    for k in saved:
        del sys.modules[k]

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        if "." in name:
            parent, _, leaf = name.rpartition(".")
            setattr(sys.modules[parent], leaf, m)
        sys.modules[name] = m
        return m

    dummy = lambda n: type(n, (), {})
    mod("quarot", __path__=[])
    mod("quarot.transformers")
    mod("vllm", __path__=[])
    mod("vllm.config", VllmConfig=dummy("VllmConfig"), CacheConfig=dummy("CacheConfig"))
    mod("vllm.model_executor", __path__=[])
    mod("vllm.model_executor.layers", __path__=[])
    sys.modules["vllm.model_executor.layers.quarot_nn"] = qnn                   # THE binding: quarot_nn -> qspec_b200.nn
    sys.modules["vllm.model_executor.layers"].quarot_nn = qnn
    mod("vllm.model_executor.layers.quantization", QuantizationConfig=dummy("QuantizationConfig"))
    mod("vllm.model_executor.layers.rotary_embedding", ERotaryEmbedding=_ERotaryEmbedding, get_rope=lambda *a, **k: None)
    mod("vllm.model_executor.sampling_metadata", SamplingMetadata=dummy("SamplingMetadata"))
    mod("vllm.platforms", current_platform=None)
    mod("vllm.sequence", IntermediateTensors=dummy("IntermediateTensors"))
    mod("vllm.attention", Attention=_Attention, AttentionMetadata=dummy("AttentionMetadata"))
    mod("vllm.model_executor.layers.sampler", SamplerOutput=dummy("SamplerOutput"), get_sampler=lambda: None)
    mod("vllm.model_executor.layers.logits_processor", LogitsProcessor=dummy("LogitsProcessor"))
    mod("vllm._custom_ops", rotary_embedding=_rope_inplace)
    mod("vllm.model_executor.models", __path__=[])
    mod("vllm.model_executor.models.utils", **{n: (lambda *a, **k: None) for n in (
        "AutoWeightsLoader", "PPMissingLayer", "extract_layer_index", "is_pp_missing_parameter",
        "make_empty_intermediate_tensors_factory", "make_layers", "maybe_prefix")})
    try:
        spec = importlib.util.spec_from_file_location("vllm.model_executor.models.quarot_llama", path)
        ref = importlib.util.module_from_spec(spec)
        sys.modules[spec.name] = ref
        spec.loader.exec_module(ref)
    finally:
        for k in [k for k in sys.modules if k == "quarot" or k.startswith("quarot.") or k == "vllm" or k.startswith("vllm.")]:
            del sys.modules[k]
        sys.modules.update(saved)
    return ref


def _random_packed(N, K, seed, device):
    g = torch.Generator().manual_seed(seed)
    w = torch.randint(0, 256, (N, K // 2), generator=g, dtype=torch.uint8)
    s = ((0.5 + torch.rand(N, 1, generator=g)) * 2.0 ** -7).to(torch.float16)
    return w.to(device), s.to(device)


def test_reference_decoder_layer_runs_unchanged_on_qspec_b200(monkeypatch=None):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    path = _reference_file()
    if path is None:
        pytest.skip("reference model file not staged (run oracle/build_ref.py where /root/reference exists)")
    from transformers import LlamaConfig
    from oracle.qspec_model_oracle import OracleModel
    import qspec_b200.nn as qnn
    from qspec_b200.model import QSpecLlamaConfig, build_cos_sin_cache
    ref = _load_reference_module(path)
    dev = torch.device(os.environ.get("QSPEC_DROPIN_DEVICE", "cuda:0"))    # "cpu" only for debugging the wiring with emulated ops
    H, I, HQ, HKV, D, S = 4096, 14336, 32, 8, 128, 64        # QuarotLlamaMLP.forward hard-codes 14336 (quarot_llama.py:283)
    cfg = LlamaConfig(hidden_size=H, intermediate_size=I, num_attention_heads=HQ, num_key_value_heads=HKV, num_hidden_layers=1,
                      vocab_size=128, max_position_embeddings=S, rope_theta=500000.0, rms_norm_eps=1e-5, attention_bias=False)
    # version skew: the fork pins a transformers whose LlamaConfig exposes `rope_theta`; transformers 5.x moved it
    # into `rope_parameters`, and the reference's `getattr(config, "rope_theta", 10000)` (:331) would silently read 10000
    cfg.rope_theta = 500000.0
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.float16)                   # as worker/model_runner.py:1117 does before building the model
    try:
        layer = ref.QuarotDecoderLayer(cfg, 0, None, None, prefix="model.layers.0")
    finally:
        torch.set_default_dtype(prev)
    assert type(layer.self_attn.q_proj) is qnn.Linear4bit and type(layer.input_layernorm) is qnn.RMSNorm
    assert type(layer.mlp.online_hadamard) is qnn.OnlineHadamard and type(layer.self_attn.quantizer) is qnn.Quantizer
    # random int4 weights in the checkpoint's format, loaded the way load_state_dict would fill the buffers
    at, mlp = layer.self_attn, layer.mlp
    packs = {}
    for i, (m, name) in enumerate(((at, "q_proj"), (at, "k_proj"), (at, "v_proj"), (at, "o_proj"), (mlp, "up_proj"),
                                   (mlp, "gate_proj"), (mlp, "down_proj"))):
        lin = getattr(m, name)
        w, s = _random_packed(lin.out_features, lin.in_features, 900 + i, "cpu")
        lin.weight.data.copy_(w)
        lin.weight_scales.data.copy_(s)
        packs[name] = (w.numpy().copy(), s.numpy().reshape(-1).copy())
    layer = layer.to(dev)
    at.fuse_qkv(None)                                          # the reference's own fusion code (:152-173)
    mlp.fuse_gate_up(None)                                     # (:301-314): rows [up ; gate]
    assert tuple(at.qkv_proj.weight.shape) == ((HQ + 2 * HKV) * D, H // 2) and at.qkv_proj.weight.dtype == torch.int8

    T = 6
    g = torch.Generator().manual_seed(5)
    x = (torch.randn(T, H, generator=g) * 0.5).to(torch.float16)
    positions = torch.arange(T, device=dev)
    # the oracle's view of the same layer
    cat = lambda names: (np.concatenate([packs[n][0] for n in names], 0), np.concatenate([packs[n][1] for n in names], 0))
    lw = dict(qkv=cat(("q_proj", "k_proj", "v_proj")), o=packs["o_proj"], gate_up=cat(("up_proj", "gate_proj")), down=packs["down_proj"])
    ocfg = QSpecLlamaConfig(num_hidden_layers=1, vocab_size=T, max_seq_len=S)
    om = OracleModel(ocfg, [lw], x.numpy(), np.zeros((1, H), np.float16), build_cos_sin_cache(D, S, 500000.0, "cpu").numpy(), 1)
    _Attention.oracle = OracleModel(ocfg, [lw], x.numpy(), np.zeros((1, H), np.float16), om.cos_sin, 1)   # separate caches

    def scratch():                                             # draft_model_runner.py:277-320
        f16, i8 = dict(dtype=torch.float16, device=dev), dict(dtype=torch.int8, device=dev)
        return {"quantized_buffer_qkv": torch.empty(T, H // 2, **i8), "quantized_buffer_mlp": torch.empty(T, I // 2, **i8),
                "act_buffer_qkv": torch.empty(T, (HQ + 2 * HKV) * D, **f16), "act_buffer_output": torch.empty(T, H, **f16),
                "act_buffer_attn": torch.empty(T, HQ, D, **f16), "act_buffer_had": torch.empty(T * D, HQ, **f16),
                "act_buffer_had_mlp": torch.empty(T * I // 512, 512, **f16), "act_buffer_gate_up": torch.empty(T, 2 * I, **f16),
                "scale_buffer": torch.empty(T, **f16), "input_sum_buffer": torch.empty(T, **f16)}

    for w4a4 in (True, False):
        kwargs = dict(scratch(), w4a4=True) if w4a4 else {}
        with torch.no_grad():
            out = layer(positions, x.to(dev).clone(), None, None, **kwargs)
        torch.cuda.synchronize()
        trace = []
        om.forward(np.arange(T), np.arange(T), np.zeros(T, dtype=np.int64), w4a4, trace=trace)
        refv = trace[-1].astype(np.float32)
        got = out.float().cpu().numpy()
        rel = np.abs(got - refv).max() / np.abs(refv).max()
        assert np.isfinite(got).all()
        # north_star: fp16 outputs of both modes within 1e-2 relative (draft: every op of the path is bit-exact, what is
        # left is exp() of the SiLU stand-in; verify: fp32 tensor-core accumulation against the oracle's float64)
        assert rel <= 1e-2, f"reference QuarotDecoderLayer on qspec_b200.nn, w4a4={w4a4}: {rel:.3e} relative from the oracle"
    assert layer.mlp.online_hadamard.qspec and layer.self_attn.o_proj_hadamard.qspec       # sticky flag set by the draft call
