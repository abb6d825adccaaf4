"""Weight-format loader for the published QSpec / QuaRot int4 checkpoint; host-side logic only.

What the reference does at load time, restated once here so that the decode engine can take real weights:

* format (third-party/QuaRot/e2e/checkpoint_utils/quantize_llama_checkpoint.py:87-119): per linear a packed `weight` uint8
  `[N, K/2]` (two's-complement nibbles, even k in the low nibble) and `weight_scales` f16 `[N, 1]`; the RTN quantiser is
  `clamp(round(W / scale), -8, 7)`; keys follow the HF Llama names with `self_attn.o_proj.1.*`, `mlp.down_proj.0.*` (the online
  rotation's buffer) and `mlp.down_proj.2.*`; norm weights are folded into the linears and dropped.
* key remapping (vllm/worker/model_runner.py:1132-1141): `o_proj.1.` -> `o_proj.`, `down_proj.0.` -> `online_hadamard.`,
  `down_proj.2.` -> `down_proj.`; two safetensors shards merged into one state dict.
* fusion (vllm/model_executor/models/quarot_llama.py:152-173, 301-314): qkv = [q ; k ; v] rows, gate_up = [up ; gate] rows, weights
  and scales alike.
* tensor parallelism: qkv / gate_up column-sharded by heads / by co-located up+gate slices, o / down row-sharded
  along the packed K dimension with replicated scales, lm_head vocab-sharded (`qspec_b200.parallel`).

Both kernels read the checkpoint bytes as they are: no repacking, no offset-binary copy.
"""
from __future__ import annotations

import glob
import os
from typing import Dict, Mapping, Optional

import torch

from .nn import pack_i4
from .parallel import shard_column, shard_gate_up, shard_qkv, shard_row

_RENAMES = (("o_proj.1.", "o_proj."), ("down_proj.0.", "online_hadamard."), ("down_proj.2.", "down_proj."))


def remap_reference_keys(state_dict: Mapping[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """model_runner.py:1137: the checkpoint's Sequential indices -> the module names of quarot_llama.py."""
    out = {}
    for k, v in state_dict.items():
        for old, new in _RENAMES:
            k = k.replace(old, new)
        out[k] = v
    return out


def to_reference_keys(state_dict: Mapping[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Inverse of `remap_reference_keys` (what quantize_llama_checkpoint.py:87-90 writes)."""
    out = {}
    for k, v in state_dict.items():
        k = k.replace("self_attn.o_proj.", "self_attn.o_proj.1.").replace("mlp.online_hadamard.", "mlp.down_proj.0.")
        if "mlp.down_proj." in k and "mlp.down_proj.0." not in k:
            k = k.replace("mlp.down_proj.", "mlp.down_proj.2.")
        out[k] = v
    return out


def quantize_weight_rtn(weight: torch.Tensor, weight_scales: torch.Tensor):
    """quantize_llama_checkpoint.py:111-119: `clamp(round(W / scale), -8, 7)` packed to int4.
    weight [N, K] float, weight_scales [N, 1] (or [N]) -> (uint8 [N, K/2], f16 [N, 1])."""
    if weight.dim() != 2 or weight.shape[1] % 2:
        raise ValueError("quantize_weight_rtn: weight must be [N, K] with even K")
    s = weight_scales.reshape(-1, 1)
    if s.shape[0] != weight.shape[0]:
        raise ValueError("quantize_weight_rtn: one scale per output channel expected")
    q = (weight / s.to(weight.dtype)).round().to(torch.int8)
    q = torch.clamp(q, -8, 7)
    return pack_i4(q), s.to(torch.float16)


def rtn_scales(weight: torch.Tensor) -> torch.Tensor:
    """Per-channel symmetric scale max|w| / 7 (fake_quant/quant_utils.py:8-28, the no-clip-search case)."""
    return (weight.abs().amax(dim=1, keepdim=True).float() / 7.0).clamp_min(1e-8).to(torch.float16)


def fuse_qkv(q, k, v, sq, sk, sv):
    """quarot_llama.py:152-173: rows [q ; k ; v], scales flattened in the same order."""
    return torch.cat((q, k, v), 0).contiguous(), torch.cat((sq.reshape(-1), sk.reshape(-1), sv.reshape(-1)), 0).contiguous()


def fuse_gate_up(gate, up, sgate, sup):
    """quarot_llama.py:301-314: rows [up ; gate] (NOT gate first)."""
    return torch.cat((up, gate), 0).contiguous(), torch.cat((sup.reshape(-1), sgate.reshape(-1)), 0).contiguous()


def _as_i8(w: torch.Tensor) -> torch.Tensor:
    if w.dtype == torch.uint8:
        return w.view(torch.int8)
    if w.dtype != torch.int8:
        raise ValueError(f"packed int4 weight must be uint8/int8, got {w.dtype}")
    return w


def layer_tensors(state_dict: Mapping[str, torch.Tensor], layer: int, cfg, tp: int = 1, rank: int = 0,
                  rerotate: bool = True) -> Dict[str, torch.Tensor]:
    """The four fused, TP-sharded (weight int8 [N, K/2], scales f16 [N]) pairs of one decoder layer, from a state dict with
    REMAPPED keys. Raises KeyError naming the missing tensor, ValueError on a shape that does not match `cfg`.
    rerotate=True (the default; only matters at tp > 1): o_proj / down_proj shards are re-rotated for the shard-local online
    rotations the TP model applies (`rerotate_row_parallel`); False slices the checkpoint bytes as they are -- right ONLY for
    synthetic weights that were never folded with the global rotation (a real checkpoint sliced that way is > 90 % off)."""
    p = f"model.layers.{layer}."

    def get(name):
        w, s = state_dict[p + name + ".weight"], state_dict[p + name + ".weight_scales"]
        if s.numel() != w.shape[0]:
            raise ValueError(f"{p}{name}: {s.numel()} scales for {w.shape[0]} channels")
        return _as_i8(w), s.reshape(-1).to(torch.float16)

    D, H, I = cfg.head_dim, cfg.hidden_size, cfg.intermediate_size
    hq, hkv = cfg.num_attention_heads, cfg.num_key_value_heads
    q, sq = get("self_attn.q_proj"); k, sk = get("self_attn.k_proj"); v, sv = get("self_attn.v_proj")
    o, so = get("self_attn.o_proj")
    g, sg = get("mlp.gate_proj"); u, su = get("mlp.up_proj")
    d, sd = get("mlp.down_proj")
    expect = {"q_proj": (q, (hq * D, H // 2)), "k_proj": (k, (hkv * D, H // 2)), "v_proj": (v, (hkv * D, H // 2)),
              "o_proj": (o, (H, hq * D // 2)), "gate_proj": (g, (I, H // 2)), "up_proj": (u, (I, H // 2)), "down_proj": (d, (H, I // 2))}
    for name, (t, shape) in expect.items():
        if tuple(t.shape) != shape:
            raise ValueError(f"{p}{name}: packed weight {tuple(t.shape)}, config expects {shape}")
    qkv, sqkv = fuse_qkv(q, k, v, sq, sk, sv)
    gu, sgu = fuse_gate_up(g, u, sg, su)
    out = {}
    out["qkv_proj.weight"], out["qkv_proj.weight_scales"] = shard_qkv(qkv, sqkv, hq, hkv, D, tp, rank)
    out["gate_up.weight"], out["gate_up.weight_scales"] = shard_gate_up(gu, sgu, I, tp, rank)
    if rerotate and tp > 1:
        wo, s_o = rerotate_row_parallel(o, so, "heads", tp, D, rank=rank)
        wd, s_d = rerotate_row_parallel(d, sd, "mlp", tp, D, rank=rank)
        out["o_proj.weight"], out["o_proj.weight_scales"] = wo.view(torch.int8), s_o.reshape(-1)
        out["down_proj.weight"], out["down_proj.weight_scales"] = wd.view(torch.int8), s_d.reshape(-1)
    else:
        out["o_proj.weight"], out["o_proj.weight_scales"] = shard_row(o, so, tp, rank)
        out["down_proj.weight"], out["down_proj.weight_scales"] = shard_row(d, sd, tp, rank)
    return out


def load_state_dict(model, state_dict: Mapping[str, torch.Tensor], reference_keys: bool = True,
                    rerotate: Optional[bool] = None) -> None:
    """Copies a QSpec checkpoint into a `QSpecLlamaModel` (any TP degree: the model's own tp_size / tp_rank select the shard).
    Shapes are validated against the model's config before anything is written. rerotate=None (default) re-rotates the
    row-parallel weights whenever the model is tensor parallel and the state dict uses the reference's key names (i.e. it is
    a real, globally rotated checkpoint); pass False explicitly for synthetic weights saved by `export_state_dict`."""
    sd = remap_reference_keys(state_dict) if reference_keys else dict(state_dict)
    cfg, tp, rank = model.cfg, model.tp_size, model.tp_rank
    if rerotate is None:
        rerotate = tp > 1 and reference_keys
    staged = [layer_tensors(sd, i, cfg, tp, rank, rerotate) for i in range(cfg.num_hidden_layers)]
    embed, head = sd["model.embed_tokens.weight"], sd.get("lm_head.weight", sd["model.embed_tokens.weight"])
    if tuple(embed.shape) != (cfg.vocab_size, cfg.hidden_size) or tuple(head.shape) != (cfg.vocab_size, cfg.hidden_size):
        raise ValueError("embed_tokens / lm_head do not match the config's [vocab, hidden]")
    dev = model.device
    with torch.no_grad():
        for layer, t in zip(model.layers, staged):
            for name in ("qkv_proj", "o_proj", "gate_up", "down_proj"):
                lin = getattr(layer, name)
                w, s = t[name + ".weight"], t[name + ".weight_scales"]
                if tuple(lin.weight.shape) != tuple(w.shape):
                    raise ValueError(f"{name}: model holds {tuple(lin.weight.shape)}, checkpoint shard is {tuple(w.shape)}")
                lin.weight = w.to(dev)
                lin.weight_scales = s.to(dev)
        model.embed = embed.to(torch.float16).to(dev)
        vs = cfg.vocab_size // tp
        model.lm_head = head[rank * vs:(rank + 1) * vs].to(torch.float16).contiguous().to(dev)


# ----------------------------------------------------------------------------------------------
# TP-aware re-rotation of the row-parallel weights (SURVEY 7 hard part 5, 8(f) rank 3)
#
# The checkpoint's o_proj / down_proj weights were folded with the GLOBAL online rotation R (o_proj: H_heads (x) I_d / sqrt(heads),
# quarot_llama.py:230-234; down_proj: (had28 (x) H_n) / sqrt(inter), hadamard.py:109-113): y = W_ckpt . (R x). At TP = t every rank
# rotates only its own slice x_r with the shard-local R_t (H_{heads/t}, had28 (x) H_{n/t}), so rank r needs
#     W'_r = (W_ckpt . R)[:, slice_r] . R_t^T      (then  sum_r W'_r (R_t x_r) = W_ckpt (R x) exactly, R_t orthogonal)
# which is dequantise -> two small Kronecker-factor products -> re-quantise (RTN, per-channel max|w|/7). The re-quantisation is
# lossy like the original one; the exact-global alternative (one reduce-scatter of the activations) needs no new weights.
# ----------------------------------------------------------------------------------------------

def _sylvester(n: int) -> torch.Tensor:
    if n < 1 or n & (n - 1):
        raise ValueError(f"Sylvester-Hadamard size must be a power of two, got {n}")
    h = torch.ones(1, 1, dtype=torch.float64)
    while h.shape[0] < n:
        h = torch.cat((torch.cat((h, h), 1), torch.cat((h, -h), 1)), 0)
    return h


def _rotation_factors(kind: str, width: int, head_dim: int):
    """(A [a, a], B [b, b] or None, norm): the online rotation of a row of `width` values viewed [a, b] is
    y[i, :] = sum_k A[i, k] (x[k, :] B) / norm."""
    if kind == "heads":                                  # H over the heads, identity over head_dim
        h = width // head_dim
        return _sylvester(h), None, float(h) ** 0.5
    if kind == "mlp":                                    # had28 over the 28 blocks, H_n inside a block
        if width % 28:
            raise ValueError(f"mlp rotation: width {width} is not 28 * 2^k")
        from ._had28 import had28
        return had28().to(torch.float64), _sylvester(width // 28), float(width) ** 0.5
    raise ValueError(f"unknown rotation kind {kind!r}")


def rotate_rows(x: torch.Tensor, kind: str, head_dim: int = 128, transpose: bool = False) -> torch.Tensor:
    """x [..., width] float64 -> R x (or R^T x) along the last axis, R the online rotation above."""
    width = x.shape[-1]
    A, B, norm = _rotation_factors(kind, width, head_dim)
    a = A.shape[0]
    v = x.reshape(*x.shape[:-1], a, width // a)
    if B is not None:
        v = v @ (B if not transpose else B.T)           # Sylvester B is symmetric; kept explicit
    v = torch.einsum("ik,...kb->...ib", A.T if transpose else A, v)
    return (v / norm).reshape(x.shape)


def rerotate_row_parallel(weight: torch.Tensor, weight_scales: torch.Tensor, kind: str, tp: int, head_dim: int = 128,
                          rank: Optional[int] = None):
    """Packed row-parallel weight [N, K/2] + scales of the checkpoint -> (packed uint8 [N, K/(2 tp)], f16 [N, 1]) for shard-local
    online rotations: the pair of `rank`, or the list over all ranks when rank is None. tp = 1 returns the input unchanged."""
    from .nn import unpack_i4
    if tp == 1:
        one = (weight.view(torch.uint8), weight_scales.reshape(-1, 1).to(torch.float16))
        return one if rank is not None else [one]
    if rank is not None and not 0 <= rank < tp:
        raise ValueError(f"rerotate_row_parallel: rank {rank} of {tp}")
    w = unpack_i4(weight).to(torch.float64) * weight_scales.reshape(-1, 1).to(torch.float64)
    K = w.shape[1]
    if K % tp or (kind == "heads" and (K // head_dim) % tp) or (kind == "mlp" and (K // 28) % tp):
        raise ValueError(f"rerotate_row_parallel: K={K} does not shard {tp} ways for the {kind} rotation")
    # W_eff = W_ckpt . R : every ROW of W is multiplied from the right by R, i.e. row <- R^T row
    w_eff = rotate_rows(w, kind, head_dim, transpose=True)
    out = []
    k = K // tp
    for r in (range(tp) if rank is None else (rank,)):
        shard = w_eff[:, r * k:(r + 1) * k]
        w_loc = rotate_rows(shard.contiguous(), kind, head_dim, transpose=False)      # row <- R_t row  ==  W_eff_r . R_t^T
        sc = rtn_scales(w_loc.to(torch.float32))
        out.append(quantize_weight_rtn(w_loc.to(torch.float32), sc.to(torch.float32)))
    return out[0] if rank is not None else out


def read_safetensors_dir(path: str) -> Dict[str, torch.Tensor]:
    """model_runner.py:1128-1136: every `model-*.safetensors` shard of the directory merged into one dict (CPU tensors)."""
    from safetensors.torch import load_file
    files = sorted(glob.glob(os.path.join(path, "*.safetensors")))
    if not files:
        raise FileNotFoundError(f"no .safetensors files under {path}")
    sd: Dict[str, torch.Tensor] = {}
    for f in files:
        sd.update(load_file(f))
    return sd


def load_qspec_checkpoint(model, path: str, rerotate: Optional[bool] = None) -> None:
    load_state_dict(model, read_safetensors_dir(path), reference_keys=True, rerotate=rerotate)


def export_state_dict(model, reference_keys: bool = True) -> Dict[str, torch.Tensor]:
    """TP=1 only: the model's fused tensors split back into the checkpoint's per-projection tensors (uint8 weights, [N, 1] scales)."""
    if model.tp_size != 1:
        raise ValueError("export_state_dict: gather the shards first (TP=1 only)")
    cfg = model.cfg
    D, I = cfg.head_dim, cfg.intermediate_size
    nq, nkv = cfg.num_attention_heads * D, cfg.num_key_value_heads * D
    sd: Dict[str, torch.Tensor] = {"model.embed_tokens.weight": model.embed.cpu(), "lm_head.weight": model.lm_head.cpu()}

    def put(key, w, s):
        sd[key + ".weight"] = w.contiguous().view(torch.uint8).cpu()
        sd[key + ".weight_scales"] = s.reshape(-1, 1).contiguous().cpu()

    for i, layer in enumerate(model.layers):
        p = f"model.layers.{i}."
        w, s = layer.qkv_proj.weight, layer.qkv_proj.weight_scales.reshape(-1)
        put(p + "self_attn.q_proj", w[:nq], s[:nq])
        put(p + "self_attn.k_proj", w[nq:nq + nkv], s[nq:nq + nkv])
        put(p + "self_attn.v_proj", w[nq + nkv:], s[nq + nkv:])
        put(p + "self_attn.o_proj", layer.o_proj.weight, layer.o_proj.weight_scales)
        w, s = layer.gate_up.weight, layer.gate_up.weight_scales.reshape(-1)
        put(p + "mlp.up_proj", w[:I], s[:I])
        put(p + "mlp.gate_proj", w[I:], s[I:])
        put(p + "mlp.down_proj", layer.down_proj.weight, layer.down_proj.weight_scales)
    return to_reference_keys(sd) if reference_keys else sd


def save_qspec_checkpoint(model, path: str, shards: int = 2) -> None:
    """Writes `model-0000i-of-0000n.safetensors` like the published checkpoint (two shards, model_runner.py:1130-1131)."""
    from safetensors.torch import save_file
    sd = export_state_dict(model, reference_keys=True)
    os.makedirs(path, exist_ok=True)
    keys = sorted(sd)
    per = (len(keys) + shards - 1) // shards
    for i in range(shards):
        part = {k: sd[k].contiguous() for k in keys[i * per:(i + 1) * per]}
        save_file(part, os.path.join(path, f"model-{i + 1:05d}-of-{shards:05d}.safetensors"))
