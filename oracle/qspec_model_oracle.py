"""CPU oracle of the QSpec decode loop (TEST INFRASTRUCTURE ONLY -- see qspec_oracle.py header).

Restates vllm/model_executor/models/quarot_llama.py (layer wiring :175-244, :263-299, :363-392,
final norm :533) and the greedy speculative step of vllm/spec_decode/spec_decode_worker.py:758-915
on top of the per-kernel oracle functions. RoPE / attention / lm_head are plain numpy fp32
(those ops are outside the bit-exact contract of the linear path).
"""
from __future__ import annotations

import numpy as np

from . import qspec_oracle as O

F32, F16 = np.float32, np.float16


class OracleModel:
    """tp == 1: the reference model. tp > 1: the Megatron sharding of SURVEY 8(e) as the reference's parallel layers run
    it (ColumnParallel qkv / gate_up without a collective, RowParallelLinear.forward for o / down,
    vllm/model_executor/layers/linear.py:1125-1149): every rank quantises and rotates only ITS K-slice -- per-(token, rank)
    activation scales, shard-local rotations H_{heads/t} and had28 (x) H_{n/t} (SURVEY 7 hard part 5) -- writes an f16
    partial output, and the partials are summed in rank order in fp32 with one rounding to f16 (the value an all-reduce
    leaves on every rank). `ranks[r]` holds rank r's shard of every layer; lm_head is vocab-sharded in rank order."""

    def __init__(self, cfg, layers, embed, lm_head, cos_sin, max_batch, ranks=None):
        """layers: list of dicts with packed weights/scales: qkv, o, gate_up, down -> (w uint8, s f16) (tp == 1), or
        ranks: list over ranks of such lists; lm_head: [vocab, H] (rank-ordered concatenation of the shards)."""
        self.cfg = cfg
        self.ranks = ranks if ranks is not None else [layers]
        self.tp = len(self.ranks)
        self.layers = self.ranks[0]
        self.embed, self.lm_head, self.cos_sin = embed, lm_head, cos_sin
        L, S = cfg.num_hidden_layers, cfg.max_seq_len
        assert cfg.num_attention_heads % self.tp == 0 and cfg.num_key_value_heads % self.tp == 0
        self.hq, self.hkv = cfg.num_attention_heads // self.tp, cfg.num_key_value_heads // self.tp
        self.inter = cfg.intermediate_size // self.tp
        # one KV cache per rank (its own kv heads)
        self.ck = [np.zeros((L, max_batch, S, self.hkv, cfg.head_dim), dtype=F16) for _ in range(self.tp)]
        self.cv = [np.zeros_like(c) for c in self.ck]

    @staticmethod
    def _layers_of(m):
        def pair(lin):
            w = (lin.weight.detach().cpu().numpy().view(np.uint8).copy(),
                 lin.weight_scales.detach().cpu().numpy().reshape(-1).copy())
            b = getattr(lin, "bias", None)
            return w + ((b.detach().cpu().numpy().reshape(-1).copy(),) if b is not None else ())
        return [dict(qkv=pair(l.qkv_proj), o=pair(l.o_proj), gate_up=pair(l.gate_up), down=pair(l.down_proj))
                for l in m.layers]

    @staticmethod
    def from_gpu_model(m, max_batch):
        return OracleModel.from_gpu_shards([m], max_batch)

    @staticmethod
    def from_gpu_shards(shards, max_batch):
        """shards: the QSpecLlamaModel of every rank, in rank order (tp == len(shards))."""
        m0 = shards[0]
        assert [s.tp_rank for s in shards] == list(range(len(shards))) and all(s.tp_size == len(shards) for s in shards)
        head = np.concatenate([s.lm_head.cpu().numpy() for s in shards], 0)
        return OracleModel(m0.cfg, None, m0.embed.cpu().numpy(), head, m0.cos_sin.cpu().numpy(), max_batch,
                           ranks=[OracleModel._layers_of(s) for s in shards])

    # -- linear in either mode ----------------------------------------------------------------
    def _linear(self, w, x_or_packed, w4a4):
        wp, sw = w[0], w[1]
        bias = w[2] if len(w) > 2 else None
        cache = self.__dict__.setdefault("_unpacked", {})
        wu = cache.get(id(wp))
        if wu is None:
            wu = cache[id(wp)] = O.unpack_i4(wp)         # int8 [N, K], unpacked once per weight
        if w4a4:
            q, s = x_or_packed
            return O.gemm_w4a4(q, s, wp, sw, bias, w_unpacked=wu)
        return O.gemm_w4a16(x_or_packed, wp, sw, bias, w_unpacked=wu)

    def _rope_cache(self, li, qkv, positions, seq_ids, r=0):
        cfg = self.cfg
        Hq, Hkv, D = self.hq, self.hkv, cfg.head_dim
        T = qkv.shape[0]
        x = qkv.reshape(T, Hq + 2 * Hkv, D).astype(F32)
        cs = self.cos_sin[positions].astype(F32)                       # [T, D]
        c, s = cs[:, None, : D // 2], cs[:, None, D // 2:]
        rot = x[:, : Hq + Hkv]
        x1, x2 = rot[..., : D // 2], rot[..., D // 2:]
        a = ((x1 * c).astype(F32) - (x2 * s).astype(F32)).astype(F32)  # same op order as the kernel
        b = ((x2 * c).astype(F32) + (x1 * s).astype(F32)).astype(F32)
        rot16 = np.concatenate((a, b), axis=-1).astype(F16)
        if getattr(cfg, "qk_hadamard", False):
            # Qwen2Attention.qk_had_trans (qwen2_quarot.py:279-288): fp32 FHT_D of the f16 rope output * 1/sqrt(D) -> f16
            rot16 = O.fht(rot16.reshape(-1, D), 1.0 / float(np.sqrt(float(D)))).reshape(rot16.shape)   # python-float scale (:219)
        q = rot16[:, :Hq]
        k = rot16[:, Hq:]
        v = qkv.reshape(T, Hq + 2 * Hkv, D)[:, Hq + Hkv:]
        for t in range(T):
            self.ck[r][li, seq_ids[t], positions[t]] = k[t]
            self.cv[r][li, seq_ids[t], positions[t]] = v[t]
        return q

    def _attention(self, li, q, positions, seq_ids, r=0):
        cfg = self.cfg
        Hq, Hkv, D = self.hq, self.hkv, cfg.head_dim
        T = q.shape[0]
        out = np.zeros((T, Hq, D), dtype=F16)
        scale = F32(1.0 / np.sqrt(D))
        for t in range(T):
            n = positions[t] + 1
            K = self.ck[r][li, seq_ids[t], :n].astype(F32)            # [n, Hkv, D]
            V = self.cv[r][li, seq_ids[t], :n].astype(F32)
            for h in range(Hq):
                hk = h // (Hq // Hkv)
                sc = K[:, hk] @ (q[t, h].astype(F32) * scale)
                p = np.exp(sc - sc.max())
                out[t, h] = ((p @ V[:, hk]) / p.sum()).astype(F16)
        return out.reshape(T, Hq * D)

    # -- rounding-sensitivity probe --------------------------------------------------------------------------
    # RoPE / attention / SiLU are outside the bit-exact contract: the GPU evaluates them in fp32 with its own summation
    # order and exp approximation, so a tiny fraction of their f16 outputs differs from this oracle's by one ulp. The int4
    # quantisers that follow can amplify such a difference into a different greedy token. `jitter = (seed, fraction)`
    # moves that fraction of the attention and SiLU*up outputs by +-1 f16 ulp; a run whose tokens change under it was
    # decided by rounding, not by the algorithm (tests use this to tell an ill-conditioned seed from a real mismatch).
    jitter = None

    def _jitter(self, a):
        if self.jitter is None:
            return a
        seed, frac = self.jitter
        self._jitter_calls = getattr(self, "_jitter_calls", 0) + 1
        rng = np.random.default_rng(seed * 1000003 + self._jitter_calls)
        pick = rng.random(a.shape) < frac
        up = rng.random(a.shape) < 0.5
        moved = np.where(up, np.nextafter(a, F16(np.inf)), np.nextafter(a, F16(-np.inf))).astype(F16)
        return np.where(pick & np.isfinite(moved), moved, a).astype(F16)

    @staticmethod
    def _all_reduce(parts):
        """Rank-ordered fp32 sum of the f16 partial outputs, one rounding (RowParallelLinear's all-reduce as
        qspec_tp_add_norm_* performs it; tp == 1: the partial itself)."""
        acc = parts[0].astype(F32)
        for p in parts[1:]:
            acc = (acc + p.astype(F32)).astype(F32)
        return acc.astype(F16)

    def forward(self, input_ids, positions, seq_ids, w4a4, trace=None):
        """trace (optional list): receives the residual stream after every block, for per-layer comparisons."""
        cfg = self.cfg
        D = cfg.head_dim
        x = self.embed[input_ids].astype(F16)
        # shard-local rotation scales as the reference's run time evaluates `1.0 / torch.tensor(n).sqrt()`: in float16, because
        # the model is built under torch.set_default_dtype(torch.float16) (worker/model_runner.py:1109; quarot_nn/hadamard.py:20-21)
        sc_h = float(F16(1.0) / np.sqrt(F16(self.hq)))
        sc_m = float(F16(1.0) / np.sqrt(F16(self.inter)))
        hadK = getattr(cfg, "had_k", 28)
        for li in range(cfg.num_hidden_layers):
            a_in = O.norm_quant_i4(x, cfg.rms_norm_eps)[:2] if w4a4 else O.norm_f16(x, cfg.rms_norm_eps)
            parts = []
            for r, layers in enumerate(self.ranks):
                lw = layers[li]
                qkv = self._linear(lw["qkv"], a_in, w4a4)
                qr = self._rope_cache(li, qkv, positions, seq_ids, r)
                attn = self._jitter(self._attention(li, qr, positions, seq_ids, r))
                if getattr(cfg, "heads_rotation", "sylvester") == "had40":
                    rot = O.hadamard_heads_hadK(attn, self.hq, D, 1.0 / float(np.sqrt(float(self.hq))))   # o_had_trans, qwen2_quarot.py:217,269-277
                else:
                    rot = O.hadamard_heads(attn, self.hq, D, sc_h)
                parts.append(self._linear(lw["o"], O.sym_quant_i4(rot) if w4a4 else rot, w4a4))
            x = (x.astype(F32) + self._all_reduce(parts).astype(F32)).astype(F16)
            if trace is not None:
                trace.append(x.copy())
            a_in = O.norm_quant_i4(x, cfg.rms_norm_eps)[:2] if w4a4 else O.norm_f16(x, cfg.rms_norm_eps)
            parts = []
            for r, layers in enumerate(self.ranks):
                lw = layers[li]
                gu = self._linear(lw["gate_up"], a_in, w4a4)
                act = self._jitter((O.silu_mul_gate_first if getattr(cfg, "gate_first", False) else O.silu_mul)(gu, self.inter))
                rot = O.hadamard_kK(act, self.inter // hadK, hadK, sc_m)
                parts.append(self._linear(lw["down"], O.sym_quant_i4(rot) if w4a4 else rot, w4a4))
            x = (x.astype(F32) + self._all_reduce(parts).astype(F32)).astype(F16)
            if trace is not None:
                trace.append(x.copy())
        return O.norm_f16(x, cfg.rms_norm_eps)

    def greedy(self, hidden):
        logits = hidden.astype(F32) @ self.lm_head.astype(F32).T
        return logits.argmax(-1), logits


def spec_step(model: OracleModel, cur_tok, seq_len, k):
    """One greedy speculative step. Returns (emitted list per seq, n_acc, new cur_tok, new seq_len, margins)."""
    B = len(cur_tok)
    seq_ids = np.arange(B)
    tok = cur_tok.copy()
    draft = np.zeros((B, k), dtype=np.int64)
    margins = []
    for i in range(k):
        h = model.forward(tok, seq_len + i, seq_ids, True)
        tok, lg = model.greedy(h)
        margins.append(_margin(lg))
        draft[:, i] = tok
    vt = np.concatenate((cur_tok[:, None], draft), 1).reshape(-1)
    vpos = (seq_len[:, None] + np.arange(k + 1)[None]).reshape(-1)
    vseq = np.repeat(seq_ids, k + 1)
    h = model.forward(vt, vpos, vseq, False)
    tgt, lg = model.greedy(h)
    margins.append(_margin(lg))
    tgt = tgt.reshape(B, k + 1)
    match = (tgt[:, :k] == draft).astype(np.int64)
    n_acc = np.cumprod(match, 1).sum(1)
    nxt = tgt[np.arange(B), n_acc]
    emitted = [list(draft[b, : n_acc[b]]) + [int(nxt[b])] for b in range(B)]
    return emitted, n_acc, nxt, seq_len + n_acc + 1, float(min(margins))


def _margin(logits):
    s = np.sort(logits, axis=-1)
    return float((s[:, -1] - s[:, -2]).min())
