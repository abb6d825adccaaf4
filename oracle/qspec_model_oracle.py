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
    def __init__(self, cfg, layers, embed, lm_head, cos_sin, max_batch):
        """layers: list of dicts with packed weights/scales: qkv, o, gate_up, down -> (w uint8, s f16)."""
        self.cfg, self.layers = cfg, layers
        self.embed, self.lm_head, self.cos_sin = embed, lm_head, cos_sin
        L, S = cfg.num_hidden_layers, cfg.max_seq_len
        self.ck = np.zeros((L, max_batch, S, cfg.num_key_value_heads, cfg.head_dim), dtype=F16)
        self.cv = np.zeros_like(self.ck)

    @staticmethod
    def from_gpu_model(m, max_batch):
        def pair(lin):
            return (lin.weight.detach().cpu().numpy().view(np.uint8).copy(),
                    lin.weight_scales.detach().cpu().numpy().reshape(-1).copy())
        layers = [dict(qkv=pair(l.qkv_proj), o=pair(l.o_proj), gate_up=pair(l.gate_up), down=pair(l.down_proj))
                  for l in m.layers]
        return OracleModel(m.cfg, layers, m.embed.cpu().numpy(), m.lm_head.cpu().numpy(), m.cos_sin.cpu().numpy(),
                           max_batch)

    # -- linear in either mode ----------------------------------------------------------------
    def _linear(self, w, x_or_packed, w4a4):
        wp, sw = w
        if w4a4:
            q, s = x_or_packed
            return O.gemm_w4a4(q, s, wp, sw)
        return O.gemm_w4a16(x_or_packed, wp, sw)

    def _rope_cache(self, li, qkv, positions, seq_ids):
        cfg = self.cfg
        Hq, Hkv, D = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
        T = qkv.shape[0]
        x = qkv.reshape(T, Hq + 2 * Hkv, D).astype(F32)
        cs = self.cos_sin[positions].astype(F32)                       # [T, D]
        c, s = cs[:, None, : D // 2], cs[:, None, D // 2:]
        rot = x[:, : Hq + Hkv]
        x1, x2 = rot[..., : D // 2], rot[..., D // 2:]
        a = ((x1 * c).astype(F32) - (x2 * s).astype(F32)).astype(F32)  # same op order as the kernel
        b = ((x2 * c).astype(F32) + (x1 * s).astype(F32)).astype(F32)
        rot16 = np.concatenate((a, b), axis=-1).astype(F16)
        q = rot16[:, :Hq]
        k = rot16[:, Hq:]
        v = qkv.reshape(T, Hq + 2 * Hkv, D)[:, Hq + Hkv:]
        for t in range(T):
            self.ck[li, seq_ids[t], positions[t]] = k[t]
            self.cv[li, seq_ids[t], positions[t]] = v[t]
        return q

    def _attention(self, li, q, positions, seq_ids):
        cfg = self.cfg
        Hq, Hkv, D = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
        T = q.shape[0]
        out = np.zeros((T, Hq, D), dtype=F16)
        scale = F32(1.0 / np.sqrt(D))
        for t in range(T):
            n = positions[t] + 1
            K = self.ck[li, seq_ids[t], :n].astype(F32)               # [n, Hkv, D]
            V = self.cv[li, seq_ids[t], :n].astype(F32)
            for h in range(Hq):
                hk = h // (Hq // Hkv)
                sc = K[:, hk] @ (q[t, h].astype(F32) * scale)
                p = np.exp(sc - sc.max())
                out[t, h] = ((p @ V[:, hk]) / p.sum()).astype(F16)
        return out.reshape(T, Hq * D)

    def forward(self, input_ids, positions, seq_ids, w4a4):
        cfg = self.cfg
        H, inter = cfg.hidden_size, cfg.intermediate_size
        Hq, D = cfg.num_attention_heads, cfg.head_dim
        x = self.embed[input_ids].astype(F16)
        sc_h = float(F32(1.0) / np.sqrt(F32(Hq)))
        sc_m = float(F32(1.0) / np.sqrt(F32(inter)))
        for li, lw in enumerate(self.layers):
            if w4a4:
                q, s, _ = O.norm_quant_i4(x, cfg.rms_norm_eps)
                qkv = self._linear(lw["qkv"], (q, s), True)
            else:
                qkv = self._linear(lw["qkv"], O.norm_f16(x, cfg.rms_norm_eps), False)
            qr = self._rope_cache(li, qkv, positions, seq_ids)
            attn = self._attention(li, qr, positions, seq_ids)
            rot = O.hadamard_heads(attn, Hq, D, sc_h)
            o = self._linear(lw["o"], O.sym_quant_i4(rot) if w4a4 else rot, w4a4)
            x = (x.astype(F32) + o.astype(F32)).astype(F16)
            if w4a4:
                q, s, _ = O.norm_quant_i4(x, cfg.rms_norm_eps)
                gu = self._linear(lw["gate_up"], (q, s), True)
            else:
                gu = self._linear(lw["gate_up"], O.norm_f16(x, cfg.rms_norm_eps), False)
            act = O.silu_mul(gu, inter)
            rot = O.hadamard_k28(act, inter // 28, sc_m)
            dn = self._linear(lw["down"], O.sym_quant_i4(rot) if w4a4 else rot, w4a4)
            x = (x.astype(F32) + dn.astype(F32)).astype(F16)
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
