"""QSpec Llama decoder built from the dual-mode layer modules -- the drop-in consumer of the hot
path, wired exactly like the reference model (vllm/model_executor/models/quarot_llama.py:
attention :175-244, MLP :263-299, decoder layer :363-392, model loop :500-533) but with the
per-layer launches fused (norm+quant, rotation+quant, SiLU*up+rotation+quant) and every
scratch buffer preallocated once (the reference allocates 11 tensors per draft step,
spec_decode/draft_model_runner.py:277-320).

One `nn.Module` instance and ONE packed int4 weight per linear serve both modes; the mode is
the `w4a4` flag of `forward`, the reference's ExecuteModelRequest.w4a4 -> kwargs["w4a4"]
switch (spec_decode_worker.py:797-812, worker_base.py:416).

Tensor parallelism: column-parallel qkv / gate_up, row-parallel o / down. The all-reduce of
RowParallelLinear.forward (layers/linear.py:1143-1144) is fused into the norm that follows (peer-mapped partial
sums, `parallel.PeerPartials`; QSPEC_TP_FUSED=0 restores one NCCL all-reduce per row-parallel GEMM); online
rotations are shard-local.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Optional

import torch

from . import ops
from .nn import Linear4bit, PackedQuantizedTensor, pack_i4, reference_hadamard_scale


@dataclass
class QSpecLlamaConfig:
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    head_dim: int = 128
    num_hidden_layers: int = 32
    vocab_size: int = 128256
    rms_norm_eps: float = 1e-5
    rope_theta: float = 500000.0
    max_seq_len: int = 1024          # KV-cache length per sequence
    name: str = "llama-3-8b"
    # Qwen2.5-14B variant of the path (vllm/model_executor/models/qwen2_quarot.py):
    qkv_bias: bool = False           # bias in qkv_proj (:172-178)
    had_k: int = 28                  # down_proj rotation had_k (x) H_{inter/had_k}: 28 (Llama-3) or 108 (13824 = 108 * 128)
    heads_rotation: str = "sylvester"   # o_proj rotation across heads: Sylvester FHT (2^k heads) or "had40" (:269-277)
    qk_hadamard: bool = False        # per-head FHT_128 of q and k after the rotary embedding (:279-288)
    gate_first: bool = False         # merged gate_up rows [gate ; up] (vLLM SiluAndMul, :80-86) instead of Llama's [up ; gate]

    @staticmethod
    def llama3_8b(**kw):
        return QSpecLlamaConfig(**kw)

    @staticmethod
    def llama3_70b(**kw):
        return QSpecLlamaConfig(hidden_size=8192, intermediate_size=28672, num_attention_heads=64,
                                num_key_value_heads=8, num_hidden_layers=80, name="llama-3-70b", **kw)

    @staticmethod
    def qwen25_14b(**kw):
        d = dict(hidden_size=5120, intermediate_size=13824, num_attention_heads=40, num_key_value_heads=8,
                 num_hidden_layers=48, vocab_size=152064, rms_norm_eps=1e-6, rope_theta=1000000.0, name="qwen2.5-14b",
                 qkv_bias=True, had_k=108, heads_rotation="had40", qk_hadamard=True, gate_first=True)
        d.update(kw)
        return QSpecLlamaConfig(**d)

    @staticmethod
    def tiny(**kw):
        d = dict(hidden_size=512, intermediate_size=28 * 64, num_attention_heads=4, num_key_value_heads=2,
                 num_hidden_layers=2, vocab_size=1024, max_seq_len=128, name="tiny")
        d.update(kw)
        return QSpecLlamaConfig(**d)

    def linear_shapes(self, tp: int = 1):
        """(N, K) per rank of the four linears (qkv, o, gate_up, down)."""
        D = self.head_dim
        hq, hkv = self.num_attention_heads // tp, max(self.num_key_value_heads // tp, 1)
        return {"qkv": ((hq + 2 * hkv) * D, self.hidden_size), "o": (self.hidden_size, hq * D),
                "gate_up": (2 * self.intermediate_size // tp, self.hidden_size),
                "down": (self.hidden_size, self.intermediate_size // tp)}


def random_int4_linear(N: int, K: int, seed: int, device, bias: bool = False) -> Linear4bit:
    """Random-init packed weights + scales as SURVEY 8(d): q ~ U{-8..7}, sW = (0.5+U)*2^-7 (bias ~ N(0, 0.1^2))."""
    device = torch.device(device)
    g = torch.Generator(device=device).manual_seed(seed)
    lin = Linear4bit(K, N, bias=False, dtype=torch.float16, device=device, empty_init=True)
    if bias:
        lin.bias = (torch.randn(N, generator=g, device=device) * 0.1).to(torch.float16)
    # generate packed bytes directly (uniform bytes == uniform nibble pairs), avoids an [N,K] int8 temp
    lin.weight = torch.randint(0, 256, (N, K // 2), generator=g, dtype=torch.uint8, device=device).view(torch.int8)
    lin.weight_scales = ((0.5 + torch.rand(N, generator=g, device=device)) * 2.0 ** -7).to(torch.float16)
    return lin


def build_cos_sin_cache(head_dim: int, max_pos: int, theta: float, device) -> torch.Tensor:
    """vllm RotaryEmbedding._compute_cos_sin_cache: [max_pos, head_dim] = [cos | sin], model dtype."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
    t = torch.arange(max_pos, dtype=torch.float32)
    fr = torch.outer(t, inv)
    return torch.cat((fr.cos(), fr.sin()), dim=-1).to(torch.float16).to(device)


class QSpecDecoderLayer(torch.nn.Module):
    def __init__(self, cfg: QSpecLlamaConfig, layer_idx: int, device, tp_size=1, tp_rank=0, seed=0):
        super().__init__()
        self.cfg = cfg
        sh = cfg.linear_shapes(tp_size)
        base = seed * 100003 + layer_idx * 16 + tp_rank * 4
        self.qkv_proj = random_int4_linear(*sh["qkv"], 1000 + base, device, bias=cfg.qkv_bias)
        self.o_proj = random_int4_linear(*sh["o"], 1001 + base, device)
        self.gate_up = random_int4_linear(*sh["gate_up"], 1002 + base, device)    # [up ; gate], quarot_llama.py:304-305
        self.down_proj = random_int4_linear(*sh["down"], 1003 + base, device)


class QSpecLlamaModel(torch.nn.Module):
    """Random-init Llama-3 architecture with QSpec dual-mode linears (no checkpoints offline)."""

    def __init__(self, cfg: QSpecLlamaConfig, max_batch: int, max_tokens: int, device="cuda",
                 tp_group=None, seed: int = 0, tp_size: Optional[int] = None, tp_rank: int = 0, peers=None):
        """tp_group: the tensor-parallel process group (one rank per GPU). tp_size / tp_rank WITHOUT a group build shard
        `tp_rank` of a TP=`tp_size` model in this process (several shards may share one device; `peers` must then be
        supplied or attached before `forward`, and the shards advanced in lock step through `forward_stages`)."""
        super().__init__()
        self.cfg = cfg
        self.device = torch.device(device)
        self.tp_group = tp_group
        if tp_group is not None:
            import torch.distributed as dist
            self.tp_size, self.tp_rank = dist.get_world_size(tp_group), dist.get_rank(tp_group)
        elif tp_size is not None:
            self.tp_size, self.tp_rank = int(tp_size), int(tp_rank)
        else:
            self.tp_size, self.tp_rank = 1, 0
        tp = self.tp_size
        assert cfg.num_attention_heads % tp == 0 and cfg.num_key_value_heads % tp == 0 and cfg.intermediate_size % (cfg.had_k * tp) == 0
        assert cfg.had_k in (28, 108) and cfg.heads_rotation in ("sylvester", "had40")
        self.qwen = cfg.had_k != 28 or cfg.heads_rotation != "sylvester" or cfg.qk_hadamard or cfg.gate_first
        if self.qwen and tp != 1:
            raise ValueError("the Qwen2.5-14B variant is TP=1 only (had108 / had40 do not shard; the reference runs it unsharded)")
        self.hq = cfg.num_attention_heads // tp
        self.hkv = cfg.num_key_value_heads // tp
        self.inter = cfg.intermediate_size // tp
        self.had_n = self.inter // cfg.had_k
        D = cfg.head_dim
        self.layers = torch.nn.ModuleList(
            [QSpecDecoderLayer(cfg, i, self.device, tp, self.tp_rank, seed) for i in range(cfg.num_hidden_layers)])
        g = torch.Generator(device="cpu").manual_seed(seed + 7)
        self.embed = (torch.randn(cfg.vocab_size, cfg.hidden_size, generator=g) * 0.02).to(torch.float16).to(self.device)
        vs = cfg.vocab_size // tp                                  # vocab-sharded lm_head
        g2 = torch.Generator(device="cpu").manual_seed(seed + 11 + self.tp_rank)
        self.lm_head = (torch.randn(vs, cfg.hidden_size, generator=g2) * 0.02).to(torch.float16).to(self.device)
        self.vocab_shard = vs
        self.cos_sin = build_cos_sin_cache(D, cfg.max_seq_len, cfg.rope_theta, self.device)
        # shard-local rotation scales (SURVEY 7 hard part 5): 1/sqrt(local heads), 1/sqrt(local inter), evaluated the way the
        # reference's run time evaluates them (float16: nn.reference_hadamard_scale)
        self.scale_heads = reference_hadamard_scale(self.hq)
        if cfg.heads_rotation == "had40":
            self.scale_heads = 1.0 / math.sqrt(self.hq)       # o_had_scale is a python float (qwen2_quarot.py:217), cast to fp32 at the multiply
        self.scale_mlp = reference_hadamard_scale(self.inter)
        self.sm_scale = 1.0 / math.sqrt(D)
        # KV cache shared by draft and verify (worker.py:309-326 ref_initilize_cache)
        L, S = cfg.num_hidden_layers, cfg.max_seq_len
        self.cache_k = torch.zeros(L, max_batch, S, self.hkv, D, dtype=torch.float16, device=self.device)
        self.cache_v = torch.zeros_like(self.cache_k)
        # scratch arena, allocated once
        T, H = max_tokens, cfg.hidden_size
        f16 = dict(dtype=torch.float16, device=self.device)
        i8 = dict(dtype=torch.int8, device=self.device)
        self.max_tokens = T
        self.buf_x = torch.empty(T, H, **f16)
        self.buf_norm = torch.empty(T, H, **f16)
        self.buf_qh = torch.empty(T, H // 2, **i8)                        # quantized_buffer_qkv (norm output)
        self.buf_qa = torch.empty(T, self.hq * D // 2, **i8)              # quantized o_proj input (local heads)
        self.buf_qm = torch.empty(T, self.inter // 2, **i8)               # quantized_buffer_mlp
        # the same three as int8 containers (one int4 value per byte) for passes of >= QSPEC_C8_MIN_T tokens: the GEMM then
        # takes the activations as TMA delivers them instead of widening the tile again for every weight tile
        self.buf_qh8 = torch.empty(T, H, **i8)
        self.buf_qa8 = torch.empty(T, self.hq * D, **i8)
        self.buf_qm8 = torch.empty(T, self.inter, **i8)
        self.buf_scale = torch.empty(T, **f16)                            # scale_buffer
        self.buf_isum = torch.empty(T, **f16)                             # input_sum_buffer
        self.buf_qkv = torch.empty(T, (self.hq + 2 * self.hkv) * D, **f16)  # act_buffer_qkv
        self.buf_attn = torch.empty(T, self.hq * D, **f16)                # act_buffer_attn
        self.buf_had = torch.empty(T, self.hq * D, **f16)                 # act_buffer_had
        self.buf_out = torch.empty(T, H, **f16)                           # act_buffer_output
        self.buf_gu = torch.empty(T, 2 * self.inter, **f16)               # act_buffer_gate_up
        self.buf_hmlp = torch.empty(T, self.inter, **f16)                 # act_buffer_had_mlp
        ops.reserve_workspaces(T, list(cfg.linear_shapes(tp).values()), self.device)
        # TP=1: o_proj / down_proj leave their k-split partial sums to the norm that follows (deferred reduction);
        # QSPEC_SPLIT_FUSED=0 keeps the reduction inside the GEMM.
        self.split_ok = tp == 1 and os.environ.get("QSPEC_SPLIT_FUSED", "1") != "0"
        self.buf_part = torch.empty(2, 8 * T * H, dtype=torch.int32, device=self.device) if self.split_ok else None
        # separate scale buffer for the o_proj / down_proj inputs: the consuming norm reads it while writing buf_scale
        self.buf_scale2 = torch.empty(T, **f16)
        # TP>1: the row-parallel partial sums live in peer-mapped memory and are summed inside the next norm
        # (qspec_tp_add_norm_*); QSPEC_TP_FUSED=0 keeps one NCCL all-reduce per row-parallel GEMM instead.
        self.peers = peers
        if peers is None and tp > 1 and tp_group is not None and os.environ.get("QSPEC_TP_FUSED", "1") != "0":
            from .parallel import PeerPartials
            self.peers = PeerPartials(T, H, torch.float16, self.device, tp_group)

    # ------------------------------------------------------------------------------------------
    def _allreduce(self, t: torch.Tensor):
        if self.tp_size > 1 and self.peers is None:
            if self.tp_group is None:
                raise RuntimeError("TP shard without a process group needs peer-mapped partial buffers (peers=...)")
            import torch.distributed as dist
            dist.all_reduce(t, group=self.tp_group)     # RowParallelLinear.forward, layers/linear.py:1143-1144

    def forward(self, input_ids: torch.Tensor, positions: torch.Tensor, seq_ids: torch.Tensor, w4a4: bool) -> torch.Tensor:
        """input_ids/positions/seq_ids: [T] int (int32 positions/seq_ids). Returns final-norm hidden [T, H]."""
        for _ in self.forward_stages(input_ids, positions, seq_ids, w4a4):
            pass
        return self.buf_norm[: input_ids.numel()]

    def forward_stages(self, input_ids: torch.Tensor, positions: torch.Tensor, seq_ids: torch.Tensor, w4a4: bool):
        """`forward` as a generator that yields after every row-parallel GEMM (the points where the ranks of a
        tensor-parallel group exchange partial sums): lets several shards that live in ONE process be advanced in lock
        step (tests: all ranks of a TP model on one GPU). The final-norm hidden state is left in `buf_norm[:T]`."""
        cfg = self.cfg
        T = input_ids.numel()
        assert T <= self.max_tokens
        D, H = cfg.head_dim, cfg.hidden_size
        x = self.buf_x[:T]
        torch.index_select(self.embed, 0, input_ids, out=x)
        nrm, qh, qa, qm = self.buf_norm[:T], self.buf_qh[:T], self.buf_qa[:T], self.buf_qm[:T]
        sc, isum = self.buf_scale[:T], self.buf_isum[:T]
        qkv, attn, had, out = self.buf_qkv[:T], self.buf_attn[:T], self.buf_had[:T], self.buf_out[:T]
        gu, hmlp = self.buf_gu[:T], self.buf_hmlp[:T]
        # attention-block and MLP-block outputs: one local buffer, or the two peer-mapped partial buffers
        peers = self.peers
        out_a, out_m = (peers.partial(0, T), peers.partial(1, T)) if peers is not None else (out, out)
        # The residual add that closes a block is fused into the norm that opens the next one (`pending`):
        # x <- x + out happens inside the norm kernel, bit-identical to ops.add followed by the norm.
        pending = False
        sc2 = self.buf_scale2[:T]
        # int8 container for the draft path's quantised activations (same values, same results; include/qspec_b200.h "_c8")
        # (not with the one-launch attention block, which writes the packed form, nor with the two-step peer reduction)
        attn_block_taken = (os.environ.get("QSPEC_ATTN_FUSED", "1") != "0" and not self.qwen
                            and T <= int(os.environ.get("QSPEC_ATTN_BLOCK_MAX_T", "32"))
                            and ops.attn_block_supported(T, self.hq, self.hkv, D, self.cache_k.shape[2]))
        c8 = (w4a4 and not self.qwen and T >= int(os.environ.get("QSPEC_C8_MIN_T", "48")) and not attn_block_taken
              and (self.split_ok or (peers is not None and not ops.tp_two_step(peers, T))))
        qh8, qa8, qm8 = self.buf_qh8[:T], self.buf_qa8[:T], self.buf_qm8[:T]

        def can_split(kind):                                     # deferred reduction possible for o_proj and down_proj at this T
            return self.split_ok and all(ops.gemm_split_factor(kind, T, n, k) in range(1, 9)
                                         for n, k in (cfg.linear_shapes(1)["o"], cfg.linear_shapes(1)["down"]))

        split = can_split(2) if c8 else False
        c8 = c8 and (split or peers is not None)
        if not c8:
            split = can_split(0 if w4a4 else 1)
        kw8 = {"w4a8": True} if c8 else {"w4a4": True}           # Linear4bit route of a draft-mode GEMM
        sp_a = sp_m = None                                                  # pending partial sums (attention / MLP block)
        # one-launch attention block for decode-sized passes; QSPEC_ATTN_BLOCK_MAX_T caps the token count it takes (larger
        # passes go through rope_kvcache + seq_attention + hadamard_heads)
        fused_attn = (os.environ.get("QSPEC_ATTN_FUSED", "1") != "0" and not self.qwen
                      and T <= int(os.environ.get("QSPEC_ATTN_BLOCK_MAX_T", "32"))
                      and ops.attn_block_supported(T, self.hq, self.hkv, D, self.cache_k.shape[2]))
        rot_heads = ops.hadamard_heads40 if cfg.heads_rotation == "had40" else ops.hadamard_heads
        # chain path (passes the one-launch block does not take): many tokens -> one CTA per (sequence, kv head)
        seq_attn = (not fused_attn and T >= 32 and os.environ.get("QSPEC_SEQ_ATTN", "1") != "0"
                    and ops.seq_attention_supported(T, self.hq, self.hkv, D, self.cache_k.shape[2]))
        # Single-GPU emulation of the two-step fused all-reduce (tests/tp_emulation.py): the rows a rank owns are reduced and
        # pushed by one launch, the other rows are consumed by a second one, with every rank's first launch before any rank's
        # second (the product does both in ONE launch per rank; kernels that wait on one another must not be queued on one GPU).
        emulate_rs = peers is not None and getattr(peers, "emulate_two_step", False) and ops.tp_two_step(peers, T)

        def staged(has_delta, call):
            if emulate_rs and has_delta:
                self._tp_phases = 1
                call()
                yield ("reduce", None)
                self._tp_phases = 2
            call()
            self._tp_phases = 3

        for li, layer in enumerate(self.layers):
            ck, cv = self.cache_k[li], self.cache_v[li]
            # ---- attention block (quarot_llama.py:374-380)
            if w4a4:
                first8 = c8 and pending                     # the very first norm of a pass has no producer GEMM: packed
                yield from staged(pending, lambda: self._norm_quant(x, qh8 if first8 else qh, sc, isum, out_m if pending else None,
                                                                    1, sp_m, c8=first8))
                if first8:
                    layer.qkv_proj(PackedQuantizedTensor(qh8, sc), qkv, w4a8=True)
                else:
                    layer.qkv_proj(PackedQuantizedTensor(qh, sc), qkv, w4a4=True)
            else:
                yield from staged(pending, lambda: self._norm_f16(nrm, x, out_m if pending else None, 1, sp_m))
                layer.qkv_proj(nrm, qkv)
            if fused_attn:
                # rope + cache append + attention + rotation over heads (+ quantiser): one launch per layer
                if w4a4:
                    ops.attn_block(qkv, self.cos_sin, positions, seq_ids, ck, cv, self.hq, self.hkv, D, self.sm_scale,
                                   self.scale_heads, out_q=qa, scale_out=sc2)
                else:
                    ops.attn_block(qkv, self.cos_sin, positions, seq_ids, ck, cv, self.hq, self.hkv, D, self.sm_scale,
                                   self.scale_heads, out=had)
            else:
                if cfg.qk_hadamard:
                    ops.rope_qkhad_kvcache(qkv, self.cos_sin, positions, seq_ids, ck, cv, self.hq, self.hkv, D,
                                           1.0 / math.sqrt(D))            # head_had_scale, qwen2_quarot.py:219
                else:
                    ops.rope_kvcache(qkv, self.cos_sin, positions, seq_ids, ck, cv, self.hq, self.hkv, D)
                if seq_attn:
                    ops.seq_attention(qkv, ck, cv, positions, seq_ids, attn, self.hq, self.hkv, D, self.sm_scale)
                else:
                    ops.decode_attention(qkv, ck, cv, positions, seq_ids, attn, self.hq, self.hkv, D, self.sm_scale)
                if w4a4 and c8:
                    rot_heads(attn, self.hq, D, self.scale_heads, out_q=qa8, scale_out=sc2, c8=True)
                elif w4a4:
                    rot_heads(attn, self.hq, D, self.scale_heads, out_q=qa, scale_out=sc2)
                else:
                    rot_heads(attn, self.hq, D, self.scale_heads, out=had)
            o_in = PackedQuantizedTensor(qa8 if c8 else qa, sc2) if w4a4 else had
            if split:
                sp_a = layer.o_proj.forward_split(o_in, self.buf_part[0], w4a4=w4a4)
            elif w4a4:
                layer.o_proj(o_in, out_a, **kw8)
            else:
                layer.o_proj(o_in, out_a)
            self._allreduce(out_a)
            yield ("o_proj", li)
            # ---- MLP block (quarot_llama.py:380,387-389): residual add fused into the norm
            if w4a4:
                yield from staged(True, lambda: self._norm_quant(x, qh8 if c8 else qh, sc, isum, out_a, 0, sp_a, c8=c8))
                layer.gate_up(PackedQuantizedTensor(qh8 if c8 else qh, sc), gu, **kw8)
                if cfg.had_k == 108:
                    ops.hadamard_k108(None, self.had_n, self.scale_mlp, gate_up=gu, gate_first=cfg.gate_first, out_q=qm, scale_out=sc2)
                else:
                    ops.hadamard_k28(None, self.had_n, self.scale_mlp, gate_up=gu, out_q=qm8 if c8 else qm, scale_out=sc2, c8=c8)
                d_in = PackedQuantizedTensor(qm8 if c8 else qm, sc2)
            else:
                yield from staged(True, lambda: self._norm_f16(nrm, x, out_a, 0, sp_a))
                layer.gate_up(nrm, gu)
                if cfg.had_k == 108:
                    ops.hadamard_k108(None, self.had_n, self.scale_mlp, gate_up=gu, gate_first=cfg.gate_first, out=hmlp)
                else:
                    ops.hadamard_k28(None, self.had_n, self.scale_mlp, gate_up=gu, out=hmlp)
                d_in = hmlp
            if split:
                sp_m = layer.down_proj.forward_split(d_in, self.buf_part[1], w4a4=w4a4)
            elif w4a4:
                layer.down_proj(d_in, out_m, **kw8)
            else:
                layer.down_proj(d_in, out_m)
            self._allreduce(out_m)
            yield ("down_proj", li)
            pending = True
        yield from staged(pending, lambda: self._norm_f16(nrm, x, out_m if pending else None, 1, sp_m))   # final norm is always fp16 (:533)

    def _norm_f16(self, nrm, x, delta, which=0, sp=None):
        if delta is None:
            ops.rms_norm_general_fuse_sum_fp16(nrm, x, self.cfg.rms_norm_eps)
        elif sp is not None:
            ops.split_add_rms_norm_fuse_sum_fp16(nrm, x, sp, self.cfg.rms_norm_eps)
        elif self.peers is not None:
            ops.tp_add_rms_norm_fuse_sum_fp16(nrm, x, self.peers, which, self.cfg.rms_norm_eps, phases=getattr(self, "_tp_phases", 3))
        else:
            ops.add_rms_norm_fuse_sum_fp16(nrm, x, delta, self.cfg.rms_norm_eps)

    def _norm_quant(self, x, q_buf, sc, isum, delta=None, which=0, sp=None, c8=False):
        """c8: q_buf is the int8 container [T, H] (only the producers fed by a deferred or peer partial sum have that form)."""
        assert not c8 or (delta is not None and (sp is not None or self.peers is not None))
        if delta is None:
            ops.rms_norm_general_fuse_sum_i4(q_buf, x, isum, sc, self.cfg.rms_norm_eps, True)
        elif sp is not None:
            ops.split_add_rms_norm_fuse_sum_i4(q_buf, x, sp, isum, sc, self.cfg.rms_norm_eps, c8=c8)
        elif self.peers is not None:
            ops.tp_add_rms_norm_fuse_sum_i4(q_buf, x, self.peers, which, isum, sc, self.cfg.rms_norm_eps,
                                            phases=getattr(self, "_tp_phases", 3), c8=c8)
        else:
            ops.add_rms_norm_fuse_sum_i4(q_buf, x, delta, isum, sc, self.cfg.rms_norm_eps)
        return PackedQuantizedTensor(q_buf, sc)

    def make_modes_agree(self, shift: Optional[int]) -> None:
        """Test / bench utility: scales the embedding table by 2^shift (relative to its initial values; None restores
        them) so that the residual stream -- hence every greedy arg max -- is dominated by the embedding of the current
        token, which both modes share. Draft (W4A4) and verify (W4A16) then agree by construction and the speculative
        step runs at a high acceptance rate (the reference checkpoint's operating point is 0.955-0.961,
        figs/image-1.png) with random-init weights. The kernels do exactly the same work; only the values change.
        In place (a captured CUDA graph keeps reading the same table). Apply the same shift on every rank of a TP group."""
        if not hasattr(self, "_embed_base"):
            self._embed_base = self.embed.clone()
        if shift is None:
            self.embed.copy_(self._embed_base)
        else:
            self.embed.copy_((self._embed_base.float() * float(2.0 ** shift)).clamp_(-65504.0, 65504.0).to(torch.float16))

    def greedy_tokens(self, hidden: torch.Tensor) -> torch.Tensor:
        """argmax over the (vocab-sharded) fp16 lm_head logits; returns int64 [T]."""
        logits = torch.matmul(hidden, self.lm_head.t())                    # library GEMM, as the reference's lm_head
        if self.tp_size == 1:
            return ops.argmax_rows(logits)
        val = torch.empty(logits.shape[0], dtype=torch.float32, device=logits.device)
        idx = ops.argmax_rows(logits, out_val=val)
        if self.peers is not None and hasattr(self.peers, "_amax_slots"):
            # pairs pushed into every peer's table over NVLink, best of the own table: one launch, no collective call
            return ops.tp_argmax_combine(val, idx, self.tp_rank * self.vocab_shard, self.peers)
        from .parallel import combine_sharded_argmax            # QSPEC_TP_FUSED=0: NCCL all-gather
        return combine_sharded_argmax(val, idx, self.tp_rank, self.vocab_shard, self.tp_group)
