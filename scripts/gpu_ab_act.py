"""Same-box timings of the activation-side kernels at large-batch sizes: python scripts/gpu_ab_act.py <tag>
(QSPEC_K28_CLUSTER_ROWS selects the down-rotation form). us per launch, CUDA graph of 32 launches."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from qspec_b200 import ops
from qspec_b200.model import build_cos_sin_cache
from gpu_microbench import time_graph
tag = sys.argv[1] if len(sys.argv) > 1 else "?"
dev = "cuda"
for M in (64, 256):
    gu = torch.randn(M, 28672, device=dev).half()
    qm = torch.empty(M, 7168, dtype=torch.int8, device=dev)
    hm = torch.empty(M, 14336, dtype=torch.float16, device=dev)
    s = torch.empty(M, dtype=torch.float16, device=dev)
    t_q = time_graph([lambda: ops.hadamard_k28(None, 512, 0.00835, gate_up=gu, out_q=qm, scale_out=s)] * 32)[1]
    t_f = time_graph([lambda: ops.hadamard_k28(None, 512, 0.00835, gate_up=gu, out=hm)] * 32)[1]
    print(json.dumps({"tag": tag, "kernel": "silu+hadamard_k28", "M": M, "quant_us": round(t_q, 2), "f16_us": round(t_f, 2)}), flush=True)
# attention: 64 sequences, ~200 cached positions, 32 layers' caches rotated (> L2)
Hq, Hkv, D, S, B, L = 32, 8, 128, 428, 64, 8
ck = [torch.randn(B, S, Hkv, D, device=dev).half() for _ in range(L)]
cv = [torch.randn(B, S, Hkv, D, device=dev).half() for _ in range(L)]
cs = build_cos_sin_cache(D, S, 500000.0, dev)
for per_seq in (1, 4):
    T = B * per_seq
    seq = torch.arange(B, dtype=torch.int32, device=dev).repeat_interleave(per_seq).contiguous()
    pos = (200 + torch.arange(per_seq, dtype=torch.int32, device=dev)).repeat(B).contiguous()
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, device=dev).half()
    out = torch.empty(T, Hq * D, dtype=torch.float16, device=dev)
    q = torch.empty(T, Hq * D // 2, dtype=torch.int8, device=dev)
    sc = torch.empty(T, dtype=torch.float16, device=dev)
    sm, hs = 1.0 / np.sqrt(D), 1.0 / np.sqrt(Hq)
    row = {"tag": tag, "kernel": "attention", "T": T, "per_seq": per_seq}
    row["seq_attention_us"] = round(time_graph([lambda l=l: ops.seq_attention(qkv, ck[l], cv[l], pos, seq, out, Hq, Hkv, D, sm) for l in range(L)] * 4)[1], 2)
    row["decode_attention_us"] = round(time_graph([lambda l=l: ops.decode_attention(qkv, ck[l], cv[l], pos, seq, out, Hq, Hkv, D, sm) for l in range(L)] * 4)[1], 2)
    row["rope_kvcache_us"] = round(time_graph([lambda l=l: ops.rope_kvcache(qkv, cs, pos, seq, ck[l], cv[l], Hq, Hkv, D) for l in range(L)] * 4)[1], 2)
    row["hadamard_heads_quant_us"] = round(time_graph([lambda: ops.hadamard_heads(out, Hq, D, hs, out_q=q, scale_out=sc)] * 32)[1], 2)
    if ops.attn_block_supported(T, Hq, Hkv, D, S):
        row["attn_block_us"] = round(time_graph([lambda l=l: ops.attn_block(qkv, cs, pos, seq, ck[l], cv[l], Hq, Hkv, D, sm, hs, out_q=q, scale_out=sc) for l in range(L)] * 4)[1], 2)
    print(json.dumps(row), flush=True)
