"""Times the attention block (rope + cache append + attention + rotation over heads + quantiser) as the chain of
separate kernels and as the one-launch qspec_attn_block, graph-replayed, at the bench's decode shapes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from qspec_b200 import ops
from qspec_b200.model import build_cos_sin_cache

dev = torch.device("cuda")
Hq, Hkv, D, S, B = 32, 8, 128, 1024, 16
L = 32                                   # distinct caches so that the cache rows come from HBM/L2 like in the model
for T, per_seq, pos0 in ((4, 1, 160), (16, 4, 160), (4, 1, 300), (4, 1, 600)):
    g = torch.Generator().manual_seed(1)
    ck = torch.randn(L, B, S, Hkv, D, generator=g).half().to(dev)
    cv = torch.randn(L, B, S, Hkv, D, generator=g).half().to(dev)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).half().to(dev)
    cos_sin = build_cos_sin_cache(D, S, 500000.0, dev)
    seq = torch.arange(T, dtype=torch.int32, device=dev) // per_seq
    pos = (pos0 + torch.arange(T, dtype=torch.int32, device=dev) % per_seq).int()
    attn = torch.empty(T, Hq * D, dtype=torch.float16, device=dev)
    q = torch.empty(T, Hq * D // 2, dtype=torch.int8, device=dev)
    sc = torch.empty(T, dtype=torch.float16, device=dev)
    sm, hs = 1.0 / np.sqrt(D), 1.0 / np.sqrt(Hq)

    def chain():
        for l in range(L):
            ops.rope_kvcache(qkv, cos_sin, pos, seq, ck[l], cv[l], Hq, Hkv, D)
            ops.decode_attention(qkv, ck[l], cv[l], pos, seq, attn, Hq, Hkv, D, sm)
            ops.hadamard_heads(attn, Hq, D, hs, out_q=q, scale_out=sc)

    def fused():
        for l in range(L):
            ops.attn_block(qkv, cos_sin, pos, seq, ck[l], cv[l], Hq, Hkv, D, sm, hs, out_q=q, scale_out=sc)

    for name, fn in (("chain", chain), ("fused", fused)):
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            fn(); torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                fn()
        for _ in range(3):
            gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            gr.replay()
        e1.record(); torch.cuda.synchronize()
        print(f"T={T} pos={pos0} {name}: {e0.elapsed_time(e1) * 1000 / (10 * L):.2f} us per layer", flush=True)
