"""Runs the small kernels of one decode layer a few times at the bench shapes (for ncu captures)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from qspec_b200 import ops
from qspec_b200.model import build_cos_sin_cache

dev = torch.device("cuda")
Hq, Hkv, D, S, B, T, H, I = 32, 8, 128, 512, 4, 4, 4096, 14336
g = torch.Generator().manual_seed(1)
ck = torch.randn(B, S, Hkv, D, generator=g).half().to(dev)
cv = torch.randn(B, S, Hkv, D, generator=g).half().to(dev)
qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).half().to(dev)
cos_sin = build_cos_sin_cache(D, S, 500000.0, dev)
seq = torch.arange(T, dtype=torch.int32, device=dev)
pos = torch.full((T,), 160, dtype=torch.int32, device=dev)
q = torch.empty(T, Hq * D // 2, dtype=torch.int8, device=dev)
sc = torch.empty(T, dtype=torch.float16, device=dev)
isum = torch.empty_like(sc)
x = torch.randn(T, H, generator=g).half().to(dev)
delta = torch.randn(T, H, generator=g).half().to(dev)
qh = torch.empty(T, H // 2, dtype=torch.int8, device=dev)
gu = torch.randn(T, 2 * I, generator=g).half().to(dev)
qm = torch.empty(T, I // 2, dtype=torch.int8, device=dev)
for _ in range(3):
    ops.attn_block(qkv, cos_sin, pos, seq, ck, cv, Hq, Hkv, D, 1 / np.sqrt(D), 1 / np.sqrt(Hq), out_q=q, scale_out=sc)
    ops.add_rms_norm_fuse_sum_i4(qh, x, delta, isum, sc, 1e-5)
    ops.hadamard_k28(None, I // 28, 1 / np.sqrt(I), gate_up=gu, out_q=qm, scale_out=sc)
torch.cuda.synchronize()
print("ok")
