"""Timed CPU baseline: the reference's own torch formulation of the dual-mode linear path,
run on the host cores (TEST/BENCH INFRASTRUCTURE ONLY -- never imported by qspec_b200/).

What is timed:
  W4A16  y = x.float() @ (unpack_i4(W).float() * sW.float()[:, None]).T   -- unpack + dequant every
         call, as the reference's torch statement does (quarot_nn/linear.py:111-119;
         unpack_i4 third-party/QuaRot/quarot/functional/quantization.py:52-74)
  W4A4   per-token scale = amax/7, q = clamp(round(x/scale), -8, 7)
         (third-party/QuaRot/fake_quant/quant_utils.py:30-38,158-169), then
         ((q.float() @ wq.float().T) * scale[:, None] * sW[None, :])
         (third-party/ao/test/test_rowwise_scaled_linear_cutlass.py:68-76)
plus the mean-subtracting norm and the online Hadamard rotations in torch fp32
(matmul_hadU, third-party/QuaRot/quarot/functional/hadamard.py:59-81).
"""
from __future__ import annotations

import os
import time

import torch


def unpack_i4(x: torch.Tensor) -> torch.Tensor:
    """quantization.py:52-74 (uint8 -> int32, interleaving low/high nibbles)."""
    x = x.view(torch.uint8)
    x0 = (x & 0x0F).to(torch.int8)
    x0[x0 >= 8] -= 16
    x1 = ((x & 0xF0) >> 4).to(torch.int8)
    x1[x1 >= 8] -= 16
    out = torch.empty(x.shape[0], x.shape[1] * 2, dtype=torch.int32)
    out[:, 0::2] = x0
    out[:, 1::2] = x1
    return out


def linear_w4a16(x, W, sW):
    return (x.float() @ (unpack_i4(W).float() * sW.float()[:, None]).T).half()


def sym_quant_fake(x):
    scale = x.float().abs().amax(dim=1, keepdim=True) / 7
    q = torch.clamp(torch.round(x.float() / scale), -8, 7)
    return q, scale


def linear_w4a4(q, scale, W, sW):
    return ((q @ unpack_i4(W).float().T) * scale * sW.float()[None, :]).half()


def norm(x, eps=1e-5):
    xf = x.float()
    xf = xf - xf.mean(-1, keepdim=True)
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).half()


def fht(x):
    """matmul_hadU butterflies (hadamard.py:63-71) along the last dim, un-normalised."""
    n = x.shape[-1]
    y = x.float().reshape(-1, n, 1)
    out = y.clone()
    while y.shape[1] > 1:
        y = y.view(y.shape[0], y.shape[1] // 2, 2, y.shape[2])
        out = out.view(y.shape)
        out[:, :, 0, :] = y[:, :, 0, :] + y[:, :, 1, :]
        out[:, :, 1, :] = y[:, :, 0, :] - y[:, :, 1, :]
        out = out.view(y.shape[0], y.shape[1], -1)
        y, out = out, y
    return y.reshape(x.shape)


class CpuLayer:
    """One decoder layer's linear path (both modes), weights as SURVEY 8(d)."""

    def __init__(self, hidden=4096, inter=14336, heads=32, kv_heads=8, head_dim=128, seed=1000):
        g = torch.Generator().manual_seed(seed)
        self.hidden, self.inter, self.heads, self.hd = hidden, inter, heads, head_dim
        shapes = {"qkv": ((heads + 2 * kv_heads) * head_dim, hidden), "o": (hidden, heads * head_dim),
                  "gate_up": (2 * inter, hidden), "down": (hidden, inter)}
        self.w = {}
        for name, (N, K) in shapes.items():
            W = torch.randint(0, 256, (N, K // 2), generator=g, dtype=torch.uint8)
            sW = ((0.5 + torch.rand(N, generator=g)) * 2.0 ** -7).half()
            self.w[name] = (W, sW)
        from qspec_b200._had28 import had28          # data only (the 28x28 sign matrix), no kernels
        self.had28 = had28()

    def _lin(self, name, x, w4a4):
        W, sW = self.w[name]
        if w4a4:
            q, s = sym_quant_fake(x)
            return linear_w4a4(q, s, W, sW)
        return linear_w4a16(x, W, sW)

    def forward(self, x, w4a4: bool):
        """Linear path of one layer; attention itself is replaced by a pass-through of q
        (attention is outside the hot path and is NOT charged to the CPU baseline)."""
        T = x.shape[0]
        qkv = self._lin("qkv", norm(x), w4a4)
        attn = qkv[:, : self.heads * self.hd]
        rot = (fht(attn.view(T, self.heads, self.hd).transpose(1, 2)) / (self.heads ** 0.5)).transpose(1, 2)
        x = x + self._lin("o", rot.reshape(T, -1).half(), w4a4)
        gu = self._lin("gate_up", norm(x), w4a4)
        act = torch.nn.functional.silu(gu[:, self.inter:].float()) * gu[:, : self.inter].float()
        n = self.inter // 28
        r = fht(act.view(T, 28, n)) / (self.inter ** 0.5)
        r = torch.einsum("ok,tkn->ton", self.had28, r).reshape(T, -1).half()
        return x + self._lin("down", r, w4a4)


def time_spec_step_sample(B=4, k=3, layers_total=32, reps=1, cfg=None, threads=None):
    """Times ONE layer's share of one speculative step (k draft passes at M=B, one verify pass
    at M=B*(k+1)) and scales by layers_total. Returns dict(seconds_per_step, cores, sample)."""
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    layer = CpuLayer(**(cfg or {}))
    g = torch.Generator().manual_seed(2000 + B)
    xd = torch.randn(B, layer.hidden, generator=g).half()
    xv = torch.randn(B * (k + 1), layer.hidden, generator=g).half()
    best = float("inf")
    for _ in range(max(1, reps)):
        t0 = time.perf_counter()
        for _ in range(k):
            layer.forward(xd, True)
        layer.forward(xv, False)
        best = min(best, time.perf_counter() - t0)
    return {"seconds_per_step": best * layers_total, "seconds_sample": best, "cores": threads,
            "sample": f"1 of {layers_total} decoder layers (linear path only: norm, 4 int4 linears, both rotations; "
                      f"attention and lm_head not charged), {k} draft passes M={B} + 1 verify pass M={B * (k + 1)}, "
                      f"scaled x{layers_total}"}
