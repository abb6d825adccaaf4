"""Tensor-parallel sharding of the dual-mode layer; host-side logic only.

Column-parallel (qkv, gate_up): shard N; scales shard with N; no collective.
Row-parallel (o, down): shard K (packed dim: k columns -> k/2 bytes); scales replicated; the
partial outputs are summed with one all-reduce (RowParallelLinear.forward, layers/linear.py:1143-1144).
"""
from __future__ import annotations

from typing import List, Tuple

import torch


def shard_column(weight: torch.Tensor, scales: torch.Tensor, tp: int, rank: int) -> Tuple[torch.Tensor, torch.Tensor]:
    N = weight.shape[0]
    if N % tp:
        raise ValueError(f"column-parallel: N={N} not divisible by tp={tp}")
    n = N // tp
    return weight[rank * n:(rank + 1) * n].contiguous(), scales.reshape(-1)[rank * n:(rank + 1) * n].contiguous()


def shard_row(weight: torch.Tensor, scales: torch.Tensor, tp: int, rank: int) -> Tuple[torch.Tensor, torch.Tensor]:
    K2 = weight.shape[1]
    if (2 * K2) % (128 * tp):
        raise ValueError(f"row-parallel: K={2 * K2} must be a multiple of 128*tp")
    k2 = K2 // tp
    return weight[:, rank * k2:(rank + 1) * k2].contiguous(), scales.reshape(-1).contiguous()


def shard_qkv(weight: torch.Tensor, scales: torch.Tensor, hq: int, hkv: int, d: int, tp: int, rank: int):
    """Fused [q; k; v] rows (quarot_llama.py:152-173) sharded by heads: rank gets hq/tp q-heads and hkv/tp kv-heads."""
    if hq % tp or hkv % tp:
        raise ValueError("heads must be divisible by tp")
    s = scales.reshape(-1)
    q, k, v = weight[: hq * d], weight[hq * d:(hq + hkv) * d], weight[(hq + hkv) * d:]
    sq, sk, sv = s[: hq * d], s[hq * d:(hq + hkv) * d], s[(hq + hkv) * d:]
    a, b = hq // tp * d, hkv // tp * d
    w = torch.cat((q[rank * a:(rank + 1) * a], k[rank * b:(rank + 1) * b], v[rank * b:(rank + 1) * b]), 0).contiguous()
    sc = torch.cat((sq[rank * a:(rank + 1) * a], sk[rank * b:(rank + 1) * b], sv[rank * b:(rank + 1) * b]), 0).contiguous()
    return w, sc


def shard_gate_up(weight: torch.Tensor, scales: torch.Tensor, inter: int, tp: int, rank: int):
    """[up ; gate] rows (quarot_llama.py:301-314) co-sharded so SiLU*up stays local."""
    if inter % tp:
        raise ValueError("intermediate size must be divisible by tp")
    s = scales.reshape(-1)
    n = inter // tp
    w = torch.cat((weight[rank * n:(rank + 1) * n], weight[inter + rank * n: inter + (rank + 1) * n]), 0).contiguous()
    sc = torch.cat((s[rank * n:(rank + 1) * n], s[inter + rank * n: inter + (rank + 1) * n]), 0).contiguous()
    return w, sc


def combine_sharded_argmax(local_max: torch.Tensor, local_idx: torch.Tensor, rank: int, shard: int, group=None):
    """Greedy token over a vocab-sharded lm_head: all-gather (max, global index) and take the best.
    Ties resolve to the lowest rank, i.e. the lowest global index, like a full argmax."""
    import torch.distributed as dist
    tp = dist.get_world_size(group)
    pack = torch.stack((local_max.float(), (local_idx + rank * shard).float()), dim=-1)
    gathered: List[torch.Tensor] = [torch.empty_like(pack) for _ in range(tp)]
    dist.all_gather(gathered, pack, group=group)
    allp = torch.stack(gathered, 0)
    best = allp[..., 0].argmax(dim=0)
    return allp[..., 1].gather(0, best.unsqueeze(0)).squeeze(0).long()
