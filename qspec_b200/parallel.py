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


# ---------------------------------------------------------------------------------------------
# Peer-mapped partial sums: the all-reduce after o_proj / down_proj fused into the next norm
# ---------------------------------------------------------------------------------------------
TP_MAX = 8    # QSPEC_TP_MAX in include/qspec_b200.h


def _map_symmetric(t: torch.Tensor, group):
    """torch symmetric memory (CUDA VMM handles over NVLink): returns every rank's device pointer of `t`."""
    import torch.distributed._symmetric_memory as symm
    hdl = symm.rendezvous(t, group)
    return [int(p) for p in hdl.buffer_ptrs], hdl


def _map_ipc(t: torch.Tensor, group):
    """Legacy CUDA IPC through torch's own storage sharing: returns every rank's device pointer of `t`."""
    import torch.distributed as dist
    from torch.multiprocessing.reductions import reduce_tensor
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    fn, fargs = reduce_tensor(t)
    gathered = [None] * world
    dist.all_gather_object(gathered, fargs, group=group)
    keep, ptrs = [], []
    for r in range(world):
        if r == rank:
            ptrs.append(t.data_ptr())
            continue
        peer = fn(*gathered[r])
        peer[..., :1].to(t.device)        # one peer copy: makes torch enable peer access between the two devices
        keep.append(peer)
        ptrs.append(peer.data_ptr())
    return ptrs, keep


class PeerPartials:
    """Two [max_tokens, hidden] partial-sum buffers per rank (attention block, MLP block), mapped into every
    peer of the tensor-parallel group, plus the arrival flags and the local epoch words that
    qspec_tp_add_norm_* (include/qspec_b200.h) needs. Replaces the all-reduce of RowParallelLinear.forward
    (layers/linear.py:1143-1144): the row-parallel GEMM writes `partial(which)`, the next norm sums all ranks'."""

    def __init__(self, max_tokens: int, hidden: int, dtype: torch.dtype, device, group):
        import ctypes
        import torch.distributed as dist
        self.group = group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > TP_MAX:
            raise ValueError(f"PeerPartials: at most {TP_MAX} ranks, got {self.world}")
        self.max_tokens, self.hidden = max_tokens, hidden
        self.how = None
        errors = []
        for how in ("symmetric", "ipc"):
            try:
                if how == "symmetric":
                    import torch.distributed._symmetric_memory as symm
                    buf = symm.empty((2, max_tokens, hidden), dtype=dtype, device=device)
                    flags = symm.empty((TP_MAX,), dtype=torch.int32, device=device)
                    buf.zero_(); flags.zero_()
                    torch.cuda.synchronize(device)
                    amax = symm.empty((2 * TP_MAX * max_tokens, 2), dtype=torch.int32, device=device)
                    aflags = symm.empty((TP_MAX,), dtype=torch.int32, device=device)
                    amax.zero_(); aflags.zero_()
                    torch.cuda.synchronize(device)
                    red = symm.empty((2, max_tokens, hidden), dtype=dtype, device=device)
                    rflags = symm.empty((max_tokens,), dtype=torch.int32, device=device)
                    red.zero_(); rflags.zero_()
                    torch.cuda.synchronize(device)
                    bptr, k0 = _map_symmetric(buf, group)
                    fptr, k1 = _map_symmetric(flags, group)
                    aptr, k2 = _map_symmetric(amax, group)
                    afptr, k3 = _map_symmetric(aflags, group)
                    rptr, k4 = _map_symmetric(red, group)
                    rfptr, k5 = _map_symmetric(rflags, group)
                else:
                    buf = torch.zeros((2, max_tokens, hidden), dtype=dtype, device=device)
                    flags = torch.zeros((TP_MAX,), dtype=torch.int32, device=device)
                    torch.cuda.synchronize(device)
                    amax = torch.zeros((2 * TP_MAX * max_tokens, 2), dtype=torch.int32, device=device)
                    aflags = torch.zeros((TP_MAX,), dtype=torch.int32, device=device)
                    torch.cuda.synchronize(device)
                    red = torch.zeros((2, max_tokens, hidden), dtype=dtype, device=device)
                    rflags = torch.zeros((max_tokens,), dtype=torch.int32, device=device)
                    torch.cuda.synchronize(device)
                    bptr, k0 = _map_ipc(buf, group)
                    fptr, k1 = _map_ipc(flags, group)
                    aptr, k2 = _map_ipc(amax, group)
                    afptr, k3 = _map_ipc(aflags, group)
                    rptr, k4 = _map_ipc(red, group)
                    rfptr, k5 = _map_ipc(rflags, group)
                ok = torch.ones(1, device=device)
            except Exception as e:           # noqa: BLE001 - any failure here means "this mapping is unavailable"
                errors.append(f"{how}: {type(e).__name__}: {e}")
                ok = torch.zeros(1, device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)      # all ranks take the same branch
            if ok.item() > 0:
                self.how, self.buf, self.flags, self._keep = how, buf, flags, (k0, k1, k2, k3, k4, k5)
                self.amax, self.amax_flags = amax, aflags
                self.reduced, self.row_flags = red, rflags
                break
        if self.how is None:
            raise RuntimeError("PeerPartials: no peer mapping available: " + " | ".join(errors))
        self.state = torch.zeros(4, dtype=torch.int32, device=device)
        stride = max_tokens * hidden * buf.element_size()
        self._partials = [(ctypes.c_void_p * self.world)(*[p + w * stride for p in bptr]) for w in (0, 1)]
        self._flags = (ctypes.c_void_p * self.world)(*fptr)
        # two-step form of the fused all-reduce (qspec_tp_rs_add_norm_*): every rank's buffers of reduced rows + per-row flags
        self._reduced = [(ctypes.c_void_p * self.world)(*[p + w * stride for p in rptr]) for w in (0, 1)]
        self._row_flags = (ctypes.c_void_p * self.world)(*rfptr)
        # vocab-sharded greedy token (qspec_tp_argmax_combine): every rank's table of (max logit, index) pairs, own flags / epoch
        self.amax_state = torch.zeros(4, dtype=torch.int32, device=device)
        self._amax_slots = (ctypes.c_void_p * self.world)(*aptr)
        self._amax_flags = (ctypes.c_void_p * self.world)(*afptr)
        torch.cuda.synchronize(device)
        dist.barrier(group=group)            # flags are zero everywhere before the first arrival is stored

    def partial(self, which: int, rows: int) -> torch.Tensor:
        return self.buf[which, :rows]

    def timeouts(self) -> int:
        return int(self.state[2].item()) + int(self.amax_state[2].item())
