"""All ranks of a tensor-parallel QSpec model on ONE GPU (test infrastructure).

The product runs one process per GPU and meets its peers through `parallel.PeerPartials` (peer-mapped partial sums +
arrival flags over NVLink). Here every shard lives in this process on the same device and the shards are advanced in
lock step through `QSpecLlamaModel.forward_stages`, which yields after every row-parallel GEMM: when a rank's fused
reduce+add+norm kernel (`qspec_tp_add_norm_*`) runs, every rank's partial of that block is already complete, so the
arrival flags it reads are preset to "arrived" and the flags it writes go to a sink. The kernels, the sharded weights,
the shard-local rotations, the per-(token, rank) activation scales and the rank-ordered fp32 sum are exactly the
product's; only the NVLink flag protocol itself is not exercised (tests/test_gpu_tp.py does that on two GPUs).
"""
from __future__ import annotations

import ctypes
from typing import List

import torch

ARRIVED = 1 << 20          # larger than any epoch a test reaches


class LocalPeers:
    """Duck-types `parallel.PeerPartials` for rank `rank` of `world` shards that share `shared` = [world, 2, T, H]."""

    emulate_two_step = True            # model.forward_stages yields between the two launches of a two-step exchange

    def __init__(self, shared: torch.Tensor, rank: int, amax_shared: torch.Tensor = None, reduced: torch.Tensor = None,
                 row_flags: torch.Tensor = None):
        self.world, self.rank = shared.shape[0], rank
        self.max_tokens, self.hidden = shared.shape[2], shared.shape[3]
        self.buf = shared[rank]                                            # [2, T, H], this rank's partial sums
        self.how = "one-gpu"
        dev = shared.device
        self._mine = torch.full((8,), ARRIVED, dtype=torch.int32, device=dev)      # read by this rank, written by nobody
        self._sink = torch.zeros(self.world, 8, dtype=torch.int32, device=dev)     # this rank's arrivals land here
        self.state = torch.zeros(4, dtype=torch.int32, device=dev)
        self._partials = [(ctypes.c_void_p * self.world)(*[shared[r, w].data_ptr() for r in range(self.world)]) for w in (0, 1)]
        self._flags = (ctypes.c_void_p * self.world)(*[(self._mine if r == rank else self._sink[r]).data_ptr()
                                                       for r in range(self.world)])

        if reduced is not None:
            # two-step fused all-reduce: every rank's [2, T, H] buffer of reduced rows and its per-row flags; the owners' launches
            # (phase 1) really push into all of them and publish the flags the consumers' launches (phase 2) wait for
            self._rs_keep = (reduced, row_flags)
            self._reduced = [(ctypes.c_void_p * self.world)(*[reduced[r, w].data_ptr() for r in range(self.world)]) for w in (0, 1)]
            self._row_flags = (ctypes.c_void_p * self.world)(*[row_flags[r].data_ptr() for r in range(self.world)])
        if amax_shared is not None:
            # greedy-token tables of every rank, [world, 2 * 8 * T, 2]: the push phase of rank r really writes into all of them
            self.amax_state = torch.zeros(4, dtype=torch.int32, device=dev)
            self._amax_keep = amax_shared                                  # the ctypes array below holds raw pointers only
            self._amax_slots = (ctypes.c_void_p * self.world)(*[amax_shared[r].data_ptr() for r in range(self.world)])
            self._amax_flags = self._flags

    def partial(self, which: int, rows: int) -> torch.Tensor:
        return self.buf[which, :rows]

    def timeouts(self) -> int:
        return int(self.state[2].item()) + (int(self.amax_state[2].item()) if hasattr(self, "amax_state") else 0)


def build_shards(cfg, tp: int, max_batch: int, max_tokens: int, device, seed: int = 0):
    """The `tp` shards of one model (same weights as `tp` processes with ranks 0..tp-1 would build)."""
    from qspec_b200.model import QSpecLlamaModel
    shared = torch.zeros(tp, 2, max_tokens, cfg.hidden_size, dtype=torch.float16, device=device)
    amax = torch.zeros(tp, 2 * 8 * max_tokens, 2, dtype=torch.int32, device=device)
    reduced = torch.zeros(tp, 2, max_tokens, cfg.hidden_size, dtype=torch.float16, device=device)
    row_flags = torch.zeros(tp, max_tokens, dtype=torch.int32, device=device)
    shards = []
    for r in range(tp):
        peers = LocalPeers(shared, r, amax, reduced, row_flags)
        shards.append(QSpecLlamaModel(cfg, max_batch=max_batch, max_tokens=max_tokens, device=device, seed=seed,
                                      tp_size=tp, tp_rank=r, peers=peers))
    return shards


@torch.no_grad()
def lockstep_forward(shards: List, input_ids, positions, seq_ids, w4a4: bool):
    """Advances every shard to its next exchange point in turn; returns the per-rank final-norm hidden states."""
    gens = [m.forward_stages(input_ids, positions, seq_ids, w4a4) for m in shards]
    alive = True
    while alive:
        alive = False
        for g in gens:
            try:
                next(g)
                alive = True
            except StopIteration:
                pass
    T = input_ids.numel()
    return [m.buf_norm[:T] for m in shards]


@torch.no_grad()
def lockstep_greedy(shards: List, hidden_per_rank) -> torch.Tensor:
    """Greedy token over the vocab-sharded lm_head with the product's peer-memory kernel (qspec_tp_argmax_combine) in two
    phases -- every rank pushes its (max, index) pairs into every rank's table, then every rank takes the best of its own --
    checked here against the rank-ordered concatenation (lowest index among equal maxima) computed with torch."""
    from qspec_b200 import ops
    best_v, best_i, local = None, None, []
    for r, (m, h) in enumerate(zip(shards, hidden_per_rank)):
        logits = torch.matmul(h, m.lm_head.t())
        val = torch.empty(logits.shape[0], dtype=torch.float32, device=logits.device)
        lidx = ops.argmax_rows(logits, out_val=val)
        local.append((val, lidx))
        idx = lidx + r * m.vocab_shard
        if best_v is None:
            best_v, best_i = val, idx
        else:
            take = val > best_v                                            # ties keep the lower rank = lower index
            best_v = torch.where(take, val, best_v)
            best_i = torch.where(take, idx, best_i)
    outs = [torch.empty_like(best_i) for _ in shards]
    for r, m in enumerate(shards):
        ops.tp_argmax_combine(local[r][0], local[r][1], r * m.vocab_shard, m.peers, out=outs[r], phases=1)
    for r, m in enumerate(shards):
        ops.tp_argmax_combine(local[r][0], local[r][1], r * m.vocab_shard, m.peers, out=outs[r], phases=2)
    for r, o in enumerate(outs):
        assert torch.equal(o, best_i), f"rank {r}: peer-memory arg max differs from the concatenated arg max"
    return best_i
