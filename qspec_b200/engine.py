"""QSpec speculative-decode loop around the dual-mode model: k draft passes in W4A4, one verify
pass in W4A16 over the same weights and KV cache, greedy acceptance.

Mirrors SpecDecodeWorker._run_speculative_decoding_step
(vllm/spec_decode/spec_decode_worker.py:758-858: w4a4=True for the proposer :797-801, w4a4=False
for the MQA scorer :812-813 which scores k+1 tokens per sequence, mqa_scorer.py:65) with greedy
rejection (draft token accepted iff it equals the target argmax), and the metrics of
spec_decode/metrics.py:171-179 (draft_acceptance_rate, system_efficiency).

The whole step (3 x 32 draft layers + 32 verify layers + sampling + acceptance + state update)
is captured once into a CUDA graph (the reference runs ~550 eager launches
per draft step with --enforce_eager); all per-step state lives on the device.
"""
from __future__ import annotations

from typing import Optional

import torch

from .model import QSpecLlamaModel


class QSpecEngine:
    def __init__(self, model: QSpecLlamaModel, batch: int, k: int = 3, use_cuda_graph: bool = True):
        self.model, self.B, self.k = model, batch, k
        dev = model.device
        B = batch
        assert model.max_tokens >= B * (k + 1)
        self.use_graph = use_cuda_graph
        self.cur_tok = torch.zeros(B, dtype=torch.int64, device=dev)
        self.seq_len = torch.zeros(B, dtype=torch.int32, device=dev)        # tokens already in the KV cache
        self.seq_ids = torch.arange(B, dtype=torch.int32, device=dev)
        self.vseq_ids = self.seq_ids.repeat_interleave(k + 1).contiguous()
        self.arange_k1 = torch.arange(k + 1, dtype=torch.int32, device=dev)
        self.draft_toks = torch.zeros(B, k, dtype=torch.int64, device=dev)
        self.out_tokens = torch.zeros(B, k + 1, dtype=torch.int64, device=dev)
        self.out_count = torch.zeros(B, dtype=torch.int64, device=dev)
        # SpecDecodeWorkerMetrics counters (metrics.py): accepted, draft, emitted
        self.counters = torch.zeros(3, dtype=torch.int64, device=dev)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.stream = torch.cuda.Stream(device=dev)
        # pinned host mirrors for the end-to-end API
        self.h_in = torch.zeros(B, dtype=torch.int64).pin_memory()
        self.h_out = torch.zeros(B, k + 2, dtype=torch.int64).pin_memory()
        self.d_out_pack = torch.zeros(B, k + 2, dtype=torch.int64, device=dev)
        # Host-side upper bound of max(seq_len): the device-resident lengths advance by at most k + 1 per step and are
        # never read back in the loop. A step touches positions seq_len .. seq_len + k of the KV cache and of the
        # rotary table; `step` refuses to run once that could pass cfg.max_seq_len (the kernels index both unchecked
        # apart from a last-resort early-out).
        self._len_bound = 0
        self._steps = 0

    # ---------------------------------------------------------------------------------------
    @torch.no_grad()
    def prefill(self, prompt_ids: torch.Tensor) -> None:
        """prompt_ids [B, P] (device int64). Verify-mode (W4A16) pass, as the reference's
        _run_no_spec prefill (spec_decode_worker.py:666-720); fills the KV cache."""
        m = self.model
        B, P = prompt_ids.shape
        assert B == self.B and P + 1 < m.cfg.max_seq_len
        pos = torch.arange(P, dtype=torch.int32, device=m.device)
        chunk = max(1, m.max_tokens)
        last_hidden = torch.empty(B, m.cfg.hidden_size, dtype=torch.float16, device=m.device)
        for b in range(B):
            for s in range(0, P, chunk):
                e = min(P, s + chunk)
                ids = prompt_ids[b, s:e].contiguous()
                h = m.forward(ids, pos[s:e].contiguous(), torch.full((e - s,), b, dtype=torch.int32, device=m.device),
                              w4a4=False)
                if e == P:
                    last_hidden[b].copy_(h[e - s - 1])
        self.cur_tok.copy_(m.greedy_tokens(last_hidden))
        self.seq_len.fill_(P)
        self._len_bound = P
        self.counters.zero_()

    # ---------------------------------------------------------------------------------------
    @torch.no_grad()
    def _spec_step(self) -> None:
        m, B, k = self.model, self.B, self.k
        tok = self.cur_tok
        # ---- proposer: k draft passes, w4a4=True (spec_decode_worker.py:797-801)
        for i in range(k):
            pos = self.seq_len + i
            hidden = m.forward(tok, pos, self.seq_ids, w4a4=True)
            tok = m.greedy_tokens(hidden)
            self.draft_toks[:, i].copy_(tok)
        # ---- scorer: one verify pass over k+1 tokens / sequence, w4a4=False (:812-813)
        vt = torch.cat((self.cur_tok.unsqueeze(1), self.draft_toks), dim=1).reshape(-1)
        vpos = (self.seq_len.unsqueeze(1) + self.arange_k1.unsqueeze(0)).reshape(-1).contiguous()
        hidden = m.forward(vt, vpos, self.vseq_ids, w4a4=False)
        tgt = m.greedy_tokens(hidden).view(B, k + 1)
        # ---- greedy rejection (spec_decode_worker.py:861-915 with temperature 0)
        match = (tgt[:, :k] == self.draft_toks).to(torch.int64)
        n_acc = match.cumprod(dim=1).sum(dim=1)                         # accepted draft tokens per sequence
        nxt = tgt.gather(1, n_acc.unsqueeze(1)).squeeze(1)              # correction / bonus token
        padded = torch.cat((self.draft_toks, nxt.unsqueeze(1)), dim=1)
        idx = self.arange_k1.unsqueeze(0).to(torch.int64)
        emitted = torch.where(idx < n_acc.unsqueeze(1), padded, torch.where(idx == n_acc.unsqueeze(1), nxt.unsqueeze(1), -1))
        self.out_tokens.copy_(emitted)
        self.out_count.copy_(n_acc + 1)
        self.cur_tok.copy_(nxt)
        self.seq_len.add_((n_acc + 1).to(torch.int32))
        self.counters[0] += n_acc.sum()
        self.counters[1] += B * k
        self.counters[2] += (n_acc + 1).sum()
        self.d_out_pack[:, : k + 1].copy_(emitted)
        self.d_out_pack[:, k + 1].copy_(n_acc + 1)

    def capture(self) -> None:
        """Warm up once eagerly (sets kernel attributes, sizes workspaces), then capture."""
        if not self.use_graph or self.graph is not None:
            return
        self._check_capacity()
        state = (self.cur_tok.clone(), self.seq_len.clone(), self.counters.clone())
        self._join_caller_stream()
        with torch.cuda.stream(self.stream):
            self._spec_step()
            torch.cuda.synchronize()
            self.cur_tok.copy_(state[0]); self.seq_len.copy_(state[1]); self.counters.copy_(state[2])
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self.stream):
                self._spec_step()
            torch.cuda.synchronize()
            self.cur_tok.copy_(state[0]); self.seq_len.copy_(state[1]); self.counters.copy_(state[2])
        torch.cuda.synchronize()
        self.graph = g

    def sync_lengths(self) -> int:
        """Reads the true max(seq_len) back (one small D2H copy + sync) and tightens the host-side bound to it."""
        self._len_bound = int(self.seq_len.max().item())
        return self._len_bound

    def _check_capacity(self) -> None:
        S, need = self.model.cfg.max_seq_len, self.k + 1
        if self._len_bound + need > S:
            self.sync_lengths()                           # the bound assumes every draft token was accepted
            if self._len_bound + need > S:
                raise RuntimeError(f"KV cache exhausted: a sequence holds {self._len_bound} tokens and a speculative step "
                                   f"writes {need} more, but max_seq_len is {S}")
        peers = getattr(self.model, "peers", None)
        if peers is not None and self._steps % 64 == 0 and peers.timeouts():
            raise RuntimeError(f"tensor-parallel peer time-out: {peers.timeouts()} arrival flag(s) never came "
                               "(a rank of the group is gone); the partial sums of those steps are stale")

    def _join_caller_stream(self) -> None:
        """self.stream is a non-blocking stream: work the caller queued on ITS stream (prefill, state edits) is ordered before
        the step only by an explicit wait. Without it the step can read seq_len / cur_tok before prefill has written them and
        append draft K/V rows at the wrong cache positions (found by scripts/gpu_determinism.py: intermittently different
        tokens when capture() followed prefill() without a synchronize)."""
        cur = torch.cuda.current_stream(self.model.device)
        if cur != self.stream:
            self.stream.wait_stream(cur)

    def step(self) -> None:
        """One speculative step, device-resident state; asynchronous on self.stream (ordered after the work already queued on
        the caller's current stream; the caller synchronises self.stream before reading the outputs)."""
        self._check_capacity()
        self._len_bound += self.k + 1
        self._steps += 1
        self._join_caller_stream()
        if self.use_graph:
            if self.graph is None:
                self.capture()
            with torch.cuda.stream(self.stream):
                self.graph.replay()
        else:
            with torch.cuda.stream(self.stream):
                self._spec_step()

    def step_host(self, host_tokens: torch.Tensor) -> torch.Tensor:
        """End-to-end API: the caller owns the last token of every sequence on the HOST (as the
        reference engine does); returns host [B, k+2] = emitted tokens (-1 padded) + count.
        H2D of the inputs and D2H of the result are part of the call."""
        self.h_in.copy_(host_tokens)
        with torch.cuda.stream(self.stream):
            self.cur_tok.copy_(self.h_in, non_blocking=True)
        self.step()
        with torch.cuda.stream(self.stream):
            self.h_out.copy_(self.d_out_pack, non_blocking=True)
        self.stream.synchronize()
        return self.h_out

    # ---------------------------------------------------------------------------------------
    def metrics(self) -> dict:
        peers = getattr(self.model, "peers", None)
        if peers is not None and peers.timeouts():
            raise RuntimeError(f"tensor-parallel peer time-out: {peers.timeouts()} arrival flag(s) never came")
        acc, draft, emitted = [int(v) for v in self.counters.tolist()]
        steps = draft // (self.B * self.k) if draft else 0
        max_emit = steps * self.B * (self.k + 1)
        return {"accepted_tokens": acc, "draft_tokens": draft, "emitted_tokens": emitted,
                "draft_acceptance_rate": acc / draft if draft else float("nan"),
                "system_efficiency": emitted / max_emit if max_emit else float("nan")}
