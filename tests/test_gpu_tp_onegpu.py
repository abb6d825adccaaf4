"""Tensor-parallel numerics against the TP-aware CPU oracle, on ONE GPU (SURVEY 8(e), 7 hard part 5).

Every rank's shard of a TP=2 / TP=8 model runs on cuda:0 in lock step (tests/tp_emulation.py): the sharded GEMMs,
shard-local rotations (H_{heads/t}, had28 (x) H_{n/t}), per-(token, rank) activation scales and the fused
reduce+add+norm kernels are the product's. The oracle (oracle/qspec_model_oracle.py, `ranks=`) restates
RowParallelLinear.forward (vllm/model_executor/layers/linear.py:1125-1149): f16 partial per rank, rank-ordered fp32 sum,
one rounding. Bars: all ranks hold the SAME hidden state bit for bit; hidden states within north_star's 1e-2 relative
(max|d| / max|ref|) of the oracle in both modes; greedy speculative decode gives the oracle's tokens and acceptance
counts on well-conditioned seeds.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

REL_TOL = 1e-2            # north_star: fp16 outputs of both modes within 1e-2 relative


def rel_err(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())


def assert_hidden_close(got, oracle_forward, what):
    """Hidden state within REL_TOL of the oracle -- or, when the case is ill-conditioned, within the oracle's OWN spread
    under rounding-level jitter of the ops outside the bit-exact contract (attention / SiLU outputs moved by 1 f16 ulp,
    OracleModel.jitter): the int4 quantisers of the draft path turn such ulps into percent-level differences of a
    random-weight model's hidden state, which says nothing about the path (whose kernels are compared bit for bit in
    tests/test_gpu_kernels.py). Verify mode has no quantiser on the activation path and must always meet REL_TOL."""
    ref = oracle_forward(None).astype(np.float32)
    rel = rel_err(got, ref)
    if rel <= REL_TOL:
        return rel
    spread = max(rel_err(oracle_forward((js, 1e-3)).astype(np.float32), ref) for js in (1, 2, 3))
    assert spread > REL_TOL and rel <= 2.0 * spread, \
        f"{what}: hidden state {rel:.3e} relative from the oracle, whose own rounding-level spread is only {spread:.3e}"
    return rel


@pytest.fixture(scope="module")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def _cfg(tp):
    from qspec_b200.model import QSpecLlamaConfig
    if tp == 8:
        # one Llama-3-8B layer: per rank 4 q heads, 1 kv head, 28 x 64 intermediate columns (what TP=8 of configs[2] runs)
        return QSpecLlamaConfig(num_hidden_layers=1, vocab_size=2048, max_seq_len=64, name="8b-1layer")
    # TP=2 / 4: 8 heads, 4 kv heads, hidden 1024, intermediate 28 * 128 * tp/2
    return QSpecLlamaConfig.tiny(hidden_size=1024, num_attention_heads=8, num_key_value_heads=4,
                                 intermediate_size=28 * 64 * tp, num_hidden_layers=2, vocab_size=2048, max_seq_len=64)


def _inputs(cfg, dev, T=10, seed=21):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, cfg.vocab_size, (T,), generator=g)
    pos = torch.arange(T, dtype=torch.int32) % 5
    seq = torch.arange(T, dtype=torch.int32) // 5
    return ids, pos, seq


@pytest.mark.parametrize("scatter_min", [None, "1"])
@pytest.mark.parametrize("tp", [2, 4, 8])
def test_tp_hidden_states_match_tp_oracle(cuda, tp, scatter_min, monkeypatch):
    """scatter_min = 1 forces the two-step form of the fused all-reduce (row r summed by rank r % tp and pushed to the peers,
    qspec_tp_rs_add_norm_*) on every exchange; the default is the one-step form."""
    if scatter_min is not None:
        monkeypatch.setenv("QSPEC_TP_SCATTER_MIN", scatter_min)
    from oracle.qspec_model_oracle import OracleModel
    from tp_emulation import build_shards, lockstep_forward
    cfg = _cfg(tp)
    shards = build_shards(cfg, tp, max_batch=2, max_tokens=16, device=cuda, seed=4)
    om = OracleModel.from_gpu_shards(shards, 2)
    assert om.tp == tp
    ids, pos, seq = _inputs(cfg, cuda)
    for w4a4 in (False, True, False, True):               # several passes: epochs advance, both partial buffers are reused
        hs = lockstep_forward(shards, ids.to(cuda), pos.to(cuda), seq.to(cuda), w4a4)
        torch.cuda.synchronize()
        got = [h.float().cpu().numpy() for h in hs]
        for r in range(1, tp):
            assert np.array_equal(got[0], got[r]), f"rank {r} disagrees with rank 0 (w4a4={w4a4})"
        assert np.isfinite(got[0]).all()

        def oracle_forward(jitter, w4a4=w4a4):
            om.jitter, om._jitter_calls = jitter, 0
            return om.forward(ids.numpy(), pos.numpy(), seq.numpy(), w4a4)

        rel = assert_hidden_close(got[0], oracle_forward, f"TP={tp} w4a4={w4a4}")
        if not w4a4:
            assert rel <= REL_TOL, f"TP={tp} verify mode: hidden state {rel:.3e} relative from the TP oracle"
    for m in shards:
        assert m.peers.timeouts() == 0


def _tp_oracle_run(shards, prompts, B, k, steps, jitter=None):
    from oracle.qspec_model_oracle import OracleModel, spec_step
    om = OracleModel.from_gpu_shards(shards, B)
    om.jitter = jitter
    P = prompts.shape[1]
    o_last = [om.forward(prompts[b].numpy(), np.arange(P), np.full(P, b), False)[-1] for b in range(B)]
    cur, _ = om.greedy(np.stack(o_last))
    first, seq_len, out = cur.copy(), np.full(B, P), []
    for _ in range(steps):
        emitted, n_acc, cur, seq_len, _ = spec_step(om, cur, seq_len, k)
        out.append([[int(t) for t in e] for e in emitted])
    return first, out


def _tp_spec_run(cuda, tp, seed, steps=4, B=2, k=3, shift=None):
    """Greedy speculative decode of a TP model in lock step beside the oracle-backed loop of the same shards. Returns
    (error or None, ill-conditioned? -- the oracle's own tokens change under rounding-level jitter of the ops outside the
    bit-exact contract, evaluated only when there is an error --, acceptance rate)."""
    from tp_emulation import build_shards, lockstep_forward, lockstep_greedy
    cfg = _cfg(tp)
    shards = build_shards(cfg, tp, max_batch=B, max_tokens=16, device=cuda, seed=seed)
    if shift is not None:
        for m in shards:
            m.make_modes_agree(shift)
    g = torch.Generator().manual_seed(9)
    P = 6
    prompts = torch.randint(0, cfg.vocab_size, (B, P), generator=g)
    base = _tp_oracle_run(shards, prompts, B, k, steps)
    last = []
    for b in range(B):
        idsb = prompts[b].to(cuda)
        posb = torch.arange(P, dtype=torch.int32, device=cuda)
        seqb = torch.full((P,), b, dtype=torch.int32, device=cuda)
        hs = lockstep_forward(shards, idsb, posb, seqb, False)
        last.append([h[P - 1].clone() for h in hs])
    cur = lockstep_greedy(shards, [torch.stack([last[b][r] for b in range(B)]) for r in range(tp)])
    err = None if np.array_equal(base[0], cur.cpu().numpy()) else "prefill token differs"
    dev_len = torch.full((B,), P, dtype=torch.int32, device=cuda)
    seq_ids = torch.arange(B, dtype=torch.int32, device=cuda)
    vseq = seq_ids.repeat_interleave(k + 1).contiguous()
    ar = torch.arange(k + 1, dtype=torch.int32, device=cuda)
    acc_total = 0
    for s in range(steps):
        emitted = base[1][s]
        tok, draft = cur, []
        for i in range(k):
            hs = lockstep_forward(shards, tok, (dev_len + i).contiguous(), seq_ids, True)
            tok = lockstep_greedy(shards, hs)
            draft.append(tok)
        draft = torch.stack(draft, 1)
        vt = torch.cat((cur.unsqueeze(1), draft), 1).reshape(-1)
        vpos = (dev_len.unsqueeze(1) + ar.unsqueeze(0)).reshape(-1).contiguous()
        hs = lockstep_forward(shards, vt, vpos, vseq, False)
        tgt = lockstep_greedy(shards, hs).view(B, k + 1)
        match = (tgt[:, :k] == draft).long()
        na = match.cumprod(1).sum(1)
        nxt = tgt.gather(1, na.unsqueeze(1)).squeeze(1)
        for b in range(B):
            mine = draft[b, : int(na[b])].tolist() + [int(nxt[b])]
            if err is None and mine != emitted[b]:
                err = f"TP={tp} seed {seed} step {s} seq {b}: tokens {mine} vs oracle {emitted[b]}"
        cur = nxt
        dev_len = dev_len + (na + 1).int()
        acc_total += int(na.sum())
    ill = False
    if err is not None:
        for js in (1, 2, 3):
            jit = _tp_oracle_run(shards, prompts, B, k, steps, jitter=(js, 1e-3))
            if not np.array_equal(jit[0], base[0]) or jit[1] != base[1]:
                ill = True
                break
    return err, ill, acc_total / (steps * B * k)


@pytest.mark.parametrize("tp,scatter_min", [(2, None), (8, None), (8, "1")])
def test_tp_greedy_spec_decode_matches_tp_oracle(cuda, tp, scatter_min, monkeypatch):
    if scatter_min is not None:
        monkeypatch.setenv("QSPEC_TP_SCATTER_MIN", scatter_min)
    _tp_greedy_spec_decode_matches_tp_oracle(cuda, tp)


def _tp_greedy_spec_decode_matches_tp_oracle(cuda, tp):
    """Identical tokens and accepted counts vs the TP oracle. As in the TP=1 test, a seed may be skipped only when it
    mismatches AND the oracle's own tokens change under rounding-level jitter of attention / SiLU (ops outside the
    bit-exact contract): then the arg max was decided by rounding, not by the algorithm."""
    checked = 0
    for seed in range(3, 12):
        err, ill, _ = _tp_spec_run(cuda, tp, seed, steps=3 if tp == 8 else 4)
        if err is not None and ill:
            continue
        assert err is None, f"{err} (and the oracle's tokens are stable under rounding-level jitter)"
        checked += 1
        if checked == 2:
            break
    assert checked >= 1, "no seed matched the oracle"


@pytest.mark.parametrize("tp", [2, 8])
def test_tp_argmax_combine_matches_concatenated_argmax(cuda, tp):
    """qspec_tp_argmax_combine (push + combine through every rank's table; both table parities; ties between ranks go to
    the lowest global index) against torch.argmax over the rank-ordered concatenation of the shard logits."""
    import ctypes
    from types import SimpleNamespace
    from qspec_b200 import ops
    T, shard, cap = 37, 512, 64
    amax = torch.zeros(tp, 2 * 8 * cap, 2, dtype=torch.int32, device=cuda)
    mine = torch.full((8,), 1 << 20, dtype=torch.int32, device=cuda)
    sink = torch.zeros(tp, 8, dtype=torch.int32, device=cuda)
    peers = [SimpleNamespace(world=tp, rank=r, max_tokens=cap, amax_state=torch.zeros(4, dtype=torch.int32, device=cuda),
                             _amax_slots=(ctypes.c_void_p * tp)(*[amax[q].data_ptr() for q in range(tp)]),
                             _amax_flags=(ctypes.c_void_p * tp)(*[(mine if q == r else sink[q]).data_ptr() for q in range(tp)]))
             for r in range(tp)]
    g = torch.Generator().manual_seed(3)
    for rnd in range(3):                                   # epochs 1, 2, 3: both tables, and the first one again
        logits = torch.randn(T, tp * shard, generator=g).half().to(cuda)
        logits[5] = 0.25                                   # a row of equal values: index 0 wins
        logits[6, shard - 1] = 9.0
        logits[6, shard] = 9.0                             # the same maximum on two ranks: the lower index wins
        logits[7, (tp - 1) * shard + 3] = 11.0             # maximum on the last rank
        ref = torch.argmax(logits.float(), dim=1)
        assert int(ref[5]) == 0 and int(ref[6]) == shard - 1
        loc = []
        for r in range(tp):
            part = logits[:, r * shard:(r + 1) * shard].contiguous()
            val = torch.empty(T, dtype=torch.float32, device=cuda)
            loc.append((val, ops.argmax_rows(part, out_val=val)))
        outs = [torch.full((T,), -1, dtype=torch.int64, device=cuda) for _ in range(tp)]
        for r in range(tp):
            ops.tp_argmax_combine(loc[r][0], loc[r][1], r * shard, peers[r], out=outs[r], phases=1)
        for r in range(tp):
            ops.tp_argmax_combine(loc[r][0], loc[r][1], r * shard, peers[r], out=outs[r], phases=2)
        for r in range(tp):
            assert torch.equal(outs[r], ref), f"round {rnd} rank {r}"
            assert int(peers[r].amax_state[0]) == rnd + 1 and int(peers[r].amax_state[2]) == 0
    with pytest.raises(ValueError):
        ops.tp_argmax_combine(loc[0][0][:0], loc[0][1][:0], 0, peers[0])


@pytest.mark.parametrize("tp,rows,K", [(2, 7, 4096), (4, 64, 4096), (8, 256, 4096), (8, 5, 8192)])
def test_tp_two_step_reduce_equals_one_step(cuda, tp, rows, K, monkeypatch):
    """qspec_tp_rs_add_norm_* (row r summed by rank r % tp, pushed to every peer, consumed from local memory) against the
    one-step qspec_tp_add_norm_* on the same partial sums: residual, packed int4 bytes, scales, sums and the f16 norm
    bit-identical on every rank, over three exchanges (both buffer parities, epochs advance, per-row flags are re-used)."""
    import ctypes
    from types import SimpleNamespace
    from oracle import qspec_oracle as O
    from qspec_b200 import ops
    T = rows
    shared = torch.zeros(tp, 2, T, K, dtype=torch.float16, device=cuda)
    reduced = torch.zeros(tp, 2, T, K, dtype=torch.float16, device=cuda)
    row_flags = torch.zeros(tp, T, dtype=torch.int32, device=cuda)
    mine = torch.full((8,), 1 << 20, dtype=torch.int32, device=cuda)
    sink = torch.zeros(tp, 8, dtype=torch.int32, device=cuda)

    def peers_of(r, two_step):
        ns = SimpleNamespace(world=tp, rank=r, max_tokens=T, hidden=K, buf=shared[r], state=torch.zeros(4, dtype=torch.int32, device=cuda),
                             _partials=[(ctypes.c_void_p * tp)(*[shared[q, w].data_ptr() for q in range(tp)]) for w in (0, 1)],
                             _flags=(ctypes.c_void_p * tp)(*[(mine if q == r else sink[q]).data_ptr() for q in range(tp)]))
        if two_step:
            ns._reduced = [(ctypes.c_void_p * tp)(*[reduced[q, w].data_ptr() for q in range(tp)]) for w in (0, 1)]
            ns._row_flags = (ctypes.c_void_p * tp)(*[row_flags[q].data_ptr() for q in range(tp)])
        return ns

    one = [peers_of(r, False) for r in range(tp)]
    two = [peers_of(r, True) for r in range(tp)]
    monkeypatch.setenv("QSPEC_TP_SCATTER_MIN", "1")
    assert ops.tp_two_step(two[0], rows) and not ops.tp_two_step(one[0], rows)
    for it in range(3):
        which = it & 1
        for r in range(tp):
            shared[r, which].copy_(torch.from_numpy(O.make_act(T, K, 700 + 10 * it + r, heavy_tail=(r == 1))))
        x0 = torch.from_numpy(O.make_act(T, K, 650 + it)).to(cuda)
        ref = []
        for r in range(tp):
            x = x0.clone()
            q = torch.empty(T, K // 2, dtype=torch.int8, device=cuda)
            sc, isum = torch.empty(T, dtype=torch.float16, device=cuda), torch.empty(T, dtype=torch.float16, device=cuda)
            ops_out = torch.empty(T, K, dtype=torch.float16, device=cuda)
            if it == 2:
                ops.tp_add_rms_norm_fuse_sum_fp16(ops_out, x, one[r], which, 1e-5)
            else:
                ops.tp_add_rms_norm_fuse_sum_i4(q, x, one[r], which, isum, sc, 1e-5)
            ref.append((x, q, sc, isum, ops_out))
        got = []
        for r in range(tp):
            got.append((x0.clone(), torch.full((T, K // 2), 77, dtype=torch.int8, device=cuda), torch.empty(T, dtype=torch.float16, device=cuda),
                        torch.empty(T, dtype=torch.float16, device=cuda), torch.empty(T, K, dtype=torch.float16, device=cuda)))
        for phases in (1, 2):                                  # every rank's owners' launch before any rank's consumers' launch
            for r in range(tp):
                x, q, sc, isum, out = got[r]
                if it == 2:
                    ops.tp_add_rms_norm_fuse_sum_fp16(out, x, two[r], which, 1e-5, phases=phases)
                else:
                    ops.tp_add_rms_norm_fuse_sum_i4(q, x, two[r], which, isum, sc, 1e-5, phases=phases)
        torch.cuda.synchronize()
        for r in range(tp):
            assert torch.equal(got[r][0], ref[r][0]), f"exchange {it} rank {r}: residual"
            if it == 2:
                assert torch.equal(got[r][4], ref[r][4]), f"exchange {it} rank {r}: f16 norm"
            else:
                assert torch.equal(got[r][1], ref[r][1]) and torch.equal(got[r][2], ref[r][2]) and torch.equal(got[r][3], ref[r][3])
            assert int(two[r].state[2]) == 0
        for r in range(1, tp):
            assert torch.equal(got[r][0], got[0][0])
