"""GPU tests of the decode loop around the hot path: RoPE/KV-cache/attention neighbours
(torch fp32 reference, tolerance stated), CUDA-graph == eager, and the end-to-end greedy
speculative decode against the CPU oracle model (identical tokens, identical acceptance)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def test_rope_kvcache_matches_numpy(cuda):
    from qspec_b200 import ops
    from qspec_b200.model import build_cos_sin_cache
    Hq, Hkv, D, S, B, T = 8, 2, 128, 64, 3, 7
    g = torch.Generator().manual_seed(0)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).half()
    pos = torch.tensor([0, 5, 9, 63, 1, 2, 3], dtype=torch.int32)
    seq = torch.tensor([0, 0, 1, 2, 2, 1, 0], dtype=torch.int32)
    cs = build_cos_sin_cache(D, S, 500000.0, "cpu")
    ck = torch.zeros(B, S, Hkv, D, dtype=torch.float16, device=cuda)
    cv = torch.zeros_like(ck)
    qd = qkv.to(cuda).clone()
    ops.rope_kvcache(qd, cs.to(cuda), pos.to(cuda), seq.to(cuda), ck, cv, Hq, Hkv, D)
    x = qkv.float().view(T, Hq + 2 * Hkv, D)
    c, s = cs[pos.long()].float()[:, None, : D // 2], cs[pos.long()].float()[:, None, D // 2:]
    r = x[:, : Hq + Hkv]
    x1, x2 = r[..., : D // 2], r[..., D // 2:]
    rot = torch.cat((x1 * c - x2 * s, x2 * c + x1 * s), -1).half()
    got = qd.cpu().view(T, Hq + 2 * Hkv, D)
    assert torch.equal(got[:, : Hq + Hkv], rot)                      # same fp32 op order => exact
    for t in range(T):
        assert torch.equal(ck[seq[t], pos[t]].cpu(), rot[t, Hq:])
        assert torch.equal(cv[seq[t], pos[t]].cpu(), qkv.view(T, Hq + 2 * Hkv, D)[t, Hq + Hkv:])


def test_decode_attention_matches_torch_fp32(cuda):
    from qspec_b200 import ops
    Hq, Hkv, D, S, B = 8, 2, 128, 300, 2
    g = torch.Generator().manual_seed(1)
    ck = torch.randn(B, S, Hkv, D, generator=g).half().to(cuda)
    cv = torch.randn(B, S, Hkv, D, generator=g).half().to(cuda)
    pos = torch.tensor([0, 17, 255, 299, 128], dtype=torch.int32, device=cuda)
    seq = torch.tensor([0, 1, 0, 1, 1], dtype=torch.int32, device=cuda)
    T = pos.numel()
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).half().to(cuda)
    out = torch.empty(T, Hq * D, dtype=torch.float16, device=cuda)
    ops.decode_attention(qkv, ck, cv, pos, seq, out, Hq, Hkv, D, 1.0 / np.sqrt(D))
    q = qkv.view(T, Hq + 2 * Hkv, D)[:, :Hq].float()
    for t in range(T):
        n = int(pos[t]) + 1
        K = ck[seq[t], :n].float().repeat_interleave(Hq // Hkv, dim=1)     # [n, Hq, D]
        V = cv[seq[t], :n].float().repeat_interleave(Hq // Hkv, dim=1)
        sc = torch.einsum("hd,nhd->hn", q[t], K) / np.sqrt(D)
        ref = torch.einsum("hn,nhd->hd", sc.softmax(-1), V)
        # fp32 math, f16 output: tolerance 2e-3 absolute on O(1) values
        assert (out[t].view(Hq, D).float() - ref).abs().max() <= 2e-3


@pytest.mark.parametrize("case", ["decode", "verify", "prefill", "tp2", "tp8", "g8", "g2", "long", "long_verify", "odd_S"])
def test_attn_block_equals_chain(cuda, case):
    """The one-launch attention block (rope + cache append + attention + rotation over heads + quantiser) against the
    chain of the separate kernels: caches, attention output, rotated f16 output, packed int4 bytes and scales are all
    bit-identical, including passes whose tokens attend to keys written in the same pass (verify, prefill)."""
    from qspec_b200 import ops
    Hq, Hkv = {"tp2": (16, 4), "tp8": (4, 1), "g8": (64, 8), "g2": (4, 2)}.get(case, (32, 8))
    D, S, B = 128, 256, 3
    if case in ("long", "long_verify"):
        S = 1536          # more keys than fit the staging area: key rows beyond it come from global memory, values in chunks
    elif case == "odd_S":
        S = 333           # cache length that is not a multiple of 8
    if case == "long":
        seq, pos = [0, 1, 2, 1], [1535, 700, 127, 128]
    elif case == "long_verify":
        seq, pos = [0, 0, 0, 0, 2, 2, 2, 2], [1200, 1201, 1202, 1203, 290, 291, 292, 293]
    elif case == "odd_S":
        seq, pos = [0, 1, 2, 1], [332, 200, 0, 129]
    elif case == "decode":
        seq, pos = [0, 1, 2, 1], [17, 200, 0, 255]
    elif case == "prefill":
        seq, pos = [1] * 12, list(range(12))
    else:
        seq, pos = [0, 0, 0, 0, 2, 2, 2, 2], [10, 11, 12, 13, 100, 101, 102, 103]
    if case in ("decode", "long", "odd_S"):
        seq[3] = 2                                                         # two tokens of sequence 2, far apart
    T = len(seq)
    assert ops.attn_block_supported(T, Hq, Hkv, D, S)
    g = torch.Generator().manual_seed(11)
    ck0 = torch.randn(B, S, Hkv, D, generator=g).half().to(cuda)
    cv0 = torch.randn(B, S, Hkv, D, generator=g).half().to(cuda)
    qkv0 = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).half().to(cuda)
    from qspec_b200.model import build_cos_sin_cache
    cos_sin = build_cos_sin_cache(D, S, 500000.0, cuda)
    posd = torch.tensor(pos, dtype=torch.int32, device=cuda)
    seqd = torch.tensor(seq, dtype=torch.int32, device=cuda)
    sm, hs = 1.0 / np.sqrt(D), 1.0 / np.sqrt(Hq)
    # chain
    qkv, ck, cv = qkv0.clone(), ck0.clone(), cv0.clone()
    attn = torch.empty(T, Hq * D, dtype=torch.float16, device=cuda)
    ops.rope_kvcache(qkv, cos_sin, posd, seqd, ck, cv, Hq, Hkv, D)
    ops.decode_attention(qkv, ck, cv, posd, seqd, attn, Hq, Hkv, D, sm)
    had = torch.empty_like(attn)
    ops.hadamard_heads(attn, Hq, D, hs, out=had)
    q = torch.empty(T, Hq * D // 2, dtype=torch.int8, device=cuda)
    sc = torch.empty(T, dtype=torch.float16, device=cuda)
    ops.hadamard_heads(attn, Hq, D, hs, out_q=q, scale_out=sc)
    # one launch, quantised output
    qkv2, ck2, cv2 = qkv0.clone(), ck0.clone(), cv0.clone()
    attn2 = torch.empty_like(attn)
    q2, sc2 = torch.empty_like(q), torch.empty_like(sc)
    ops.attn_block(qkv2, cos_sin, posd, seqd, ck2, cv2, Hq, Hkv, D, sm, hs, out_q=q2, scale_out=sc2, attn_out=attn2)
    assert torch.equal(qkv2, qkv0)                                         # input left untouched
    assert torch.equal(ck2, ck) and torch.equal(cv2, cv)
    assert torch.equal(attn2, attn)
    assert torch.equal(sc2, sc) and torch.equal(q2, q)
    # one launch, f16 output (verify mode)
    ck3, cv3 = ck0.clone(), cv0.clone()
    had3 = torch.empty_like(had)
    ops.attn_block(qkv0, cos_sin, posd, seqd, ck3, cv3, Hq, Hkv, D, sm, hs, out=had3)
    assert torch.equal(ck3, ck) and torch.equal(cv3, cv)
    assert torch.equal(had3, had)


@pytest.mark.parametrize("rows,V", [(4, 128256), (16, 128256), (3, 1024), (5, 4100), (1, 8)])
def test_argmax_rows_matches_numpy(cuda, rows, V):
    """Greedy token kernel: index of the row maximum, lowest index among equal maxima (numpy's rule), plus the value."""
    from qspec_b200 import ops
    g = torch.Generator().manual_seed(rows * 7 + V)
    ld = (V + 7) // 8 * 8
    buf = torch.randn(rows, ld, generator=g).half()
    buf[0, V // 3] = 9.0
    buf[0, V - 1] = 9.0                                                    # tie: the lower index wins
    if rows > 1:
        buf[1, V - 1] = 11.0                                               # maximum in the (possibly ragged) tail
    logits = buf.to(cuda)[:, :V]
    val = torch.empty(rows, dtype=torch.float32, device=cuda)
    for _ in range(2):                                                     # twice: the counters clean themselves
        idx = ops.argmax_rows(logits, out_val=val)
    ref = buf[:, :V].float().numpy()
    assert np.array_equal(idx.cpu().numpy(), ref.argmax(axis=1))
    assert np.array_equal(val.cpu().numpy(), ref.max(axis=1))


def _tiny_engine(cuda, graph, seed=0, B=2, k=3, cfg=None, shift=None):
    from qspec_b200.engine import QSpecEngine
    from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel
    cfg = cfg or QSpecLlamaConfig.tiny()
    model = QSpecLlamaModel(cfg, max_batch=B, max_tokens=max(16, B * (k + 1)), device=cuda, seed=seed)
    if shift is not None:
        model.make_modes_agree(shift)
    eng = QSpecEngine(model, B, k, use_cuda_graph=graph)
    return cfg, model, eng


def test_cuda_graph_step_equals_eager(cuda):
    B, k, steps = 2, 3, 6
    outs = []
    for graph in (False, True):
        cfg, model, eng = _tiny_engine(cuda, graph)
        g = torch.Generator().manual_seed(5)
        prompts = torch.randint(0, cfg.vocab_size, (B, 12), generator=g).to(cuda)
        eng.prefill(prompts)
        toks = []
        for _ in range(steps):
            eng.step()
            torch.cuda.synchronize()
            toks.append(eng.d_out_pack.cpu().clone())
        outs.append((torch.stack(toks), eng.metrics()))
    assert torch.equal(outs[0][0], outs[1][0])
    assert outs[0][1] == outs[1][1]


JITTER_FRACTION = 1e-3      # fraction of attention / SiLU outputs moved by 1 f16 ulp in the rounding-sensitivity probe


def _oracle_run(model, prompts, B, k, steps, jitter=None):
    """The oracle-backed loop for `model`'s weights: prefill (verify mode, one sequence at a time like the engine), then
    `steps` speculative steps. Returns (first tokens, [(emitted, n_acc, seq_len)] per step, min top-2 logit margin)."""
    from oracle.qspec_model_oracle import OracleModel, spec_step
    om = OracleModel.from_gpu_model(model, B)
    om.jitter = jitter
    P = prompts.shape[1]
    last = [om.forward(prompts[b].numpy(), np.arange(P), np.full(P, b), False)[-1] for b in range(B)]
    cur, _ = om.greedy(np.stack(last))
    first = cur.copy()
    seq_len = np.full(B, P)
    out, min_margin = [], 1e9
    for _ in range(steps):
        emitted, n_acc, cur, seq_len, margin = spec_step(om, cur, seq_len, k)
        min_margin = min(min_margin, margin)
        out.append(([[int(t) for t in e] for e in emitted], n_acc.copy(), seq_len.copy()))
    return first, out, min_margin


def _decided_by_rounding(model, prompts, B, k, steps, base):
    """True when the oracle's OWN tokens change under a +-1-ulp perturbation of a JITTER_FRACTION of the attention / SiLU
    outputs (the ops outside the bit-exact contract): such a run is ill-conditioned, a GPU / oracle difference on it says
    nothing about the path."""
    for seed in (1, 2, 3):
        first, out, _ = _oracle_run(model, prompts, B, k, steps, jitter=(seed, JITTER_FRACTION))
        if not np.array_equal(first, base[0]) or any(a[0] != b[0] for a, b in zip(out, base[1])):
            return True
    return False


def _e2e_case(cuda, seed, shift=None, cfg=None, B=2, k=3, steps=5, P=10, graph=True):
    """One seeded run of the engine beside the oracle-backed loop. Returns (error or None, ill-conditioned?, histogram of
    accepted-draft counts per (step, sequence), engine metrics). `ill-conditioned` is only evaluated when there IS an
    error: the oracle loop is repeated with rounding-level jitter (see _decided_by_rounding)."""
    cfg, model, eng = _tiny_engine(cuda, graph=graph, seed=seed, B=B, k=k, cfg=cfg, shift=shift)
    g = torch.Generator().manual_seed(9)
    prompts = torch.randint(0, cfg.vocab_size, (B, P), generator=g)
    eng.prefill(prompts.to(cuda))
    base = _oracle_run(model, prompts, B, k, steps)
    first, ref_steps, _ = base
    err = None
    if not np.array_equal(first, eng.cur_tok.cpu().numpy()):
        err = "prefill token differs"
    acc = draft = 0
    hist = np.zeros(k + 1, dtype=np.int64)
    for s, (emitted, n_acc, seq_len) in enumerate(ref_steps):
        out = eng.step_host(eng.cur_tok.cpu()).clone()
        for b in range(B):
            cnt = int(out[b, k + 1])
            if err is None and cnt != len(emitted[b]):
                err = f"seed {seed} step {s} seq {b}: accepted count differs ({cnt} vs {len(emitted[b])})"
            if err is None and out[b, :cnt].tolist() != emitted[b]:
                err = f"seed {seed} step {s} seq {b}: tokens differ"
            if err is None and out[b, cnt:k + 1].tolist() != [-1] * (k + 1 - cnt):
                err = f"seed {seed} step {s} seq {b}: padding after the emitted tokens is not -1"
        if err is None and not np.array_equal(eng.seq_len.cpu().numpy(), seq_len):
            err = f"seed {seed} step {s}: sequence lengths differ after the step"
        for v in n_acc:
            hist[int(v)] += 1
        acc += int(n_acc.sum()); draft += B * k
    m = eng.metrics()
    if err is None and not (m["accepted_tokens"] == acc and m["draft_tokens"] == draft
                            and abs(m["draft_acceptance_rate"] - acc / draft) < 1e-12):
        err = f"seed {seed}: acceptance counters differ"
    ill = err is not None and _decided_by_rounding(model, prompts, B, k, steps, base)
    return err, ill, hist, m


def test_e2e_greedy_tokens_and_acceptance_match_oracle(cuda):
    """north_star: end-to-end greedy decode gives identical tokens and the same draft acceptance
    rate as the oracle-backed run of the same loop (tiny Llama config, random int4 weights).

    Attention, RoPE and SiLU are outside the bit-exact contract (fp32 with the GPU's own summation order / exp), and the
    W4A16 GEMMs accumulate in fp32 in tensor-core order against the oracle's float64: a tiny fraction of f16 values differs
    by one ulp, which the int4 quantisers of the draft path can amplify into a different arg max. A seed may therefore be
    skipped ONLY when it mismatches AND the oracle's own tokens change under a rounding-level perturbation of those ops
    (_decided_by_rounding); every other seed must match exactly, and at least two seeds must."""
    checked = 0
    for seed in range(3, 15):
        err, ill, _, _ = _e2e_case(cuda, seed)
        if err is not None and ill:
            continue
        assert err is None, f"{err} (and the oracle's tokens are stable under rounding-level jitter)"
        checked += 1
        if checked == 3:
            break
    assert checked >= 2, "fewer than two seeds matched the oracle"


def test_accept_path_matches_oracle_for_every_accepted_count(cuda):
    """The accepted-token path (n_acc > 0: several tokens emitted, seq_len += n_acc + 1, KV rows of rejected drafts
    overwritten by the next step). With random weights draft and verify almost never agree, so the model is made to
    agree by construction (`make_modes_agree`: the embedding dominates the residual stream); the run then sits near the
    reference's operating point (draft acceptance 0.955-0.961, vllm/spec_decode/metrics.py:171-179, figs/image-1.png).
    Bars: tokens, emitted counts, sequence lengths and the acceptance counters equal the oracle-backed loop on every
    seed that matches or is well-conditioned; acceptance >= 0.8 over the checked seeds; every accepted count
    0..k occurred. A seed may only be skipped when it diverges AND the oracle's own tokens change under rounding-level
    jitter of the ops outside the contract (_decided_by_rounding)."""
    k = 3
    hist = np.zeros(k + 1, dtype=np.int64)
    acc = draft = checked = 0
    for seed in range(20, 36):
        err, ill, h, m = _e2e_case(cuda, seed, shift=8, B=4, k=k, steps=6)
        if err is not None and ill:
            continue
        assert err is None, f"{err} (and the oracle's tokens are stable under rounding-level jitter)"
        hist += h
        acc += m["accepted_tokens"]; draft += m["draft_tokens"]
        checked += 1
        if checked >= 3 and (hist > 0).all():
            break
    assert checked >= 3, "fewer than three seeds could be checked"
    assert acc / draft >= 0.8, f"draft acceptance {acc / draft:.3f} < 0.8 (histogram {hist.tolist()})"
    assert (hist > 0).all(), f"accepted counts seen: {hist.tolist()} -- not every n_acc in 0..{k} occurred"


def test_e2e_one_layer_at_llama3_8b_shapes_matches_oracle(cuda):
    """The speculative loop on ONE decoder layer of Llama-3-8B's real shapes (6144/4096/28672 x 4096, 4096 x 14336:
    the split-k plans, deferred reductions and the 7-CTA rotation of configs[1]) against the oracle, with and without the
    agree-by-construction shift (8B shapes need 2^14: the block outputs grow with sqrt(K))."""
    from qspec_b200.model import QSpecLlamaConfig
    for shift in (None, 14):
        checked = 0
        for seed in range(40, 46):
            cfg = QSpecLlamaConfig(num_hidden_layers=1, vocab_size=2048, max_seq_len=64, name="8b-1layer")
            err, ill, hist, m = _e2e_case(cuda, seed, shift=shift, cfg=cfg, B=2, steps=3, P=6)
            if err is not None and ill:
                continue
            assert err is None, f"shift={shift}: {err} (and the oracle's tokens are stable under rounding-level jitter)"
            checked += 1
            if shift is not None:
                assert m["draft_acceptance_rate"] >= 0.5, f"acceptance {m['draft_acceptance_rate']:.2f} with agreeing modes"
            break
        assert checked == 1, f"shift={shift}: no seed could be checked"


def test_step_refuses_to_run_past_the_kv_cache(cuda):
    """positions are bounded on the host. A step touches positions seq_len .. seq_len + k; the engine raises
    before any of them could reach max_seq_len, and everything decoded before that equals a run with a larger cache."""
    from qspec_b200.model import QSpecLlamaConfig
    B, k, P = 2, 3, 10
    runs = []
    for S in (40, 128):
        cfg, model, eng = _tiny_engine(cuda, graph=False, seed=3, B=B, k=k, cfg=QSpecLlamaConfig.tiny(max_seq_len=S), shift=8)
        g = torch.Generator().manual_seed(5)
        eng.prefill(torch.randint(0, cfg.vocab_size, (B, P), generator=g).to(cuda))
        toks = []
        try:
            for _ in range(12):
                eng.step()
                torch.cuda.synchronize()
                toks.append(eng.d_out_pack.cpu().clone())
                assert int(eng.seq_len.max()) + 0 <= S
            raised = False
        except RuntimeError as e:
            assert "KV cache exhausted" in str(e)
            raised = True
        runs.append((toks, raised, int(eng.seq_len.max())))
    small, large = runs
    assert small[1] and not large[1]
    assert small[2] + k + 1 > 40 and small[2] <= 40            # stopped only when the next step could not fit
    assert len(small[0]) >= 3
    for a, b in zip(small[0], large[0]):
        assert torch.equal(a, b)


@pytest.mark.parametrize("Hq,Hkv,S,per_seq", [(32, 8, 300, 4), (8, 8, 64, 1), (40, 8, 200, 3), (64, 8, 130, 9), (4, 2, 1100, 2)])
def test_seq_attention_matches_torch_fp32_and_decode_attention(cuda, Hq, Hkv, S, per_seq):
    """The (sequence, kv head) attention kernel of large-batch passes: every token of a sequence in the pass (verify: k + 1,
    here also more than one chunk) against the cache, causal by position; fp32 math, f16 output: <= 2e-3 absolute on O(1)
    values against torch fp32, and the same against the per-(token, head) kernel it replaces."""
    from qspec_b200 import ops
    D, B = 128, 5
    g = torch.Generator().manual_seed(Hq * 7 + S)
    ck = torch.randn(B, S, Hkv, D, generator=g).half().to(cuda)
    cv = torch.randn(B, S, Hkv, D, generator=g).half().to(cuda)
    seqs, poss = [], []
    for s_ in (3, 0, 4):                                      # sequence 1 and 2 have no token in the pass
        base = int(torch.randint(0, S - per_seq, (1,), generator=g))
        for i in range(per_seq):
            seqs.append(s_); poss.append(base + i)
    perm = torch.randperm(len(seqs), generator=g).tolist()    # tokens of a sequence need not be adjacent
    seq = torch.tensor([seqs[i] for i in perm], dtype=torch.int32, device=cuda)
    pos = torch.tensor([poss[i] for i in perm], dtype=torch.int32, device=cuda)
    T = pos.numel()
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).half().to(cuda)
    assert ops.seq_attention_supported(T, Hq, Hkv, D, S)
    out = torch.empty(T, Hq * D, dtype=torch.float16, device=cuda)
    ops.seq_attention(qkv, ck, cv, pos, seq, out, Hq, Hkv, D, 1.0 / np.sqrt(D))
    old = torch.empty_like(out)
    ops.decode_attention(qkv, ck, cv, pos, seq, old, Hq, Hkv, D, 1.0 / np.sqrt(D))
    q = qkv.view(T, Hq + 2 * Hkv, D)[:, :Hq].float()
    for t in range(T):
        n = int(pos[t]) + 1
        K = ck[seq[t], :n].float().repeat_interleave(Hq // Hkv, dim=1)
        V = cv[seq[t], :n].float().repeat_interleave(Hq // Hkv, dim=1)
        sc = torch.einsum("hd,nhd->hn", q[t], K) / np.sqrt(D)
        ref = torch.einsum("hn,nhd->hd", sc.softmax(-1), V)
        assert (out[t].view(Hq, D).float() - ref).abs().max() <= 2e-3
    assert (out.float() - old.float()).abs().max() <= 2e-3


def test_engine_is_deterministic_without_caller_synchronisation(cuda, monkeypatch):
    """Size-independent property at BASELINE configs[2] width (64 sequences, k = 3, Llama-3-8B layer shapes, 2 layers): decoding
    the same prompts twice -- prefill, capture and steps queued back to back with NO synchronize in between, under the CUDA graph
    and eagerly -- emits the same tokens. (The engine's stream is non-blocking: before QSpecEngine._join_caller_stream the first
    step could overtake prefill() and append draft K/V rows at stale cache positions; scripts/gpu_determinism.py.)"""
    from qspec_b200.engine import QSpecEngine
    from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel
    B, k, steps = 64, 3, 6
    cfg = QSpecLlamaConfig.llama3_8b(num_hidden_layers=2, vocab_size=4096, max_seq_len=256)
    model = QSpecLlamaModel(cfg, max_batch=B, max_tokens=B * (k + 1), device=cuda, seed=5)
    prompts = torch.randint(0, cfg.vocab_size, (B, 48), generator=torch.Generator().manual_seed(1)).to(cuda)
    for shift in (None, 14):
        model.make_modes_agree(shift)
        runs = []
        for graph, c8_min in ((True, None), (True, None), (False, None), (True, "100000")):
            # last run: the draft passes through the packed int4 route instead of the int8 container (QSPEC_C8_MIN_T)
            if c8_min is None:
                monkeypatch.delenv("QSPEC_C8_MIN_T", raising=False)
            else:
                monkeypatch.setenv("QSPEC_C8_MIN_T", c8_min)
            eng = QSpecEngine(model, B, k, use_cuda_graph=graph)
            eng.prefill(prompts)
            eng.capture()
            outs = []
            for _ in range(steps):
                eng.step()
                with torch.cuda.stream(eng.stream):
                    outs.append(eng.d_out_pack.clone())
            eng.stream.synchronize()
            runs.append((torch.stack(outs).cpu(), eng.metrics()))
        monkeypatch.delenv("QSPEC_C8_MIN_T", raising=False)
        for name, (o, m) in zip(("second graph run", "eager run", "packed-activation run"), runs[1:]):
            assert torch.equal(o, runs[0][0]), f"shift {shift}: {name} emitted different tokens"
        if shift is not None:
            assert runs[0][1]["draft_acceptance_rate"] >= 0.5
    model.make_modes_agree(None)
