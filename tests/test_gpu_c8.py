"""int8 container of the draft path's quantised activations ("_c8" entry points, include/qspec_b200.h): the same int4 VALUES one
per byte instead of two per byte, consumed by the GEMM's int8-activation instantiation. Bars: every producer writes exactly the
values of its packed twin (which the other tests tie to the oracle bit for bit) with the same scales / sums / residual; the
GEMM on the container gives the packed GEMM's output bit for bit (plain, and deferred-reduction + consumer); a whole pass of
the decoder gives the same hidden state with and without the container."""
import numpy as np
import pytest
import torch

from oracle import qspec_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from qspec_b200 import ops as _ops
    return _ops


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def unpack(p):
    """packed int4 [rows, K/2] (even k in the low nibble, two's complement) -> int8 [rows, K]"""
    p = p.view(torch.int8)
    lo = (p << 4) >> 4
    hi = p >> 4
    return torch.stack((lo, hi), dim=-1).reshape(p.shape[0], -1)


@pytest.mark.parametrize("rows,H,D", [(64, 32, 128), (5, 8, 128), (3, 64, 128)])
def test_hadamard_heads_c8(ops, rows, H, D):
    x = cu(O.make_act(rows, H * D, 910 + rows, heavy_tail=True))
    q = torch.empty(rows, H * D // 2, dtype=torch.int8, device="cuda")
    q8 = torch.full((rows, H * D), 99, dtype=torch.int8, device="cuda")
    s, s8 = (torch.empty(rows, dtype=torch.float16, device="cuda") for _ in range(2))
    ops.hadamard_heads(x, H, D, 0.1767578125, out_q=q, scale_out=s)
    ops.hadamard_heads(x, H, D, 0.1767578125, out_q=q8, scale_out=s8, c8=True)
    assert torch.equal(q8, unpack(q)) and torch.equal(s8, s)
    assert int(q8.min()) >= -8 and int(q8.max()) <= 7
    with pytest.raises(ValueError):
        ops.hadamard_heads(x, H, D, 0.1767578125, out_q=q, scale_out=s, c8=True)       # packed-sized buffer for the container


@pytest.mark.parametrize("rows,n", [(64, 512), (4, 512), (3, 128), (2, 1024), (5, 64)])
def test_hadamard_k28_c8(ops, rows, n):
    L = 28 * n
    gu = cu((O.make_act(rows, 2 * L, 920 + n).astype(np.float32) * 1.5).astype(np.float16))
    q = torch.empty(rows, L // 2, dtype=torch.int8, device="cuda")
    q8 = torch.full((rows, L), 99, dtype=torch.int8, device="cuda")
    s, s8 = (torch.empty(rows, dtype=torch.float16, device="cuda") for _ in range(2))
    ops.hadamard_k28(None, n, 0.00835418701171875, gate_up=gu, out_q=q, scale_out=s)
    ops.hadamard_k28(None, n, 0.00835418701171875, gate_up=gu, out_q=q8, scale_out=s8, c8=True)
    assert torch.equal(q8, unpack(q)) and torch.equal(s8, s)


@pytest.mark.parametrize("M,N,K", [(64, 4096, 4096), (64, 4096, 14336), (48, 512, 1792), (256, 4096, 4096)])
def test_gemm_on_container_equals_packed_gemm(ops, M, N, K):
    """W4A4 through the int8-activation instantiation (acc = 16 sum, >> 4) == the packed route (acc = 256 sum, >> 8): plain GEMM,
    and deferred reduction (qspec_gemm_w4a8_split) + consumer (kind 2, packed and container outputs)."""
    rng = np.random.default_rng(5 + M + N + K)
    w = cu(rng.integers(-128, 128, size=(N, K // 2), dtype=np.int8))
    sW = cu(((0.5 + rng.random(N)) * 2.0 ** -7).astype(np.float16))
    bias = cu((rng.standard_normal(N) * 0.05).astype(np.float16))
    a = cu(rng.integers(-128, 128, size=(M, K // 2), dtype=np.int8))
    a8 = unpack(a).contiguous()
    sA = cu((0.01 + rng.random(M) * 0.1).astype(np.float16))
    out, out8 = (torch.empty(M, N, dtype=torch.float16, device="cuda") for _ in range(2))
    ops.rowwise_scaled_linear_cutlass_s4s4_unified(a, sA, w, sW, bias, out)
    ops.rowwise_scaled_linear_cutlass_s8s4(a8, sA, w, sW, bias, out8)
    assert torch.equal(out8, out)
    split = ops.gemm_split_factor(2, M, N, K)
    if split not in range(1, 9):
        return
    part = torch.empty(8 * M * N, dtype=torch.int32, device="cuda")
    x0 = cu(O.make_act(M, N, 31 + M))
    # reference: plain packed GEMM, then add + norm + quant
    x1 = x0.clone()
    q1 = torch.empty(M, N // 2, dtype=torch.int8, device="cuda")
    s1, i1 = (torch.empty(M, dtype=torch.float16, device="cuda") for _ in range(2))
    ops.add_rms_norm_fuse_sum_i4(q1, x1, out, i1, s1, 1e-5)
    sp = ops.gemm_w4a4_c8_split(a8, sA, w, sW, bias, part)
    assert sp.split == split and sp.w4a16 == 2
    x2 = x0.clone()
    q2 = torch.empty(M, N, dtype=torch.int8, device="cuda")
    s2, i2 = (torch.empty(M, dtype=torch.float16, device="cuda") for _ in range(2))
    ops.split_add_rms_norm_fuse_sum_i4(q2, x2, sp, i2, s2, 1e-5, c8=True)
    assert torch.equal(x2, x1) and torch.equal(q2, unpack(q1)) and torch.equal(s2, s1) and torch.equal(i2, i1)
    x3 = x0.clone()
    q3 = torch.empty_like(q1)
    ops.split_add_rms_norm_fuse_sum_i4(q3, x3, sp, i2, s2, 1e-5)                  # kind-2 partials, packed output
    assert torch.equal(x3, x1) and torch.equal(q3, q1)
    nrm, nrm_ref = torch.empty_like(x0), torch.empty_like(x0)
    x4, x5 = x0.clone(), x0.clone()
    ops.split_add_rms_norm_fuse_sum_fp16(nrm, x4, sp, 1e-5)
    ops.add_rms_norm_fuse_sum_fp16(nrm_ref, x5, out, 1e-5)
    assert torch.equal(nrm, nrm_ref)


@pytest.mark.parametrize("world,rank", [(2, 1), (8, 3)])
def test_tp_add_norm_c8(ops, world, rank):
    from test_gpu_kernels import _OneGpuPeers
    rows, K = 64, 4096
    parts = [cu(O.make_act(rows, K, 2600 + r, heavy_tail=(r == 1))) for r in range(world)]
    x0 = cu(O.make_act(rows, K, 2500))
    res = []
    for c8 in (False, True):
        peers = _OneGpuPeers(parts, rank)
        x = x0.clone()
        q = torch.empty(rows, K if c8 else K // 2, dtype=torch.int8, device="cuda")
        s, isum = (torch.empty(rows, dtype=torch.float16, device="cuda") for _ in range(2))
        ops.tp_add_rms_norm_fuse_sum_i4(q, x, peers, 0, isum, s, 1e-5, c8=c8)
        assert peers.state.tolist() == [1, 0, 0, 0]
        res.append((x, q, s, isum))
    assert torch.equal(res[1][0], res[0][0]) and torch.equal(res[1][1], unpack(res[0][1]))
    assert torch.equal(res[1][2], res[0][2]) and torch.equal(res[1][3], res[0][3])


@pytest.mark.parametrize("T,w4a4", [(64, True), (96, True), (256, True)])
def test_decoder_pass_same_with_and_without_container(T, w4a4, monkeypatch):
    """Two layers of Llama-3-8B shapes, one draft-mode pass of T tokens: the hidden state with the container (default from 48
    tokens) equals the packed route (QSPEC_C8_MIN_T beyond T) bit for bit."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel
    dev = torch.device("cuda:0")
    cfg = QSpecLlamaConfig.llama3_8b(num_hidden_layers=2, vocab_size=2048, max_seq_len=64)
    m = QSpecLlamaModel(cfg, max_batch=T, max_tokens=T, device=dev, seed=3)
    m.cache_k.zero_(); m.cache_v.zero_()
    g = torch.Generator().manual_seed(T)
    ids = torch.randint(0, cfg.vocab_size, (T,), generator=g).to(dev)
    pos = torch.zeros(T, dtype=torch.int32, device=dev)
    seq = torch.arange(T, dtype=torch.int32, device=dev)
    monkeypatch.setenv("QSPEC_C8_MIN_T", "100000")
    ref = m.forward(ids, pos, seq, w4a4=w4a4).clone()
    monkeypatch.setenv("QSPEC_C8_MIN_T", "48")
    from qspec_b200 import ops
    l0 = ops.launch_count()
    got = m.forward(ids, pos, seq, w4a4=w4a4).clone()
    assert torch.equal(got, ref)
    assert torch.isfinite(got.float()).all()


@pytest.mark.parametrize("tp", [2, 8])
def test_tp_decoder_pass_same_with_and_without_container(tp, monkeypatch):
    """All ranks of a TP model on one GPU (tests/tp_emulation.py), one draft-mode pass of 64 tokens: with the container the
    peer-reduction norm writes [T, H] int8 (qspec_tp_add_norm_quant_i4_c8) and every sharded GEMM takes the int8 route; hidden
    states equal the packed route bit for bit on every rank."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tp_emulation import build_shards, lockstep_forward
    from qspec_b200.model import QSpecLlamaConfig
    dev = torch.device("cuda:0")
    T = 64
    cfg = QSpecLlamaConfig.llama3_8b(num_hidden_layers=2, vocab_size=2048, max_seq_len=64, name="8b-2layer")
    shards = build_shards(cfg, tp, max_batch=T, max_tokens=T, device=dev, seed=4)
    for m in shards:
        m.cache_k.zero_(); m.cache_v.zero_()
    g = torch.Generator().manual_seed(7)
    ids = torch.randint(0, cfg.vocab_size, (T,), generator=g).to(dev)
    pos = torch.zeros(T, dtype=torch.int32, device=dev)
    seq = torch.arange(T, dtype=torch.int32, device=dev)
    monkeypatch.setenv("QSPEC_C8_MIN_T", "100000")
    ref = [h.clone() for h in lockstep_forward(shards, ids, pos, seq, True)]
    monkeypatch.setenv("QSPEC_C8_MIN_T", "48")
    got = [h.clone() for h in lockstep_forward(shards, ids, pos, seq, True)]
    torch.cuda.synchronize()
    for r in range(tp):
        assert torch.equal(got[r], ref[r]), f"rank {r}"
        assert torch.equal(got[r], got[0])
    for m in shards:
        assert m.peers.timeouts() == 0
