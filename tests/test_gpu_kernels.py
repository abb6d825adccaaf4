"""GPU parity tests: every CUDA kernel, called through the C-ABI (qspec_b200.ops -> ctypes ->
libqspec_b200.so), against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): quantised activations (packed bytes + f16 scales) and W4A4
int32 accumulators bit-exact; f16/bf16 GEMM outputs within 1e-2 relative (max|d|/max|ref|).
"""
import os

import numpy as np
import pytest
import torch

from oracle import qspec_oracle as O

pytestmark = pytest.mark.gpu

REL_TOL = 1e-2   # north_star: fp16/bf16 outputs of both modes within 1e-2 relative


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from qspec_b200 import ops as _ops
    return _ops


def cu(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.view(dtype)
    return t.cuda()


def special_rows(K, seed):
    """heavy-tail rows, an all-zero row, a row with +-65504, a constant row."""
    x = O.make_act(6, K, seed, heavy_tail=True)
    x[2] = 0
    x[3, 5] = 65504.0
    x[3, 9] = -65504.0
    x[4] = np.float16(0.37)
    return x


# ---------------------------------------------------------------------------------- (A) norm
@pytest.mark.parametrize("K", [128, 2048, 4096, 8192, 14336])
def test_norm_quant_i4_bit_exact(ops, K):
    x = special_rows(K, 2000 + K)
    x[3, 5] = 300.0; x[3, 9] = -300.0      # keep (x-mean)^2 finite in fp32
    ref_q, ref_s, ref_sum = O.norm_quant_i4(x, 1e-5)
    xd = cu(x)
    q = torch.empty(x.shape[0], K // 2, dtype=torch.int8, device="cuda")
    s = torch.empty(x.shape[0], dtype=torch.float16, device="cuda")
    isum = torch.empty_like(s)
    ops.rms_norm_general_fuse_sum_i4(q, xd, isum, s, 1e-5, True)
    assert np.array_equal(q.cpu().numpy().view(np.uint8), ref_q)
    assert np.array_equal(s.cpu().numpy(), ref_s)
    assert np.array_equal(isum.cpu().numpy(), ref_sum)


@pytest.mark.parametrize("K", [64, 4096, 8192])
def test_norm_f16_bit_exact(ops, K):
    x = special_rows(K, 2100 + K)
    x[3, 5] = 300.0; x[3, 9] = -300.0
    xd = cu(x)
    out = torch.empty_like(xd)
    ops.rms_norm_general_fuse_sum_fp16(out, xd, 1e-5)
    assert np.array_equal(out.cpu().numpy(), O.norm_f16(x, 1e-5))


@pytest.mark.parametrize("K", [64, 4096, 8192])
def test_add_norm_fused_equals_add_then_norm(ops, K):
    """Fused residual add + norm (quarot_llama.py:380,389): x is updated in place and both outputs are bit-identical to
    the oracle's f16 add followed by the oracle norm."""
    x = O.make_act(6, K, 2300 + K)
    d = O.make_act(6, K, 2400 + K, heavy_tail=True)
    xs = (x.astype(np.float32) + d.astype(np.float32)).astype(np.float16)
    ref_q, ref_s, ref_sum = O.norm_quant_i4(xs, 1e-5)
    xd, dd = cu(x), cu(d)
    q = torch.empty(6, K // 2, dtype=torch.int8, device="cuda")
    s = torch.empty(6, dtype=torch.float16, device="cuda")
    isum = torch.empty_like(s)
    ops.add_rms_norm_fuse_sum_i4(q, xd, dd, isum, s, 1e-5)
    assert np.array_equal(xd.cpu().numpy(), xs)
    assert np.array_equal(q.cpu().numpy().view(np.uint8), ref_q)
    assert np.array_equal(s.cpu().numpy(), ref_s)
    assert np.array_equal(isum.cpu().numpy(), ref_sum)
    xd2 = cu(x)
    out = torch.empty_like(xd2)
    ops.add_rms_norm_fuse_sum_fp16(out, xd2, dd, 1e-5)
    assert np.array_equal(xd2.cpu().numpy(), xs)
    assert np.array_equal(out.cpu().numpy(), O.norm_f16(xs, 1e-5))


class _OneGpuPeers:
    """Stand-in for qspec_b200.parallel.PeerPartials with every rank's partials on ONE GPU and all arrival flags
    already set: exercises the rank-ordered sum, the epoch words and the C-ABI of qspec_tp_add_norm_* without
    any kernel waiting on another one."""

    def __init__(self, parts, rank):
        import ctypes
        self.world, self.rank = len(parts), rank
        self.max_tokens, self.hidden = parts[0].shape
        self.buf = torch.stack(parts, 0)                                    # [world, rows, K]; which=0 only
        self.flag_arrays = torch.full((self.world, 8), 1 << 20, dtype=torch.int32, device="cuda")
        self.state = torch.zeros(4, dtype=torch.int32, device="cuda")
        self._partials = [(ctypes.c_void_p * self.world)(*[self.buf[r].data_ptr() for r in range(self.world)])]
        self._flags = (ctypes.c_void_p * self.world)(*[self.flag_arrays[r].data_ptr() for r in range(self.world)])


@pytest.mark.parametrize("world,rank,K", [(2, 0, 4096), (2, 1, 4096), (4, 2, 8192), (8, 7, 64)])
def test_tp_fused_reduce_add_norm(ops, world, rank, K):
    """Peer partial sums -> rank-ordered fp32 sum, one rounding (the all-reduced o_proj/down_proj output) -> residual
    add -> norm, in one launch (layers/linear.py:1143-1144 + quarot_llama.py:380,389)."""
    rows = 5
    x = O.make_act(rows, K, 2500 + K)
    parts = [O.make_act(rows, K, 2600 + K + r, heavy_tail=(r == 1)) for r in range(world)]
    acc = parts[0].astype(np.float32)
    for r in range(1, world):
        acc = acc + parts[r].astype(np.float32)
    delta = acc.astype(np.float16)
    xs = (x.astype(np.float32) + delta.astype(np.float32)).astype(np.float16)
    ref_q, ref_s, ref_sum = O.norm_quant_i4(xs, 1e-5)
    peers = _OneGpuPeers([cu(p) for p in parts], rank)
    xd = cu(x)
    q = torch.empty(rows, K // 2, dtype=torch.int8, device="cuda")
    s = torch.empty(rows, dtype=torch.float16, device="cuda")
    isum = torch.empty_like(s)
    ops.tp_add_rms_norm_fuse_sum_i4(q, xd, peers, 0, isum, s, 1e-5)
    assert np.array_equal(xd.cpu().numpy(), xs)
    assert np.array_equal(q.cpu().numpy().view(np.uint8), ref_q)
    assert np.array_equal(s.cpu().numpy(), ref_s)
    assert np.array_equal(isum.cpu().numpy(), ref_sum)
    assert peers.state.tolist() == [1, 0, 0, 0]                             # epoch advanced, no time-out
    for r in range(world):                                                  # the arrival was stored to every peer
        if r != rank:
            assert int(peers.flag_arrays[r, rank]) == 1
    xd2 = cu(x)
    out = torch.empty_like(xd2)
    ops.tp_add_rms_norm_fuse_sum_fp16(out, xd2, peers, 0, 1e-5)
    assert np.array_equal(xd2.cpu().numpy(), xs)
    assert np.array_equal(out.cpu().numpy(), O.norm_f16(xs, 1e-5))
    assert peers.state.tolist() == [2, 0, 0, 0]


@pytest.mark.parametrize("M,N,K", [(4, 4096, 4096), (16, 4096, 14336), (5, 512, 1792), (13, 512, 256), (4, 8192, 8192)])
@pytest.mark.parametrize("w4a16", [False, True])
def test_deferred_reduction_equals_gemm_then_add_norm(ops, M, N, K, w4a16):
    """qspec_gemm_*_split (raw per-k-slice partial sums) + qspec_split_add_norm_* against the plain GEMM followed by the
    fused add+norm: residual, packed int4 bytes, scales and sums bit-identical, in both modes (at these decode shapes the
    plain call splits k the same way, so the fp32 sums of the W4A16 mode are identical too)."""
    split = ops.gemm_split_factor(w4a16, M, N, K)
    assert 1 <= split <= 8
    rng = np.random.default_rng(77 + M + N + K)
    w = cu(rng.integers(-128, 128, size=(N, K // 2), dtype=np.int8))
    sW = cu(((0.5 + rng.random(N)) * 2.0 ** -7).astype(np.float16))
    x0 = cu(O.make_act(M, N, 31 + M))
    part = torch.empty(8 * M * N, dtype=torch.int32, device="cuda")
    out = torch.empty(M, N, dtype=torch.float16, device="cuda")
    if w4a16:
        a = cu(O.make_act(M, K, 5 + K))
        ops.gemm_w4a16(a, w, sW, None, out)
        sp = ops.gemm_w4a16_split(a, w, sW, None, part)
    else:
        a = cu(rng.integers(-128, 128, size=(M, K // 2), dtype=np.int8))
        sA = cu((0.01 + rng.random(M) * 0.1).astype(np.float16))
        ops.rowwise_scaled_linear_cutlass_s4s4_unified(a, sA, w, sW, None, out)
        sp = ops.gemm_w4a4_split(a, sA, w, sW, None, part)
    assert sp.split == split
    q1, q2 = (torch.empty(M, N // 2, dtype=torch.int8, device="cuda") for _ in range(2))
    s1, s2, i1, i2 = (torch.empty(M, dtype=torch.float16, device="cuda") for _ in range(4))
    x1, x2 = x0.clone(), x0.clone()
    ops.add_rms_norm_fuse_sum_i4(q1, x1, out, i1, s1, 1e-5)
    ops.split_add_rms_norm_fuse_sum_i4(q2, x2, sp, i2, s2, 1e-5)
    assert torch.equal(x1, x2) and torch.equal(q1, q2) and torch.equal(s1, s2) and torch.equal(i1, i2)
    x1, x2 = x0.clone(), x0.clone()
    o1, o2 = torch.empty_like(x0), torch.empty_like(x0)
    ops.add_rms_norm_fuse_sum_fp16(o1, x1, out, 1e-5)
    ops.split_add_rms_norm_fuse_sum_fp16(o2, x2, sp, 1e-5)
    assert torch.equal(x1, x2) and torch.equal(o1, o2)


@pytest.mark.parametrize("M,N,K,bf16,with_bias", [(64, 4096, 4096, False, True), (128, 512, 1792, True, False),
                                                  (7, 4096, 4096, True, True), (33, 1024, 512, False, False)])
@pytest.mark.parametrize("w4a16", [False, True])
def test_deferred_reduction_bf16_bias_tall(ops, M, N, K, bf16, with_bias, w4a16):
    """Deferred reduction with bias, bf16 storage and token tiles up to 128 rows: the f16/bf16 norm output and the updated
    residual equal the plain GEMM followed by the fused add+norm, bit for bit."""
    dt = torch.bfloat16 if bf16 else torch.float16
    split = ops.gemm_split_factor(w4a16, M, N, K)
    assert 1 <= split <= 8
    g = torch.Generator().manual_seed(M * 31 + N + K)
    w = torch.randint(-128, 128, (N, K // 2), generator=g, dtype=torch.int8).cuda()
    sW = ((0.5 + torch.rand(N, generator=g)) * 2.0 ** -7).to(dt).cuda()
    bias = (torch.randn(N, generator=g) * 0.1).to(dt).cuda() if with_bias else None
    x0 = torch.randn(M, N, generator=g).to(dt).cuda()
    part = torch.empty(8 * M * N, dtype=torch.int32, device="cuda")
    out = torch.empty(M, N, dtype=dt, device="cuda")
    if w4a16:
        a = torch.randn(M, K, generator=g).to(dt).cuda()
        ops.gemm_w4a16(a, w, sW, bias, out)
        sp = ops.gemm_w4a16_split(a, w, sW, bias, part)
    else:
        a = torch.randint(-128, 128, (M, K // 2), generator=g, dtype=torch.int8).cuda()
        sA = (0.01 + torch.rand(M, generator=g) * 0.1).to(dt).cuda()
        ops.rowwise_scaled_linear_cutlass_s4s4_unified(a, sA, w, sW, bias, out)
        sp = ops.gemm_w4a4_split(a, sA, w, sW, bias, part)
    x1, x2 = x0.clone(), x0.clone()
    o1, o2 = torch.empty_like(x0), torch.empty_like(x0)
    ops.add_rms_norm_fuse_sum_fp16(o1, x1, out, 1e-5)
    ops.split_add_rms_norm_fuse_sum_fp16(o2, x2, sp, 1e-5)
    if not w4a16:
        assert torch.equal(x1, x2) and torch.equal(o1, o2)      # int32 partial sums: exact whatever the k-split
    else:
        # fp32 partial sums: identical only when both calls split k the same way (the plain call may keep a tile whole)
        for u, v in ((x1, x2), (o1, o2)):
            tol = float(v.float().abs().max()) * (2.0 ** -6 if bf16 else 2.0 ** -9)       # two ulps of the largest value
            assert (u.float() - v.float()).abs().max() <= tol


def test_norm_bf16_close(ops):
    x = torch.randn(5, 4096, device="cuda").to(torch.bfloat16)
    out = torch.empty_like(x)
    ops.rms_norm_general_fuse_sum_fp16(out, x, 1e-5)
    xf = x.float()
    ref = (xf - xf.mean(-1, keepdim=True)) * torch.rsqrt(xf.var(-1, unbiased=False, keepdim=True) + 1e-5)
    assert (out.float() - ref).abs().max() <= 2e-2


# ------------------------------------------------------------------------------ quantiser
@pytest.mark.parametrize("K", [32, 4096, 14336])
def test_sym_quant_bit_exact(ops, K):
    x = special_rows(K, 2200 + K)
    ref_q, ref_s = O.sym_quant_i4(x, 1.0)
    q = torch.empty(x.shape[0], K // 2, dtype=torch.int8, device="cuda")
    s = torch.empty(x.shape[0], dtype=torch.float16, device="cuda")
    ops.fuse_sym_quant(cu(x), s, q, 1.0)
    assert np.array_equal(s.cpu().numpy(), ref_s)
    assert np.array_equal(q.cpu().numpy().view(np.uint8), ref_q)


def test_sym_quant_clip_ratio(ops):
    x = O.make_act(4, 1024, 77)
    ref_q, ref_s = O.sym_quant_i4(x, 0.9)
    q = torch.empty(4, 512, dtype=torch.int8, device="cuda")
    s = torch.empty(4, dtype=torch.float16, device="cuda")
    ops.fuse_sym_quant(cu(x), s, q, 0.9)
    assert np.array_equal(s.cpu().numpy(), ref_s)
    assert np.array_equal(q.cpu().numpy().view(np.uint8), ref_q)


# ------------------------------------------------------------------------------ (B) Hadamard
@pytest.mark.parametrize("n", [8, 16, 32, 64, 128, 256, 512, 1024, 2048])
def test_fht_bit_exact(ops, n):
    rows = 37
    x = O.make_act(rows, n, 2300 + n)
    scale = float(np.float32(1.0) / np.sqrt(np.float32(n)))
    out = torch.empty(rows, n, dtype=torch.float16, device="cuda")
    ops.faster_fast_hadamard_transform(cu(x), scale, out)
    assert np.array_equal(out.cpu().numpy(), O.fht(x, scale))


def test_fht_o_proj_shape_matches_had_test(ops):
    # third-party/QuaRot/had_test.py:14-28 shape [T*128, 32]
    x = O.make_act(16 * 128, 32, 5)
    scale = float(1.0 / torch.tensor(32.0).sqrt())
    out = torch.empty(16 * 128, 32, dtype=torch.float16, device="cuda")
    ops.faster_fast_hadamard_transform(cu(x), scale, out)
    assert np.array_equal(out.cpu().numpy(), O.fht(x, scale))


@pytest.mark.parametrize("H,D", [(32, 128), (64, 128), (16, 128), (8, 128), (4, 128), (2, 128), (32, 64)])
def test_hadamard_heads_fused_bit_exact(ops, H, D):
    rows = 5
    x = O.make_act(rows, H * D, 2400 + H)
    x[1] = 0
    scale = float(np.float32(1.0) / np.sqrt(np.float32(H)))
    ref = O.hadamard_heads(x, H, D, scale)
    ref_q, ref_s = O.sym_quant_i4(ref, 1.0)
    out = torch.empty(rows, H * D, dtype=torch.float16, device="cuda")
    q = torch.empty(rows, H * D // 2, dtype=torch.int8, device="cuda")
    s = torch.empty(rows, dtype=torch.float16, device="cuda")
    ops.hadamard_heads(cu(x), H, D, scale, out=out, out_q=q, scale_out=s)
    assert np.array_equal(out.cpu().numpy(), ref)
    assert np.array_equal(s.cpu().numpy(), ref_s)
    assert np.array_equal(q.cpu().numpy().view(np.uint8), ref_q)
    # quant-only variant (the draft path writes no f16 intermediate)
    q2 = torch.empty_like(q); s2 = torch.empty_like(s)
    ops.hadamard_heads(cu(x), H, D, scale, out_q=q2, scale_out=s2)
    assert torch.equal(q2, q) and torch.equal(s2, s)


@pytest.mark.parametrize("n", [64, 128, 256, 512, 1024])
def test_hadamard_k28_fused_bit_exact(ops, n):
    rows = 4
    L = 28 * n
    x = O.make_act(rows, L, 2500 + n)
    x[2] = 0
    scale = float(np.float32(1.0) / np.sqrt(np.float32(L)))
    ref = O.hadamard_k28(x, n, scale)
    ref_q, ref_s = O.sym_quant_i4(ref, 1.0)
    out = torch.empty(rows, L, dtype=torch.float16, device="cuda")
    q = torch.empty(rows, L // 2, dtype=torch.int8, device="cuda")
    s = torch.empty(rows, dtype=torch.float16, device="cuda")
    ops.hadamard_k28(cu(x), n, scale, out=out, out_q=q, scale_out=s)
    assert np.array_equal(out.cpu().numpy(), ref)
    assert np.array_equal(s.cpu().numpy(), ref_s)
    assert np.array_equal(q.cpu().numpy().view(np.uint8), ref_q)


def test_hadamard_k28_with_silu_prologue(ops):
    rows, n = 3, 512
    inter = 28 * n
    gu = (O.make_act(rows, 2 * inter, 31).astype(np.float32) * 1.5).astype(np.float16)
    scale = float(np.float32(1.0) / np.sqrt(np.float32(inter)))
    # exp differs by ulps between libm and CUDA: run SiLU*up on the GPU kernel, then compare the
    # fused prologue against rotation of THAT (bit-exact), and the SiLU itself with tolerance.
    act = torch.empty(rows, inter, dtype=torch.float16, device="cuda")
    ops.silu_mul(cu(gu), act)
    np.testing.assert_allclose(act.cpu().numpy().astype(np.float32), O.silu_mul(gu, inter).astype(np.float32),
                               rtol=2e-3, atol=1e-3)
    ref = O.hadamard_k28(act.cpu().numpy(), n, scale)
    out = torch.empty(rows, inter, dtype=torch.float16, device="cuda")
    ops.hadamard_k28(None, n, scale, gate_up=cu(gu), out=out)
    assert np.array_equal(out.cpu().numpy(), ref)


# ------------------------------------------------------------------------------ W4A4 GEMM
W4A4_SHAPES = [(1, 4096, 4096), (4, 4096, 4096), (16, 4096, 4096), (16, 6144, 4096), (7, 4096, 14336),
               (64, 4096, 4096), (200, 512, 1024), (300, 640, 256), (16, 136, 128)]


@pytest.mark.parametrize("M,N,K", W4A4_SHAPES)
def test_w4a4_acc_int32_bit_exact(ops, M, N, K):
    rng = np.random.default_rng(M * 7 + N + K)
    a = O.pack_i4(rng.integers(-8, 8, (M, K), dtype=np.int8))
    w = O.pack_i4(rng.integers(-8, 8, (N, K), dtype=np.int8))
    acc = ops.gemm_w4a4_acc_i32(cu(a, torch.int8), cu(w, torch.int8))
    torch.cuda.synchronize()
    assert np.array_equal(acc.cpu().numpy(), O.gemm_w4a4_acc(a, w))
    # second call: the self-cleaning split-K workspace must give the same answer
    acc2 = ops.gemm_w4a4_acc_i32(cu(a, torch.int8), cu(w, torch.int8))
    assert torch.equal(acc, acc2)


def test_w4a4_acc_extremes(ops):
    M, N, K = 16, 256, 14336
    a = O.pack_i4(np.full((M, K), -8, dtype=np.int8))
    w = O.pack_i4(np.full((N, K), -8, dtype=np.int8))
    acc = ops.gemm_w4a4_acc_i32(cu(a, torch.int8), cu(w, torch.int8))
    assert (acc == 64 * K).all()
    w2 = O.pack_i4(np.full((N, K), 7, dtype=np.int8))
    acc = ops.gemm_w4a4_acc_i32(cu(a, torch.int8), cu(w2, torch.int8))
    assert (acc == -56 * K).all()


@pytest.mark.parametrize("dtype", ["f16", "bf16"])
@pytest.mark.parametrize("use_bias", [False, True])
@pytest.mark.parametrize("M,N,K", [(2, 512, 128), (3, 2048, 2048), (4, 3584, 640), (13, 8704, 8576),
                                   (26, 18944, 1664), (67, 6656, 1408)])
def test_w4a4_output_torchao_shapes(ops, M, N, K, use_bias, dtype):
    """Shapes + formula of third-party/ao/test/test_rowwise_scaled_linear_cutlass.py:14-21,68-76;
    same fp32 epilogue order as the kernel => outputs must be bit-identical to the oracle."""
    rng = np.random.default_rng(N + K + M)
    a = O.pack_i4(rng.integers(-8, 8, (M, K), dtype=np.int8))
    w = O.pack_i4(rng.integers(-8, 8, (N, K), dtype=np.int8))
    tdt = torch.float16 if dtype == "f16" else torch.bfloat16
    sA = ((0.5 + torch.rand(M)) * 2.0 ** -3).to(tdt)
    sW = ((0.5 + torch.rand(N)) * 2.0 ** -7).to(tdt)
    bias = torch.rand(N).to(tdt) if use_bias else None
    out = torch.empty(M, N, dtype=tdt, device="cuda")
    ops.rowwise_scaled_linear_cutlass_s4s4_unified(cu(a, torch.int8), sA.cuda(), cu(w, torch.int8), sW.cuda(),
                                                   None if bias is None else bias.cuda(), out)
    acc = torch.from_numpy(O.gemm_w4a4_acc(a, w)).float()
    ref = (acc * sA.float()[:, None]) * sW.float()[None, :]
    if bias is not None:
        ref = ref + bias.float()[None, :]
    ref = ref.to(tdt)
    assert torch.equal(out.cpu(), ref)


def test_w4a4_golden_reference_formula(ops, golden_dir):
    g = np.load(os.path.join(golden_dir, "w4a4_output_ref.npz"))
    for i in range(6):
        xp, sx = g[f"s{i}_seed_xp"], g[f"s{i}_sx"]
        wp, sw, bias = g[f"s{i}_wp_cols"], g[f"s{i}_sw_cols"], g[f"s{i}_bias_cols"]
        out = torch.empty(xp.shape[0], wp.shape[0], dtype=torch.float16, device="cuda")
        ops.rowwise_scaled_linear_cutlass_s4s4_unified(cu(xp), cu(sx), cu(wp), cu(sw), cu(bias), out)
        assert np.array_equal(out.cpu().numpy(), g[f"s{i}_refb_cols"]), f"golden shape {i}"


# ------------------------------------------------------------------------------ W4A16 GEMM
def rel_err(got, ref):
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-12))


@pytest.mark.parametrize("dtype", ["f16", "bf16"])
@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16])
def test_w4a16_q_proj_config1(ops, M, dtype):
    """BASELINE.json configs[0]: K=N=4096, M=1..16."""
    N = K = 4096
    wp, sW = O.make_weight(N, K, 1000)
    x = O.make_act(M, K, 2000 + M)
    tdt = torch.float16 if dtype == "f16" else torch.bfloat16
    xd = cu(x).to(tdt); sWd = cu(sW).to(tdt)
    out = torch.empty(M, N, dtype=tdt, device="cuda")
    ops.gemm_w4a16(xd, cu(wp, torch.int8), sWd, None, out)
    ref = xd.double().cpu() @ (torch.from_numpy(O.unpack_i4(wp)).double() * sWd.double().cpu()[:, None]).T
    err = rel_err(out.double().cpu().numpy(), ref.numpy())
    assert err <= REL_TOL, f"rel err {err}"
    # fp32 accumulation: in practice two orders tighter than the bar
    assert err <= (2e-3 if dtype == "f16" else 1e-2)


@pytest.mark.parametrize("M,N,K", [(16, 6144, 4096), (16, 28672, 4096), (16, 4096, 14336), (64, 4096, 4096),
                                   (100, 512, 1024), (5, 144, 384), (33, 1024, 1792),
                                   # CTA pairs (N % 256 == 0, M >= 128): 128-token tiles, then ONE 129..256-token tile (UMMA N up to 256)
                                   (128, 512, 1024), (129, 512, 1024), (200, 1024, 512), (256, 4096, 4096), (256, 6144, 4096),
                                   (300, 512, 1024), (256, 640, 1024)])
def test_w4a16_shapes_bias(ops, M, N, K):
    wp, sW = O.make_weight(N, K, 1001)
    x = O.make_act(M, K, 2000 + M, heavy_tail=True)
    bias = (np.random.default_rng(3).standard_normal(N)).astype(np.float16)
    out = torch.empty(M, N, dtype=torch.float16, device="cuda")
    ops.gemm_w4a16(cu(x), cu(wp, torch.int8), cu(sW), cu(bias), out)
    ref = O.gemm_w4a16(x, wp, sW, bias).astype(np.float32)
    assert rel_err(out.float().cpu().numpy(), ref) <= REL_TOL
    # deterministic split-K: a second run is bit-identical
    out2 = torch.empty_like(out)
    ops.gemm_w4a16(cu(x), cu(wp, torch.int8), cu(sW), cu(bias), out2)
    assert torch.equal(out, out2)


@pytest.mark.parametrize("np256", [0, 2])
@pytest.mark.parametrize("M", [16, 128, 200, 256, 384])
def test_w4a16_exact_on_integers(ops, M, np256, monkeypatch):
    """With integer activations and unit scales every product and sum is exact in fp32:
    the kernel must reproduce the integer GEMM exactly (catches any fragment/layout slip, in every tile form:
    cta_group::1, CTA pairs with 128-token tiles, CTA pairs with one 256-token tile, two tiles of 256 / 128).
    QSPEC_GEMM_NP256 is read at every call: 0 = 128-token tiles, 2 = 256-token tiles on every pair-mode shape (the default
    uses them only where every pair owns whole tiles, e.g. gate_up: covered by the full-size tests)."""
    monkeypatch.setenv("QSPEC_GEMM_NP256", str(np256))
    N, K = 512, 2048
    rng = np.random.default_rng(9)
    wq = rng.integers(-8, 8, (N, K), dtype=np.int8)
    xi = rng.integers(-4, 5, (M, K)).astype(np.float16)
    out = torch.empty(M, N, dtype=torch.float16, device="cuda")
    sW = np.ones(N, dtype=np.float16) * np.float16(2.0 ** -6)
    ops.gemm_w4a16(cu(xi), cu(O.pack_i4(wq), torch.int8), cu(sW), None, out)
    ref = (xi.astype(np.int64) @ wq.astype(np.int64).T).astype(np.float64) * 2.0 ** -6
    assert np.array_equal(out.cpu().numpy().astype(np.float64), ref.astype(np.float16).astype(np.float64))


# ------------------------------------------------------------------------------ errors
def test_error_behaviour(ops):
    x = torch.zeros(2, 100, dtype=torch.float16, device="cuda")     # K % 16 != 0
    q = torch.empty(2, 50, dtype=torch.int8, device="cuda")
    s = torch.empty(2, dtype=torch.float16, device="cuda")
    with pytest.raises(ValueError):
        ops.fuse_sym_quant(x, s, q, 1.0)
    with pytest.raises(ValueError):
        ops.fuse_sym_quant(x.cpu(), s, q, 1.0)                      # no CPU path
    with pytest.raises(ValueError):
        ops.faster_fast_hadamard_transform(torch.zeros(2, 24, dtype=torch.float16, device="cuda"), 1.0,
                                           torch.zeros(2, 24, dtype=torch.float16, device="cuda"))
    # empty batch is a no-op
    ops.fuse_sym_quant(x[:0, :96].contiguous(), s[:0], q[:0, :48].contiguous(), 1.0)
