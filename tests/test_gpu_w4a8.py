"""W4A8 mode (the reference's kwargs["w4a8"] path): int8 activation quantisers and the int8 x int4 GEMM through
the C ABI against the CPU oracle. Bars as for W4A4: quantised activations, int32 accumulators and f16/bf16 outputs bit-exact."""
import numpy as np
import pytest
import torch

from oracle import qspec_oracle as O

pytestmark = pytest.mark.gpu
EPS = 1e-5


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from qspec_b200 import ops as _ops
    return _ops


def cu(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    return (t.view(dtype) if dtype is not None else t).cuda()


def special_rows(K, seed):
    x = O.make_act(6, K, seed, heavy_tail=True)
    x[2] = 0
    x[3, 5], x[3, 9] = 65504.0, -65504.0
    x[4] = np.float16(0.37)
    return x


@pytest.mark.parametrize("K", [32, 4096, 14336])
def test_sym_quant_i8_bit_exact(ops, K):
    x = special_rows(K, 5200 + K)
    rq, rs = O.sym_quant_i8(x, 1.0)
    q = torch.empty(x.shape[0], K, dtype=torch.int8, device="cuda")
    s = torch.empty(x.shape[0], dtype=torch.float16, device="cuda")
    ops.fuse_sym_quant_i8(cu(x), s, q, 1.0)
    assert np.array_equal(q.cpu().numpy(), rq) and np.array_equal(s.cpu().numpy().view(np.uint16), rs.view(np.uint16))


@pytest.mark.parametrize("K", [128, 4096, 8192])
def test_norm_quant_i8_bit_exact(ops, K):
    x = special_rows(K, 5300 + K)
    x[3, 5], x[3, 9] = 300.0, -300.0
    rq, rs, rsum = O.norm_quant_i8(x, EPS)
    q = torch.empty(x.shape[0], K, dtype=torch.int8, device="cuda")
    s = torch.empty(x.shape[0], dtype=torch.float16, device="cuda")
    isum = torch.empty_like(s)
    ops.rms_norm_general_fuse_sum_i8(q, cu(x), isum, s, EPS, True)
    assert np.array_equal(q.cpu().numpy(), rq)
    assert np.array_equal(s.cpu().numpy(), rs) and np.array_equal(isum.cpu().numpy(), rsum)


W4A8_SHAPES = [(1, 4096, 4096), (4, 4096, 4096), (16, 6144, 4096), (7, 4096, 14336), (64, 4096, 4096), (200, 512, 1024),
               (300, 640, 256), (16, 136, 128), (5, 144, 384), (4, 28672, 4096)]


@pytest.mark.parametrize("M,N,K", W4A8_SHAPES)
def test_w4a8_acc_int32_bit_exact(ops, M, N, K):
    rng = np.random.default_rng(M * 7 + N + K)
    a = rng.integers(-128, 128, (M, K), dtype=np.int8)
    w = O.pack_i4(rng.integers(-8, 8, (N, K), dtype=np.int8))
    acc = ops.gemm_w4a8_acc_i32(cu(a), cu(w, torch.int8))
    torch.cuda.synchronize()
    assert np.array_equal(acc.cpu().numpy(), O.gemm_w4a8_acc(a, w))
    acc2 = ops.gemm_w4a8_acc_i32(cu(a), cu(w, torch.int8))                 # self-cleaning split-K workspace
    assert torch.equal(acc, acc2)


def test_w4a8_acc_extremes_and_k_order(ops):
    M, N, K = 16, 256, 14336
    a = np.full((M, K), -128, dtype=np.int8)
    acc = ops.gemm_w4a8_acc_i32(cu(a), cu(O.pack_i4(np.full((N, K), -8, dtype=np.int8)), torch.int8))
    assert (acc == 1024 * K).all()
    acc = ops.gemm_w4a8_acc_i32(cu(np.full((M, K), 127, dtype=np.int8)), cu(O.pack_i4(np.full((N, K), -8, dtype=np.int8)), torch.int8))
    assert (acc == -1016 * K).all()
    # one-hot activations pick single weight columns: any permutation of k between the operands shows up here
    K = 512
    a = np.zeros((M, K), dtype=np.int8)
    ks = np.random.default_rng(3).choice(K, size=M, replace=False)
    a[np.arange(M), ks] = 1
    wq = np.random.default_rng(4).integers(-8, 8, (N, K), dtype=np.int8)
    acc = ops.gemm_w4a8_acc_i32(cu(a), cu(O.pack_i4(wq), torch.int8))
    assert np.array_equal(acc.cpu().numpy(), wq[:, ks].T.astype(np.int32))


@pytest.mark.parametrize("dtype", ["f16", "bf16"])
@pytest.mark.parametrize("M,N,K", [(4, 4096, 4096), (16, 4096, 14336), (67, 6656, 1408), (13, 512, 384)])
def test_w4a8_output_bit_exact(ops, M, N, K, dtype):
    rng = np.random.default_rng(M + N + K)
    a = rng.integers(-128, 128, (M, K), dtype=np.int8)
    wp, sW = O.make_weight(N, K, 1234)
    sA = ((0.5 + rng.random(M, dtype=np.float32)) * np.float32(2.0 ** -6)).astype(np.float16)
    bias = rng.standard_normal(N).astype(np.float16)
    if dtype == "f16":
        out = torch.empty(M, N, dtype=torch.float16, device="cuda")
        ops.rowwise_scaled_linear_cutlass_s8s4(cu(a), cu(sA), cu(wp, torch.int8), cu(sW), cu(bias), out)
        assert np.array_equal(out.cpu().numpy(), O.gemm_w4a8(a, sA, wp, sW, bias))
    else:
        tb = lambda v: torch.from_numpy(v.astype(np.float32)).to(torch.bfloat16)
        sAb, sWb, bb = tb(sA), tb(sW), tb(bias)
        out = ops.rowwise_scaled_linear_cutlass_s8s4(cu(a), sAb.cuda(), cu(wp, torch.int8), sWb.cuda(), bb.cuda())
        ref = O.gemm_w4a8(a, sAb.float().numpy(), wp, sWb.float().numpy(), bb.float().numpy(), out_dtype="bf16")
        got = out.view(torch.int16).cpu().numpy().view(np.uint16)
        assert np.array_equal(got, np.asarray(ref).view(np.uint16))


def test_module_chain_w4a8(ops):
    from qspec_b200 import nn as Q
    T, H = 5, 4096
    x = O.make_act(T, H, 61, heavy_tail=True)
    wp, sW = O.make_weight(H, H, 9)
    lin = Q.Linear4bit(H, H)
    lin.weight.copy_(torch.from_numpy(wp)); lin.weight_scales.copy_(torch.from_numpy(sW).view(-1, 1))
    lin = lin.cuda()
    p = Q.RMSNorm(H, EPS, fuse=True)(cu(x), w4a8=True)
    rq, rs, _ = O.norm_quant_i8(x, EPS)
    assert p.quantized_x.dtype == torch.int8 and tuple(p.quantized_x.shape) == (T, H)
    assert np.array_equal(p.quantized_x.cpu().numpy(), rq) and np.array_equal(p.scales_x.cpu().numpy(), rs)
    assert np.array_equal(lin(p, w4a8=True).cpu().numpy(), O.gemm_w4a8(rq, rs, wp, sW))
    pq = Q.Quantizer()(cu(x), w4a8=True)
    rq2, rs2 = O.sym_quant_i8(x, 1.0)
    assert np.array_equal(pq.quantized_x.cpu().numpy(), rq2)
    assert np.array_equal(lin(pq, w4a8=True).cpu().numpy(), O.gemm_w4a8(rq2, rs2, wp, sW))
