"""CPU tests: pin the oracle against the reference-generated golden vectors (tests/golden/,
made by tests/golden/make_golden.py from /root/reference's own python helpers)."""
import os

import numpy as np
import pytest

from oracle import qspec_oracle as O


def test_pack_unpack_matches_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "pack_i4.npz"))
    assert np.array_equal(O.pack_i4(g["q"]), g["packed"])
    assert np.array_equal(O.unpack_i4(g["packed"]), g["q"])


def test_pack_all_nibbles_roundtrip():
    q = np.arange(-8, 8, dtype=np.int8)
    qq = np.stack(np.meshgrid(q, q), -1).reshape(-1, 2)
    assert np.array_equal(O.unpack_i4(O.pack_i4(qq)), qq)
    # even index -> low nibble (quantization.py:47)
    assert O.pack_i4(np.array([[1, -1]], dtype=np.int8))[0, 0] == 0xF1


def test_fht_is_sylvester_hadamard():
    from scipy.linalg import hadamard
    rng = np.random.default_rng(0)
    for n in (8, 32, 512):
        x = rng.standard_normal((15, n)).astype(np.float32)
        ref = x.astype(np.float64) @ hadamard(n).astype(np.float64)
        got = O.fht_f32(x)
        np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-4)


def test_hadamard_matches_reference_matmul_hadU(golden_dir):
    g = np.load(os.path.join(golden_dir, "matmul_hadU.npz"))
    # o_proj rotation: FHT_32 / sqrt(32)
    got = O.fht(g["x32"].astype(np.float32), 1.0 / np.sqrt(32.0))
    np.testing.assert_allclose(got, g["y32"], rtol=1e-5, atol=1e-5)
    got = O.fht(g["x512"].astype(np.float32), 1.0 / np.sqrt(512.0))
    np.testing.assert_allclose(got, g["y512"], rtol=1e-5, atol=1e-5)
    # down_proj rotation: the reference's matmul_hadU(n=14336) views [..., 512, 28]-style
    # recursion; the fused CUDA path is had28 (x) H512 on the [28, 512] view.  Both are
    # (had28 kron H512)/sqrt(14336) up to the index convention checked here.
    x = g["x14336"].astype(np.float32)
    got = O.hadamard_k28(x.astype(np.float32), 512)
    had = O.had28_matrix().astype(np.float64)
    from scipy.linalg import hadamard
    full = np.kron(had, hadamard(512).astype(np.float64)) / np.sqrt(14336.0)
    ref = g["x14336"] @ full.T
    np.testing.assert_allclose(got, ref, rtol=1e-4, atol=2e-4)
    # and matmul_hadU itself equals that Kronecker matrix
    np.testing.assert_allclose(g["y14336"], ref, rtol=1e-6, atol=1e-6)  # ref divides by a float32 sqrt


def test_hadamard_heads_equals_transposed_fht():
    rng = np.random.default_rng(1)
    rows, H, D = 3, 32, 128
    x = rng.standard_normal((rows, H * D)).astype(np.float16)
    y = O.hadamard_heads(x, H, D)
    from scipy.linalg import hadamard
    ref = np.einsum("gh,rhd->rgd", hadamard(H).astype(np.float64),
                    x.reshape(rows, H, D).astype(np.float64)) / np.sqrt(H)
    np.testing.assert_allclose(y.astype(np.float64).reshape(rows, H, D), ref, atol=2e-3, rtol=2e-3)


def test_sym_quant_matches_reference_fake_quant(golden_dir):
    g = np.load(os.path.join(golden_dir, "sym_quant_fake.npz"))
    x16 = g["x"].astype(np.float16)
    packed, scale = O.sym_quant_i4(x16)
    q = O.unpack_i4(packed)
    # the reference fake-quant works in fp32 on the un-rounded x: compare on fp16 inputs with
    # its own formula (functional/quantization.py:32-35), allowing the fp16 x/scale rounding
    ref = np.clip(np.rint(x16.astype(np.float32) / scale.astype(np.float32)[:, None]), -8, 7)
    assert np.abs(q.astype(np.int32) - ref.astype(np.int32)).max() <= 1
    assert (q != ref).mean() < 2e-3
    # scale = amax/7 in fp16
    assert np.array_equal(scale, (np.abs(x16).max(1).astype(np.float32) / np.float32(7)).astype(np.float16))
    # amax element always maps to +-7
    assert (np.abs(q).max(axis=1) == 7).all()


def test_sym_quant_zero_row_and_extremes():
    x = np.zeros((2, 32), dtype=np.float16)
    x[1, 3] = 65504.0
    x[1, 4] = -65504.0
    packed, scale = O.sym_quant_i4(x)
    q = O.unpack_i4(packed)
    assert scale[0] == 0 and (q[0] == 0).all()          # 0/0 -> NaN -> 0 (quant.cu:150-152)
    assert q[1, 3] == 7 and q[1, 4] == -7


def test_w4a4_oracle_matches_reference_formula(golden_dir):
    g = np.load(os.path.join(golden_dir, "w4a4_output_ref.npz"))
    for i in range(6):
        xp, sx = g[f"s{i}_seed_xp"], g[f"s{i}_sx"]
        wp, sw = g[f"s{i}_wp_cols"], g[f"s{i}_sw_cols"]
        out = O.gemm_w4a4(xp, sx, wp, sw)
        assert np.array_equal(out, g[f"s{i}_ref_cols"]), f"shape {i}"
        outb = O.gemm_w4a4(xp, sx, wp, sw, bias=g[f"s{i}_bias_cols"])
        assert np.array_equal(outb, g[f"s{i}_refb_cols"]), f"shape {i} (bias)"


def test_w4a4_acc_exact_int():
    rng = np.random.default_rng(3)
    a = rng.integers(-8, 8, (5, 256), dtype=np.int8)
    w = rng.integers(-8, 8, (24, 256), dtype=np.int8)
    acc = O.gemm_w4a4_acc(O.pack_i4(a), O.pack_i4(w))
    assert np.array_equal(acc, a.astype(np.int64) @ w.astype(np.int64).T)
    # extreme: all -8 x all -8 over K
    a[:] = -8
    w[:] = -8
    assert (O.gemm_w4a4_acc(O.pack_i4(a), O.pack_i4(w)) == 64 * 256).all()


def test_w4a16_oracle_matches_torch_statement():
    import torch
    wp, sW = O.make_weight(64, 256, 7)
    x = O.make_act(5, 256, 8)
    ref = (torch.from_numpy(x).float() @
           (torch.from_numpy(O.unpack_i4(wp)).float() * torch.from_numpy(sW).float()[:, None]).T)
    got = O.gemm_w4a16(x, wp, sW).astype(np.float32)
    np.testing.assert_allclose(got, ref.numpy(), rtol=2e-3, atol=2e-3)


def test_norm_quant_properties():
    x = O.make_act(6, 4096, 11, heavy_tail=True)
    packed, scale, isum = O.norm_quant_i4(x, 1e-5)
    q = O.unpack_i4(packed)
    y = O.norm_f16(x, 1e-5).astype(np.float32)
    # textbook mean-subtracting layernorm (layernorm_kernels.cu:40-48), no gamma
    xf = x.astype(np.float64)
    ref = (xf - xf.mean(1, keepdims=True)) / np.sqrt(xf.var(1, keepdims=True) + 1e-5)
    np.testing.assert_allclose(y, ref, rtol=2e-3, atol=2e-3)
    assert np.array_equal(scale, (np.abs(y).max(1).astype(np.float32) / np.float32(7)).astype(np.float16))
    deq = q.astype(np.float32) * scale.astype(np.float32)[:, None]
    assert np.abs(deq - y).max() <= 0.51 * scale.astype(np.float32).max() + 1e-3
    assert (np.abs(q).max(axis=1) == 7).all()


def test_tree_sum_is_a_sum():
    rng = np.random.default_rng(5)
    for K in (16, 128, 4096, 8192, 14336):
        v = rng.standard_normal((3, K)).astype(np.float32)
        np.testing.assert_allclose(O.tree_sum_f32(v), v.astype(np.float64).sum(1), rtol=1e-5, atol=1e-3)


def test_bf16_helpers_roundtrip():
    import torch
    v = np.random.default_rng(2).standard_normal(1000).astype(np.float32)
    ref = torch.from_numpy(v).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(O.f32_to_bf16_bits(v), ref)
    assert np.array_equal(O.bf16_bits_to_f32(ref), torch.from_numpy(v).to(torch.bfloat16).float().numpy())


# ----------------------------------------------------------------------------------------------
# oracle vs outputs of the reference's OWN CUDA kernels (K1-K4 + the had28 mix), produced on a B200 by
# tests/golden/make_golden_gpu.py from oracle/_ref (the reference sources compiled where they lie)
# ----------------------------------------------------------------------------------------------
def _ulp16(a, b):
    def key(v):
        u = v.view(np.uint16).astype(np.int32)
        return np.where(u & 0x8000, -(u & 0x7FFF), u & 0x7FFF)
    return np.abs(key(np.ascontiguousarray(a)) - key(np.ascontiguousarray(b)))


@pytest.fixture(scope="module")
def ref_kernel_vectors(golden_dir):
    return np.load(os.path.join(golden_dir, "ref_kernels_b200.npz"))


def test_norm_quant_i4_matches_reference_kernel_output(ref_kernel_vectors):
    """K1 rms_norm_general_fuse_sum_i4 (layernorm_kernels.cu:569-720, --use_fast_math build): packed int4 bytes and f16 scales are
    identical, including the all-zero row (amax floor 1e-6) and the outlier row."""
    g = ref_kernel_vectors
    q, s, isum = O.norm_quant_i4(g["k12_x"], 1e-5)
    assert np.array_equal(q, g["k1_q"])
    assert np.array_equal(s.view(np.uint16), g["k1_scale"].view(np.uint16))
    # input_sum: sum of mean-subtracted values, i.e. rounding noise around 0; the reference accumulates it per thread in f16
    # (unused downstream) -- only the magnitude is comparable.
    assert np.abs(isum.astype(np.float32) - g["k1_input_sum"].astype(np.float32)).max() < 0.1


def test_norm_f16_matches_reference_kernel_output(ref_kernel_vectors):
    """K2 rms_norm_general_fuse_sum_fp16: <= 1 f16 ulp (the reference reduces in a block-size-dependent order with fast-math rsqrt);
    on this fixture 2 of 49152 values differ."""
    g = ref_kernel_vectors
    u = _ulp16(O.norm_f16(g["k12_x"], 1e-5), g["k2_y"])
    assert u.max() <= 1
    assert (u != 0).sum() <= 1e-3 * u.size


@pytest.mark.parametrize("tag", ["a", "b"])
def test_sym_quant_matches_reference_kernel_output(ref_kernel_vectors, tag):
    """K3 fuse_sym_quant (quant.cu:102-186) at K = 4096 and 14336, incl. the amax = 0 row (scale 0 -> NaN -> 0) and +-65504."""
    g = ref_kernel_vectors
    q, s = O.sym_quant_i4(g[f"k3{tag}_x"], 1.0)
    assert np.array_equal(q, g[f"k3{tag}_q"])
    assert np.array_equal(s.view(np.uint16), g[f"k3{tag}_scale"].view(np.uint16))


def test_fht_matches_reference_kernel_output(ref_kernel_vectors):
    """K4 faster_fast_hadamard_transform at the two QSpec sizes ([T*128, 32] * 1/sqrt(32), [T*28, 512] * 1/sqrt(14336))."""
    g = ref_kernel_vectors
    ya = O.fht(g["k4a_x"], float(g["k4a_scale"]))
    assert np.array_equal(ya.view(np.uint16), g["k4a_y"].view(np.uint16))
    yb = O.fht(g["k4b_x"].reshape(-1, 512), float(g["k4b_scale"]))
    assert np.array_equal(yb.view(np.uint16), np.ascontiguousarray(g["k4b_fht"].reshape(-1, 512)).view(np.uint16))


def test_hadamard_k28_matches_reference_gpu_path(ref_kernel_vectors):
    """down_proj rotation as the reference runs it on the GPU: FHT kernel -> f16 -> torch fp16 matmul with had28 (hadamard.py:109-113)."""
    g = ref_kernel_vectors
    y = O.hadamard_k28(g["k4b_x"], 512, float(g["k4b_scale"]))
    assert np.array_equal(y.view(np.uint16), np.ascontiguousarray(g["k5_y"]).view(np.uint16))


def test_w4a8_quantisers_match_reference_kernel_output(golden_dir):
    """W4A8 mode: norm + int8 quantiser (layernorm_kernels.cu:350-470) and fuse_sym_quant_i8 (quant.cu:188-245) as the reference's own
    kernels computed them on a B200 (tests/golden/make_golden_gpu.py): identical values and scales."""
    g = np.load(os.path.join(golden_dir, "ref_kernels_b200_w4a8.npz"))
    q, s, _ = O.norm_quant_i8(g["k1i8_x"], 1e-5)
    assert np.array_equal(q, g["k1i8_q"]) and np.array_equal(s.view(np.uint16), g["k1i8_scale"].view(np.uint16))
    q, s = O.sym_quant_i8(g["k3i8_x"], 1.0)
    assert np.array_equal(q, g["k3i8_q"]) and np.array_equal(s.view(np.uint16), g["k3i8_scale"].view(np.uint16))


def test_w4a8_gemm_oracle_is_the_integer_product():
    rng = np.random.default_rng(8)
    a = rng.integers(-128, 128, (5, 384), dtype=np.int8)
    wq = rng.integers(-8, 8, (24, 384), dtype=np.int8)
    acc = O.gemm_w4a8_acc(a, O.pack_i4(wq))
    assert np.array_equal(acc, a.astype(np.int64) @ wq.astype(np.int64).T)
    sA, sW = np.full(5, 0.5, np.float16), np.full(24, 0.25, np.float16)
    out = O.gemm_w4a8(a, sA, O.pack_i4(wq), sW)
    assert np.array_equal(out, (acc.astype(np.float32) * np.float32(0.125)).astype(np.float16))


def test_qwen_rotations_match_reference_matmul_hadU(golden_dir):
    """had108 (x) H_128 (intermediate 13824) and had40 (40 heads) against the reference's own float64 matmul_hadU
    (third-party/QuaRot/quarot/functional/hadamard.py:59-81; vectors from tests/golden/make_golden.py). Neither matrix is
    symmetric, so the orientation (hadK @ x, row o of hadK -> output block o) is pinned here."""
    g = np.load(os.path.join(golden_dir, "matmul_hadU_qwen.npz"))
    h108, h40 = O.hadK_matrix(108), O.hadK_matrix(40)
    assert not np.array_equal(h108, h108.T) and not np.array_equal(h40, h40.T)
    assert np.array_equal(h108.astype(np.int32) @ h108.T.astype(np.int32), 108 * np.eye(108, dtype=np.int32))
    x, y = g["x13824"], g["y13824"]
    got = O.hadamard_kK(x.astype(np.float16), 128, 108).astype(np.float64)
    assert np.abs(got - y).max() <= 2e-3 * np.abs(y).max()                      # f16 input + two f16 roundings
    wrong = O.hadamard_kK(x.astype(np.float16), 128, 108, had=h108.T).astype(np.float64)
    assert np.abs(wrong - y).max() > 0.1 * np.abs(y).max()
    x40, y40 = g["x40"], g["y40"]
    got = O.hadamard_heads_hadK(x40.astype(np.float16), 40, 1).astype(np.float64)
    assert np.abs(got - y40).max() <= 2e-3 * np.abs(y40).max()


def test_qwen_get_hadK_precedence_matches_reference():
    """nn.get_hadK follows the reference's order of tests (hadamard.py:6-56): 13824 -> K = 108 (not 28-divisible anyway),
    14336 -> 28, 40 -> 40, 32 -> 1; unsupported sizes raise."""
    from qspec_b200.nn import get_hadK
    assert get_hadK(13824)[1] == 108 and get_hadK(14336)[1] == 28 and get_hadK(40)[1] == 40 and get_hadK(32) == (None, 1)
    assert np.array_equal(get_hadK(13824)[0].numpy().astype(np.int8), O.hadK_matrix(108))
    assert np.array_equal(get_hadK(40)[0].numpy().astype(np.int8), O.hadK_matrix(40))
    with pytest.raises(NotImplementedError):
        get_hadK(172 * 64)
