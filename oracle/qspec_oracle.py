"""CPU oracle for the QSpec dual-mode int4 linear path (numpy).

TEST INFRASTRUCTURE ONLY.  Nothing under ``qspec_b200/`` may import this module; only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs use it, and only as the checker / reported baseline.

It restates, function by function, the arithmetic of the reference's hot path
(hku-netexplo-lab/QSpec, paths relative to that tree).  Pinning status:

* pack/unpack, Hadamard matrices, ``matmul_hadU`` and the W4A4 ``output_ref`` formula are
  checked against the reference's own pure-python helpers (imported from /root/reference by
  ``tests/golden/make_golden.py``; vectors committed under ``tests/golden/``).
* norm+quant (K1/K2), fuse_sym_quant (K3) and the FHT entry point (K4) have NO test or fixture in the
  reference and are CUDA-only there, so they are pinned against the reference's OWN kernels:
  ``oracle/build_ref.py`` compiles the reference sources where they lie into ``oracle/_ref/`` and
  ``tests/golden/make_golden_gpu.py`` ran them on a B200 (vectors committed as
  ``tests/golden/ref_kernels_b200.npz``; live comparison in ``tests/test_gpu_ref_kernels.py``).
  Result: K1 packed bytes + scales, K3, K4 and the had28 mix are IDENTICAL; K2 is within 1 f16 ulp
  (the reference is a ``--use_fast_math`` build with a block-size-dependent reduction order).  The oracle
  follows the kernel source line by line with IEEE-rounded fp32 ops and a fixed reduction tree,
  which the new CUDA kernels mirror bit-for-bit.
* W4A16 lives in BitBLAS (not vendored, unpinned version) and cannot be run; the oracle is the torch
  statement of intent at quarot_nn/linear.py:111-119, checked on a B200 against the reference's own
  Triton kernel for the op (quarot_nn/qspec_gemm.py, AOT-compiled by oracle/build_ref.py): 4-7e-4
  relative on six shapes, inside north_star's 1e-2 (tests/test_gpu_ref_kernels.py).
"""
from __future__ import annotations

import numpy as np

F32 = np.float32
F16 = np.float16

# ----------------------------------------------------------------------------------------------
# a1: nibble format  (third-party/QuaRot/quarot/functional/quantization.py:41-74)
# ----------------------------------------------------------------------------------------------

def pack_i4(q: np.ndarray) -> np.ndarray:
    """[R, C] signed ints in [-8, 7] -> [R, C/2] uint8; even index -> low nibble
    (quantization.py:41-49: ``q_i8[:, 0::2] | (q_i8[:, 1::2] << 4)``)."""
    q = np.asarray(q)
    assert q.shape[-1] % 2 == 0
    assert q.min(initial=0) >= -8 and q.max(initial=0) <= 7
    u = (q.astype(np.int16) & 0xF).astype(np.uint8)
    return (u[..., 0::2] | (u[..., 1::2] << 4)).astype(np.uint8)


def unpack_i4(p: np.ndarray) -> np.ndarray:
    """[R, C/2] uint8/int8 -> [R, C] int8 in [-8, 7] (quantization.py:52-74)."""
    p = np.asarray(p).view(np.uint8)
    lo = (p & 0xF).astype(np.int8)
    hi = (p >> 4).astype(np.int8)
    lo = np.where(lo >= 8, lo - 16, lo).astype(np.int8)
    hi = np.where(hi >= 8, hi - 16, hi).astype(np.int8)
    out = np.empty(p.shape[:-1] + (p.shape[-1] * 2,), dtype=np.int8)
    out[..., 0::2] = lo
    out[..., 1::2] = hi
    return out


# ----------------------------------------------------------------------------------------------
# fixed fp32 reduction tree shared with qspec_b200/csrc/act_kernels.cu (block_sum_256)
# ----------------------------------------------------------------------------------------------
_NT = 256  # threads per row in the norm kernels


def _xor_tree(v: np.ndarray) -> np.ndarray:
    """v [..., 32] fp32 -> [...]: butterfly sum with lane offsets 16, 8, 4, 2, 1."""
    lanes = np.arange(32)
    for off in (16, 8, 4, 2, 1):
        v = (v + v[..., lanes ^ off]).astype(F32)
    return v[..., 0]


def tree_sum_f32(vals: np.ndarray) -> np.ndarray:
    """Sum ``vals`` [rows, K] (fp32, K % 8 == 0) exactly as the CUDA norm kernel does:
    each 8-vector is summed left to right; thread t (of 256) adds its vectors t, t+256, ...
    in order starting from 0.0f; warps reduce with the xor butterfly; the 8 warp sums
    (padded with 0 to 32 lanes) go through the same butterfly."""
    vals = np.asarray(vals, dtype=F32)
    rows, K = vals.shape
    assert K % 8 == 0
    nv = K // 8
    v8 = vals.reshape(rows, nv, 8)
    s8 = v8[..., 0].copy()
    for i in range(1, 8):
        s8 = (s8 + v8[..., i]).astype(F32)
    slots = -(-nv // _NT)
    pad = np.zeros((rows, slots * _NT), dtype=F32)
    pad[:, :nv] = s8
    valid = np.zeros(slots * _NT, dtype=bool)
    valid[:nv] = True
    pad = pad.reshape(rows, slots, _NT)
    valid = valid.reshape(slots, _NT)
    part = np.zeros((rows, _NT), dtype=F32)
    for s in range(slots):
        # threads without a vector in this slot skip the add (keeps -0.0 handling identical)
        part = np.where(valid[s][None, :], (part + pad[:, s, :]).astype(F32), part)
    wsum = _xor_tree(part.reshape(rows, _NT // 32, 32))  # [rows, 8]
    w32 = np.zeros((rows, 32), dtype=F32)
    w32[:, : _NT // 32] = wsum
    return _xor_tree(w32)


# ----------------------------------------------------------------------------------------------
# a5/a6: mean-subtracting "RMSNorm" (+ int4 quant)
#   third-party/kernels/csrc/layernorm_kernels.cu:569-720 (kernel), :890-957 (hosts)
# ----------------------------------------------------------------------------------------------

def _norm_core(x: np.ndarray, eps: float):
    x = np.asarray(x)
    assert x.dtype == F16 or x.dtype == F32
    xf = x.astype(F32)
    rows, K = xf.shape
    mean = (tree_sum_f32(xf) / F32(K)).astype(F32)                       # :587-627
    d = (xf - mean[:, None]).astype(F32)                                 # :633-637
    var = tree_sum_f32((d * d).astype(F32))
    v = ((var / F32(K)).astype(F32) + F32(eps)).astype(F32)              # :641
    r = (F32(1.0) / np.sqrt(v, dtype=F32)).astype(F32)                   # rsqrtf -> IEEE 1/sqrt
    y = (d * r[:, None]).astype(F32)                                     # :40-48
    return y


def norm_f16(x: np.ndarray, eps: float = 1e-5) -> np.ndarray:
    """rms_norm_general_fuse_sum_fp16: out = f16((x-mean)*rsqrt(var+eps)) (:679-682)."""
    return _norm_core(x, eps).astype(F16)


def norm_quant_i4(x: np.ndarray, eps: float = 1e-5):
    """rms_norm_general_fuse_sum_i4 (per-token branch, :655-719).
    Returns (packed uint8 [rows,K/2], scale f16 [rows], input_sum f16 [rows])."""
    y = _norm_core(x, eps)
    yh = y.astype(F16)                                                   # :661 cuda_cast<T>
    amax_h = np.maximum(np.abs(yh).max(axis=1), F16(1e-6)).astype(F16)   # :655,665
    amax = amax_h.astype(F32)                                            # :685
    s = (F32(7.0) / amax).astype(F32)                                    # :687
    t = (y * s[:, None]).astype(F32)                                     # :698 (un-rounded fp32 y)
    t = np.clip(t, F32(-8.0), F32(7.0))                                  # clamp_int4
    q = np.rint(t).astype(np.int8)                                       # cvt.rni.sat.s8.f32
    scale = (amax / F32(7.0)).astype(F32).astype(F16)                    # :716
    # Deviation (documented in DESIGN.md): the reference accumulates `sum` per thread in fp16;
    # input_sum is unused downstream, so the oracle defines it as the fp32 tree sum of yh.
    isum = tree_sum_f32(yh.astype(F32)).astype(F16)                      # :717
    return pack_i4(q), scale, isum


# ----------------------------------------------------------------------------------------------
# a9: per-token symmetric quant  (third-party/QuaRot/quarot/kernels/quant.cu:102-166)
# ----------------------------------------------------------------------------------------------

def sym_quant_i4(x: np.ndarray, clip_ratio: float = 1.0):
    """Returns (packed uint8 [rows,K/2], scale f16 [rows])."""
    x = np.asarray(x)
    assert x.dtype == F16
    amax = np.abs(x).max(axis=1).astype(F16)                             # :112-125
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        s0 = (amax.astype(F32) / F32(7.0)).astype(F32).astype(F16)       # __hdiv(max_abs, 7)
        scale = (s0.astype(F32) * F16(clip_ratio).astype(F32)).astype(F32).astype(F16)  # :128
        qh = (x.astype(F32) / scale.astype(F32)[:, None]).astype(F32).astype(F16)       # :150
        qf = qh.astype(F32)
        qf = np.where(np.isnan(qf), F32(0.0), qf)                        # __half2int_rn(NaN) == 0
        q = np.clip(np.rint(qf), -8, 7).astype(np.int8)                  # :152
    return pack_i4(q), scale


# ----------------------------------------------------------------------------------------------
# W4A8 mode: int8 activation quantisers and the int8 x int4 GEMM
#   third-party/kernels/csrc/layernorm_kernels.cu:350-470 (generalLayerNorm_fuse_sum_i8)
#   third-party/QuaRot/quarot/kernels/quant.cu:188-223 (rowAbsMaxQuantizeKernel_i8)
#   quarot_nn/linear.py:86-99 (rowwise_scaled_linear_cutlass_s8s4; same epilogue as a3)
# ----------------------------------------------------------------------------------------------

def norm_quant_i8(x: np.ndarray, eps: float = 1e-5):
    """Returns (int8 [rows,K], scale f16 [rows], input_sum f16 [rows]): as norm_quant_i4 with 127 in place of 7 and
    a saturating round-to-nearest-even to int8 (cvt.rni.sat.s8.f32, utils.cuh:225-235)."""
    y = _norm_core(x, eps)
    yh = y.astype(F16)
    amax = np.maximum(np.abs(yh).max(axis=1), F16(1e-6)).astype(F16).astype(F32)      # :434,441
    s = (F32(127.0) / amax).astype(F32)                                               # :449
    t = np.clip((y * s[:, None]).astype(F32), F32(-128.0), F32(127.0))
    q = np.rint(t).astype(np.int8)                                                    # :461-462
    scale = (amax / F32(127.0)).astype(F32).astype(F16)                               # :466
    isum = tree_sum_f32(yh.astype(F32)).astype(F16)
    return q, scale, isum


def sym_quant_i8(x: np.ndarray, clip_ratio: float = 1.0):
    """Returns (int8 [rows,K], scale f16 [rows]); all-f16 arithmetic like sym_quant_i4, range [-128, 127]."""
    x = np.asarray(x)
    assert x.dtype == F16
    amax = np.abs(x).max(axis=1).astype(F16)
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        s0 = (amax.astype(F32) / F32(127.0)).astype(F32).astype(F16)
        scale = (s0.astype(F32) * F16(clip_ratio).astype(F32)).astype(F32).astype(F16)
        qh = (x.astype(F32) / scale.astype(F32)[:, None]).astype(F32).astype(F16)
        qf = qh.astype(F32)
        qf = np.where(np.isnan(qf), F32(0.0), qf)
        q = np.clip(np.rint(qf), -128, 127).astype(np.int8)
    return q, scale


def gemm_w4a8_acc(a_i8: np.ndarray, w_packed: np.ndarray) -> np.ndarray:
    """Exact int32 accumulators [M,N] of int8 activations x int4 weights (|acc| <= 1024*K: float64 BLAS is exact)."""
    a = np.asarray(a_i8).astype(np.float64)
    w = unpack_i4(w_packed).astype(np.float64)
    return np.rint(a @ w.T).astype(np.int32)


def gemm_w4a8(a_i8, sA, w_packed, sW, bias=None, out_dtype=F16) -> np.ndarray:
    acc = gemm_w4a8_acc(a_i8, w_packed).astype(F32)
    t = (acc * np.asarray(sA).astype(F32)[:, None]).astype(F32)
    t = (t * np.asarray(sW).astype(F32)[None, :]).astype(F32)
    if bias is not None:
        t = (t + np.asarray(bias).astype(F32)[None, :]).astype(F32)
    return _to16(t, out_dtype)


# ----------------------------------------------------------------------------------------------
# a7/a8: online Hadamard
#   fast-hadamard-transform/csrc/fast_hadamard_transform_common.h:86-124,146-161
#   third-party/QuaRot/quarot/functional/hadamard.py:98-126
# ----------------------------------------------------------------------------------------------

def fht_f32(v: np.ndarray) -> np.ndarray:
    """Un-normalised Sylvester FHT along the last axis in fp32, radix-2 butterflies in
    ascending stride (a+b, a-b) -- the order every thread/warp/chunk stage of the reference
    kernel follows (common.h:86-124)."""
    v = np.array(v, dtype=F32, copy=True)
    n = v.shape[-1]
    assert n & (n - 1) == 0
    lead = v.shape[:-1]
    stride = 1
    while stride < n:
        w = v.reshape(lead + (n // (2 * stride), 2, stride))
        a = w[..., 0, :].copy()
        b = w[..., 1, :].copy()
        w[..., 0, :] = (a + b).astype(F32)
        w[..., 1, :] = (a - b).astype(F32)
        v = w.reshape(lead + (n,))
        stride *= 2
    return v


def fht(x: np.ndarray, scale: float) -> np.ndarray:
    """faster_fast_hadamard_transform(x, scale, out): out = T(fht(x) * scale) (common.h:146-161)."""
    x = np.asarray(x)
    return (fht_f32(x.astype(F32)) * F32(scale)).astype(F32).astype(x.dtype)


def hadamard_heads(x: np.ndarray, H: int, D: int, scale: float | None = None) -> np.ndarray:
    """o_proj online Hadamard incl. both transposes (quarot_llama.py:230-234):
    x [rows, H*D] viewed [rows,H,D] -> transpose -> [rows*D, H] -> FHT_H * scale -> back."""
    rows = x.shape[0]
    if scale is None:
        scale = float(F32(1.0) / np.sqrt(F32(H)))
    xt = np.ascontiguousarray(x.reshape(rows, H, D).transpose(0, 2, 1)).reshape(rows * D, H)
    yt = fht(xt, scale)
    return np.ascontiguousarray(yt.reshape(rows, D, H).transpose(0, 2, 1)).reshape(rows, H * D)


def hadK_matrix(K: int) -> np.ndarray:
    """K x K +-1 Hadamard matrix of third-party/QuaRot/quarot/functional/hadamard.py:get_had{K} (28: Llama-3
    intermediate sizes; 108 and 40: the Qwen2.5-14B variant, qwen2_quarot.py:114,216), committed as
    tests/golden/had{K}.npy by tests/golden/make_golden.py and loaded from there."""
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    return np.load(os.path.join(here, "..", "tests", "golden", f"had{K}.npy")).astype(np.int8)


def had28_matrix() -> np.ndarray:
    return hadK_matrix(28)


def hadamard_kK(x: np.ndarray, n: int, K: int = 28, scale: float | None = None,
                had: np.ndarray | None = None) -> np.ndarray:
    """down_proj online Hadamard (hadamard.py:109-113 / :116-126): view [rows,K,n]; FHT_n * scale -> f16;
    then hadK @ . with fp32 accumulation in ascending k, f16 out."""
    if had is None:
        had = hadK_matrix(K)
    assert had.shape == (K, K)
    rows = x.shape[0]
    if scale is None:
        scale = float(F32(1.0) / np.sqrt(F32(K * n)))
    y = fht(x.reshape(rows * K, n), scale).reshape(rows, K, n).astype(F32)
    out = np.zeros((rows, K, n), dtype=F32)
    for k in range(K):
        sgn = had[:, k].astype(F32)[None, :, None]
        out = (out + sgn * y[:, k:k + 1, :]).astype(F32)
    return out.astype(x.dtype).reshape(rows, K * n)


def hadamard_k28(x: np.ndarray, n: int, scale: float | None = None,
                 had: np.ndarray | None = None) -> np.ndarray:
    return hadamard_kK(x, n, 28, scale, had)


def hadamard_heads_hadK(x: np.ndarray, H: int, D: int, scale: float | None = None, had: np.ndarray | None = None) -> np.ndarray:
    """Qwen2Attention.o_had_trans for a head count that is not a power of two (o_K = 40,
    vllm/model_executor/models/qwen2_quarot.py:269-277): `(o_hadK.to(f16) @ x.reshape(-1, H, D)) * o_had_scale` -- the f16
    matmul accumulates in fp32 (ascending k here) and rounds to f16, the multiplication by the python float (fp32 opmath)
    rounds a second time."""
    if had is None:
        had = hadK_matrix(H)
    if scale is None:
        scale = float(F32(1.0) / np.sqrt(F32(H)))
    rows = x.shape[0]
    xv = x.reshape(rows, H, D).astype(F32)
    acc = np.zeros((rows, H, D), dtype=F32)
    for k in range(H):
        acc = (acc + had[:, k].astype(F32)[None, :, None] * xv[:, k:k + 1, :]).astype(F32)
    y16 = acc.astype(F16)
    return (y16.astype(F32) * F32(scale)).astype(F32).astype(F16).reshape(rows, H * D)


def silu_mul_gate_first(gate_up: np.ndarray, inter: int) -> np.ndarray:
    """vLLM SiluAndMul on the merged [gate ; up] projection (qwen2_quarot.py:80-86,122): f16(f16(silu(gate)) * up)."""
    swapped = np.concatenate((gate_up[:, inter:], gate_up[:, :inter]), axis=1)
    return silu_mul(swapped, inter)


def silu_mul(gate_up: np.ndarray, inter: int) -> np.ndarray:
    """quarot_llama.py:279-284: up = gate_up[:, :inter], gate = gate_up[:, inter:];
    f16(f16(silu(gate)) * up).  (exp is libm here, expf on the GPU: compare with tolerance.)"""
    up = gate_up[:, :inter]
    g = gate_up[:, inter:].astype(F32)
    s = (g / (F32(1.0) + np.exp(-g, dtype=F32))).astype(F32).astype(gate_up.dtype)
    return (s.astype(F32) * up.astype(F32)).astype(F32).astype(gate_up.dtype)


# ----------------------------------------------------------------------------------------------
# a3: W4A4 GEMM  (third-party/ao/test/test_rowwise_scaled_linear_cutlass.py:68-76;
#                 epilogue order rowwise_scaled_linear_cutlass_unified.cuh:342-377)
# ----------------------------------------------------------------------------------------------

def gemm_w4a4_acc(a_packed: np.ndarray, w_packed: np.ndarray, w_unpacked: np.ndarray | None = None) -> np.ndarray:
    """Exact int32 accumulators [M,N].  |acc| <= 64*K < 2^24 for K <= 262143, so an fp32
    BLAS matmul of the small integers is exact; asserted below. (w_unpacked: unpack_i4(w_packed) computed by the
    caller once, for loops that reuse a large weight.)"""
    a = unpack_i4(a_packed).astype(F32)
    w = (unpack_i4(w_packed) if w_unpacked is None else w_unpacked).astype(F32)
    assert 64 * a.shape[1] < (1 << 24)
    return (a @ w.T).astype(np.int32)


def gemm_w4a4(a_packed, sA, w_packed, sW, bias=None, out_dtype=F16, w_unpacked=None) -> np.ndarray:
    """D = T(((float(acc)*sA[m])*sW[n]) + bias[n]) -- fp32 epilogue in the EVT's order."""
    acc = gemm_w4a4_acc(a_packed, w_packed, w_unpacked).astype(F32)
    t = (acc * np.asarray(sA).astype(F32)[:, None]).astype(F32)
    t = (t * np.asarray(sW).astype(F32)[None, :]).astype(F32)
    if bias is not None:
        t = (t + np.asarray(bias).astype(F32)[None, :]).astype(F32)
    return _to16(t, out_dtype)


# ----------------------------------------------------------------------------------------------
# a4: W4A16 GEMM  (quarot_nn/linear.py:111-119; quarot_nn/qspec_gemm.py:65-88)
# ----------------------------------------------------------------------------------------------

def gemm_w4a16(x, w_packed, sW, bias=None, out_dtype=F16, w_unpacked=None) -> np.ndarray:
    """C = x @ (unpack_i4(W) * sW[:,None]).T (+bias).  Evaluated in float64 so the oracle is
    the *exact* value of the reference's statement of intent; the 1e-2 relative budget covers
    the reference's fp16-accumulate BitBLAS kernel and our fp32-accumulate one."""
    xf = _from16(x).astype(np.float64)
    # x @ (q * s).T == (x @ q.T) * s up to float64 rounding (~1e-16 relative, far below the f16 result's 5e-4)
    q = (unpack_i4(w_packed) if w_unpacked is None else w_unpacked).astype(np.float64)
    c = (xf @ q.T) * _from16(sW).astype(np.float64)[None, :]
    if bias is not None:
        c = c + _from16(bias).astype(np.float64)[None, :]
    return _to16(c.astype(F32), out_dtype)


# ----------------------------------------------------------------------------------------------
# bf16 helpers (numpy has no bfloat16: carried as uint16 bit patterns)
# ----------------------------------------------------------------------------------------------

def f32_to_bf16_bits(v: np.ndarray) -> np.ndarray:
    u = np.asarray(v, dtype=F32).view(np.uint32).astype(np.uint64)
    rounded = (u + 0x7FFF + ((u >> 16) & 1)) >> 16          # round to nearest even
    return rounded.astype(np.uint16)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.asarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(F32)


def _to16(v, out_dtype):
    if out_dtype == "bf16":
        return f32_to_bf16_bits(v)
    return np.asarray(v, dtype=F32).astype(F16)


def _from16(v):
    v = np.asarray(v)
    if v.dtype == np.uint16:
        return bf16_bits_to_f32(v)
    return v.astype(F32)


# ----------------------------------------------------------------------------------------------
# synthetic data of SURVEY 8(d)
# ----------------------------------------------------------------------------------------------

def make_weight(N: int, K: int, seed: int):
    rng = np.random.default_rng(seed)
    q = rng.integers(-8, 8, size=(N, K), dtype=np.int8)
    sW = ((0.5 + rng.random(N, dtype=F32)) * F32(2.0 ** -7)).astype(F16)
    return pack_i4(q), sW


def make_act(M: int, K: int, seed: int, heavy_tail: bool = False) -> np.ndarray:
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((M, K), dtype=F32)
    if heavy_tail:
        x = x * (1.0 + 20.0 * (rng.random((M, K)) < 0.001))
    return x.astype(F16)
