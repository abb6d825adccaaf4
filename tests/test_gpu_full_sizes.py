"""BASELINE.json full sizes (configs[2]-[4]: 64 / 32 concurrent sequences, k = 3 => draft M = 64 / 32, verify M = 256 / 128; Llama-3-8B
and 70B shapes, TP=8 shards of 70B), checked through size-independent properties so the CPU side stays cheap:

* W4A4 int32 accumulators: checksum of checksums -- the column sums must equal (sum_m a[m]) . w[n] and the row sums a[m] . (sum_n w[n]),
  both exact integers -- plus the full dot products of 96 sampled channels against the oracle. Any misplaced tile, dropped k-unit, wrong
  split-K reduction or token-tile mix-up changes at least one of them.
* W4A16 f16 outputs: 96 sampled channels against the float64 oracle within north_star's 1e-2 relative, and run-to-run determinism of the
  whole output.
"""
import numpy as np
import pytest
import torch

from oracle import qspec_oracle as O

pytestmark = pytest.mark.gpu

FULL = [
    pytest.param(64, 6144, 4096, id="8b-qkv-draft64"),
    pytest.param(256, 6144, 4096, id="8b-qkv-verify256"),
    pytest.param(256, 28672, 4096, id="8b-gate_up-verify256"),
    pytest.param(256, 4096, 14336, id="8b-down-verify256"),
    pytest.param(128, 57344, 8192, id="70b-gate_up-verify128"),
    pytest.param(32, 8192, 28672, id="70b-down-draft32"),
    pytest.param(128, 7168, 8192, id="70b-gate_up-tp8-shard"),
    pytest.param(128, 8192, 3584, id="70b-down-tp8-shard"),
]


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from qspec_b200 import ops as _ops
    return _ops


def _rand_packed(rows, K, seed):
    """uniform bytes == uniform nibble pairs; returns packed uint8 [rows, K/2]."""
    return np.random.default_rng(seed).integers(0, 256, (rows, K // 2), dtype=np.uint8)


def _rows_dot(wu: np.ndarray, v: np.ndarray) -> np.ndarray:
    """wu int8 [N, K] times int64 vector v [K] -> int64 [N], exact (float64 BLAS on row chunks: |sum| << 2^53)."""
    out = np.empty(wu.shape[0], dtype=np.int64)
    vf = v.astype(np.float64)
    for r in range(0, wu.shape[0], 4096):
        out[r:r + 4096] = np.rint(wu[r:r + 4096].astype(np.float64) @ vf).astype(np.int64)
    return out


@pytest.mark.parametrize("M,N,K", FULL)
def test_w4a4_accumulators_full_size_checksums(ops, M, N, K):
    a, w = _rand_packed(M, K, 11 + M), _rand_packed(N, K, 13 + N)
    acc = ops.gemm_w4a4_acc_i32(torch.from_numpy(a).view(torch.int8).cuda(), torch.from_numpy(w).view(torch.int8).cuda())
    got = acc.cpu().numpy().astype(np.int64)
    au, wu = O.unpack_i4(a).astype(np.int64), O.unpack_i4(w)               # wu stays int8 (up to 470 M values)
    assert np.array_equal(got.sum(axis=0), _rows_dot(wu, au.sum(axis=0)))                  # per-channel checksum over tokens
    assert np.array_equal(got.sum(axis=1), au @ wu.sum(axis=0, dtype=np.int64))            # per-token checksum over channels
    sel = np.random.default_rng(5).choice(N, size=96, replace=False)
    sel[:3] = (0, N // 2 + 127, N - 1)
    assert np.array_equal(got[:, sel], au @ wu[sel].astype(np.int64).T)


@pytest.mark.parametrize("M,N,K", FULL)
def test_w4a16_outputs_full_size_sampled(ops, M, N, K):
    w = _rand_packed(N, K, 17 + N)
    rng = np.random.default_rng(19 + M)
    sW = ((0.5 + rng.random(N, dtype=np.float32)) * np.float32(2.0 ** -7)).astype(np.float16)
    x = O.make_act(M, K, 3000 + M, heavy_tail=True)
    xd, wd, sd = torch.from_numpy(x).cuda(), torch.from_numpy(w).view(torch.int8).cuda(), torch.from_numpy(sW).cuda()
    out = torch.empty(M, N, dtype=torch.float16, device="cuda")
    ops.gemm_w4a16(xd, wd, sd, None, out)
    out2 = torch.empty_like(out)
    ops.gemm_w4a16(xd, wd, sd, None, out2)
    assert torch.equal(out, out2)                                          # fixed reduction order
    sel = np.random.default_rng(7).choice(N, size=96, replace=False)
    sel[:3] = (0, N // 2 + 127, N - 1)
    ref = O.gemm_w4a16(x, w[sel], sW[sel]).astype(np.float32)
    got = out.float().cpu().numpy()[:, sel]
    assert np.abs(got - ref).max() / np.abs(ref).max() <= 1e-2
