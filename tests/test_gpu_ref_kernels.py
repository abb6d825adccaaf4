"""Pins the activation-side kernels (and the numpy oracle) against the reference's OWN CUDA kernels.

oracle/build_ref.py compiles the reference's layernorm / fuse_sym_quant / fast-hadamard-transform sources, from where they lie under
/root/reference, into oracle/_ref/ (git-ignored; travels to the GPU box). tests/ref_kernels_worker.py runs them in a subprocess on
the same seeded inputs as this library's kernels and the oracle. Bars:

* K3 fuse_sym_quant (quant.cu:102-186, all-fp16, order-independent): packed bytes and scales IDENTICAL to the reference kernel.
* K4 FHT (fast_hadamard_transform_cuda.cu:124-198, fp32 butterflies in a fixed order): outputs IDENTICAL.
* K2 norm (layernorm_kernels.cu:569-720, built --use_fast_math, block-size-dependent fp32 reduction order): <= 1 f16 ulp.
* K1 norm + int4 quant: scale <= 1 f16 ulp; quantised values differ by at most 1 LSB in at most 0.5 % of the elements (values that sit
  on a rounding boundary move with the last bit of mean / rsqrt; the reference's own result depends on its fast-math build).

Skipped (not failed) when oracle/_ref was not built or a module does not load on the box.
"""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref_report():
    if not os.path.isdir(REF_DIR) or not any(f.endswith((".so", ".cubin")) for f in os.listdir(REF_DIR)):
        pytest.skip("oracle/_ref not built (python oracle/build_ref.py needs /root/reference)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ref_kernels_worker.py")], capture_output=True, text=True,
                       timeout=600, cwd=ROOT)
    lines = [l for l in r.stdout.splitlines() if l.startswith("REF_KERNELS_JSON ")]
    if r.returncode != 0 or not lines:
        pytest.skip(f"reference kernels did not run on this box (rc={r.returncode}): {r.stderr[-400:]}")
    rep = json.loads(lines[-1][len("REF_KERNELS_JSON "):])
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "ref_kernels_report.json"), "w") as f:
            json.dump(rep, f, indent=1)
    return rep


def _need(rep, module, key):
    if rep["modules"].get(module) != "loaded" or key not in rep:
        pytest.skip(f"{module}: {rep['modules'].get(module)}")
    return rep[key]


def test_fuse_sym_quant_identical_to_reference_kernel(ref_report):
    for case in _need(ref_report, "quarot_quant_ref", "K3_fuse_sym_quant"):
        for who in ("ours_vs_ref", "oracle_vs_ref"):
            c = case[who]
            assert c["bytes_equal"] and c["scales_equal"], (who, case)


def test_fht_identical_to_reference_kernel(ref_report):
    for case in _need(ref_report, "fast_hadamard_transform_cuda", "K4_fht"):
        for who in ("ours_vs_ref", "oracle_vs_ref"):
            assert case[who]["differ"] == 0, (who, case)


def test_norm_f16_within_one_ulp_of_reference_kernel(ref_report):
    for case in _need(ref_report, "layernorm_ops", "K2_norm_f16"):
        for who in ("ours_vs_ref", "oracle_vs_ref"):
            assert case[who]["max_ulp"] <= 1, (who, case)


def test_norm_quant_i4_against_reference_kernel(ref_report):
    for case in _need(ref_report, "layernorm_ops", "K1_norm_quant_i4"):
        assert case["ours_vs_oracle_equal"], case
        for who in ("ours_vs_ref", "oracle_vs_ref"):
            c = case[who]
            assert c["scale_max_ulp"] <= 1, (who, case)
            assert c["max_abs_diff"] <= 1, (who, case)
            assert c["differ"] <= 0.005 * c["values"], (who, case)


def test_w4a8_quantisers_against_reference_kernels(ref_report):
    """W4A8 mode: fuse_sym_quant_i8 identical; norm + int8 quant: scale <= 1 ulp, values within 1 LSB (at 127 steps per amax the
    fast-math rsqrt of the reference moves a few values across a rounding boundary)."""
    if "K3_fuse_sym_quant_i8" in ref_report:
        for case in ref_report["K3_fuse_sym_quant_i8"]:
            for who in ("ours_vs_ref", "oracle_vs_ref"):
                assert case[who]["q_equal"] and case[who]["scales_equal"], (who, case)
    for case in _need(ref_report, "layernorm_ops", "K1_norm_quant_i8"):
        for who in ("ours_vs_ref", "oracle_vs_ref"):
            c = case[who]
            assert c["scale_max_ulp"] <= 1 and c["max_abs_diff"] <= 1 and c["differ"] <= 0.02 * c["values"], (who, case)


def test_w4a16_within_tolerance_of_reference_triton_kernel(ref_report):
    """north_star: fp16 outputs within 1e-2 relative (max|d| / max|ref|), stated per layer shape. The checker is the reference's own Triton
    W4A16 kernel (quarot_nn/qspec_gemm.py, fp32 accumulate, weights pre-scaled in f16); BitBLAS itself is absent from the reference tree."""
    for case in _need(ref_report, "triton_w4a16_ref", "W4A16_triton"):
        assert case["ref_absmax"] > 0.1, case                            # the reference kernel really wrote an output
        assert case["ours_vs_ref_rel"] <= 1e-2, case
        assert case["oracle_vs_ref_rel"] <= 1e-2, case


# ----------------------------------------------------------------------------------------------
# this library's kernels (through the C ABI) vs the COMMITTED outputs of the reference's kernels
# (tests/golden/ref_kernels_b200.npz, made on a B200 by tests/golden/make_golden_gpu.py): runs on any
# GPU box, whether or not oracle/_ref travelled.
# ----------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def vec():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_kernels_b200.npz"))


def _bits(a):
    import numpy as np
    return np.ascontiguousarray(a).view(np.uint16)


def test_cuda_kernels_reproduce_reference_kernel_vectors(vec):
    import numpy as np
    import torch
    from qspec_b200 import ops
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    # K1
    x = cu(vec["k12_x"])
    q = torch.empty(12, 2048, dtype=torch.int8, device="cuda")
    s = torch.empty(12, dtype=torch.float16, device="cuda")
    isum = torch.empty_like(s)
    ops.rms_norm_general_fuse_sum_i4(q, x, isum, s, 1e-5, True)
    assert np.array_equal(q.cpu().numpy().view(np.uint8), vec["k1_q"])
    assert np.array_equal(_bits(s.cpu().numpy()), _bits(vec["k1_scale"]))
    # K2: <= 1 f16 ulp
    y = torch.empty_like(x)
    ops.rms_norm_general_fuse_sum_fp16(y, x, 1e-5)
    a, b = _bits(y.cpu().numpy()).astype(np.int32), _bits(vec["k2_y"]).astype(np.int32)
    key = lambda u: np.where(u & 0x8000, -(u & 0x7FFF), u & 0x7FFF)
    assert np.abs(key(a) - key(b)).max() <= 1
    # K3
    for tag in ("a", "b"):
        x3 = cu(vec[f"k3{tag}_x"])
        q3 = torch.empty(x3.shape[0], x3.shape[1] // 2, dtype=torch.int8, device="cuda")
        s3 = torch.empty(x3.shape[0], dtype=torch.float16, device="cuda")
        ops.fuse_sym_quant(x3, s3, q3, 1.0)
        assert np.array_equal(q3.cpu().numpy().view(np.uint8), vec[f"k3{tag}_q"])
        assert np.array_equal(_bits(s3.cpu().numpy()), _bits(vec[f"k3{tag}_scale"]))
    # K4 + the fused down_proj rotation (FHT_512 -> f16 -> had28 mix in one kernel)
    x4 = cu(vec["k4a_x"])
    o4 = torch.empty_like(x4)
    ops.faster_fast_hadamard_transform(x4, float(vec["k4a_scale"]), o4)
    assert np.array_equal(_bits(o4.cpu().numpy()), _bits(vec["k4a_y"]))
    x5 = cu(vec["k4b_x"])
    o5 = torch.empty_like(x5)
    ops.hadamard_k28(x5, 512, float(vec["k4b_scale"]), out=o5)
    assert np.array_equal(_bits(o5.cpu().numpy()), _bits(vec["k5_y"]))
