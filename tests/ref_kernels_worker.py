"""Runs the reference's OWN CUDA kernels (oracle/_ref, built by oracle/build_ref.py from the sources under /root/reference) next
to this library's kernels and the numpy oracle on the same seeded inputs; prints one JSON object.

Launched in a subprocess by tests/test_gpu_ref_kernels.py so that a fault inside third-party kernel code cannot poison the CUDA
context of the test session. Needs a GPU; does not read /root/reference (only the prebuilt modules).

Reference call sites reproduced here:
  K1  layernorm_ops.rms_norm_general_fuse_sum_i4(out_q, x, input_sum, scaling, eps, True)   quarot_nn/normalization.py:53-60
  K2  layernorm_ops.rms_norm_general_fuse_sum_fp16(out, x, eps)                              quarot_nn/normalization.py:75-79
  K3  quarot._CUDA.fuse_sym_quant(x, scale, out, clip)                                       third-party/QuaRot/quarot/__init__.py:119-144
  K4  fast_hadamard_transform_cuda.faster_fast_hadamard_transform(x, scale, out)             quarot/functional/hadamard.py:98-126
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))

from oracle import qspec_oracle as O  # noqa: E402
from qspec_b200 import ops  # noqa: E402

EPS = 1e-5


def _load(name):
    try:
        return __import__(name), None
    except Exception as e:  # missing build, ABI mismatch, ...
        return None, f"{type(e).__name__}: {str(e)[:300]}"


def _ulp16(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """distance in f16 representable steps (sign-magnitude -> ordered ints)."""
    def key(v):
        u = v.view(np.uint16).astype(np.int32)
        return np.where(u & 0x8000, -(u & 0x7FFF), u & 0x7FFF)
    return np.abs(key(a) - key(b))


def _nib_stats(pa: np.ndarray, pb: np.ndarray):
    qa, qb = O.unpack_i4(pa).astype(np.int32), O.unpack_i4(pb).astype(np.int32)
    d = np.abs(qa - qb)
    return {"values": int(d.size), "differ": int((d != 0).sum()), "max_abs_diff": int(d.max(initial=0))}


def acts(T, K, seed, edge=False):
    x = O.make_act(T, K, seed, heavy_tail=True)
    if edge and T >= 3:
        x[1, :] = 0                       # amax = 0 row (K3: scale 0 -> NaN -> 0; K1: amax floor 1e-6)
        x[2, 5] = np.float16(65504.0)     # extreme element
        x[2, 6] = np.float16(-65504.0)
    return x


def run_k3(mod, res):
    out = []
    for T, K, seed, edge in ((1, 4096, 11, False), (4, 4096, 12, True), (16, 4096, 13, True), (33, 4096, 14, True),
                             (4, 14336, 15, True), (16, 14336, 16, False)):
        x = acts(T, K, seed, edge)
        xd = torch.from_numpy(x).cuda()
        rq = torch.zeros(T, K // 2, dtype=torch.int8, device="cuda")
        rs = torch.zeros(T, dtype=torch.float16, device="cuda")
        mod.fuse_sym_quant(xd, rs, rq, 1.0)
        mq = torch.zeros_like(rq)
        ms = torch.zeros_like(rs)
        ops.fuse_sym_quant(xd, ms, mq, 1.0)
        torch.cuda.synchronize()
        oq, osc = O.sym_quant_i4(x, 1.0)
        rqn, rsn = rq.cpu().numpy().view(np.uint8), rs.cpu().numpy()
        mqn, msn = mq.cpu().numpy().view(np.uint8), ms.cpu().numpy()
        out.append({"T": T, "K": K, "edge": edge,
                    "ours_vs_ref": {"bytes_equal": bool(np.array_equal(mqn, rqn)), "scales_equal": bool(np.array_equal(msn.view(np.uint16), rsn.view(np.uint16))),
                                    **_nib_stats(mqn, rqn)},
                    "oracle_vs_ref": {"bytes_equal": bool(np.array_equal(oq, rqn)), "scales_equal": bool(np.array_equal(osc.view(np.uint16), rsn.view(np.uint16))),
                                      **_nib_stats(oq, rqn)}})
    res["K3_fuse_sym_quant"] = out
    if not hasattr(mod, "fuse_sym_quant_i8"):
        return
    out8 = []                                            # W4A8 quantiser (quant.cu:188-245)
    for T, K, seed, edge in ((4, 4096, 17, True), (16, 4096, 18, True), (6, 14336, 19, True)):
        x = acts(T, K, seed, edge)
        xd = torch.from_numpy(x).cuda()
        rq = torch.zeros(T, K, dtype=torch.int8, device="cuda")
        rs = torch.zeros(T, dtype=torch.float16, device="cuda")
        mod.fuse_sym_quant_i8(xd, rs, rq, 1.0)
        mq, ms = torch.zeros_like(rq), torch.zeros_like(rs)
        ops.fuse_sym_quant_i8(xd, ms, mq, 1.0)
        torch.cuda.synchronize()
        oq, osc = O.sym_quant_i8(x, 1.0)
        rqn, rsn = rq.cpu().numpy(), rs.cpu().numpy().view(np.uint16)
        out8.append({"T": T, "K": K,
                     "ours_vs_ref": {"q_equal": bool(np.array_equal(mq.cpu().numpy(), rqn)), "scales_equal": bool(np.array_equal(ms.cpu().numpy().view(np.uint16), rsn))},
                     "oracle_vs_ref": {"q_equal": bool(np.array_equal(oq, rqn)), "scales_equal": bool(np.array_equal(osc.view(np.uint16), rsn))}})
    res["K3_fuse_sym_quant_i8"] = out8


def run_k1k2(mod, res):
    k1, k2 = [], []
    for T, K, seed, edge in ((1, 4096, 21, False), (4, 4096, 22, True), (16, 4096, 23, True), (33, 4096, 24, True), (64, 4096, 25, False),
                             (4, 8192, 26, False)):
        x = acts(T, K, seed, edge)
        if edge:
            x[2, 5] = np.float16(300.0)   # keep the row finite in fp32 variance; still a heavy outlier
            x[2, 6] = np.float16(-300.0)
        xd = torch.from_numpy(x).cuda()
        # --- K1
        rq = torch.zeros(T, K // 2, dtype=torch.int8, device="cuda")
        rsum = torch.zeros(T, dtype=torch.float16, device="cuda")
        rs = torch.zeros(T, dtype=torch.float16, device="cuda")
        mod.rms_norm_general_fuse_sum_i4(rq, xd, rsum, rs, EPS, True)
        mq, msum, ms = torch.zeros_like(rq), torch.zeros_like(rsum), torch.zeros_like(rs)
        ops.rms_norm_general_fuse_sum_i4(mq, xd, msum, ms, EPS, True)
        torch.cuda.synchronize()
        oq, osc, _ = O.norm_quant_i4(x, EPS)
        rqn, rsn = rq.cpu().numpy().view(np.uint8), rs.cpu().numpy()
        mqn, msn = mq.cpu().numpy().view(np.uint8), ms.cpu().numpy()
        k1.append({"T": T, "K": K, "edge": edge,
                   "ours_vs_ref": {**_nib_stats(mqn, rqn), "scale_max_ulp": int(_ulp16(msn, rsn).max())},
                   "oracle_vs_ref": {**_nib_stats(oq, rqn), "scale_max_ulp": int(_ulp16(osc, rsn).max())},
                   "ours_vs_oracle_equal": bool(np.array_equal(mqn, oq) and np.array_equal(msn.view(np.uint16), osc.view(np.uint16)))})
        # --- K2
        ry = torch.zeros_like(xd)
        mod.rms_norm_general_fuse_sum_fp16(ry, xd, EPS)
        my = torch.zeros_like(xd)
        ops.rms_norm_general_fuse_sum_fp16(my, xd, EPS)
        torch.cuda.synchronize()
        oy = O.norm_f16(x, EPS)
        ryn, myn = ry.cpu().numpy(), my.cpu().numpy()
        u_m, u_o = _ulp16(myn, ryn), _ulp16(oy, ryn)
        k2.append({"T": T, "K": K, "edge": edge,
                   "ours_vs_ref": {"values": int(u_m.size), "differ": int((u_m != 0).sum()), "max_ulp": int(u_m.max())},
                   "oracle_vs_ref": {"values": int(u_o.size), "differ": int((u_o != 0).sum()), "max_ulp": int(u_o.max())}})
    res["K1_norm_quant_i4"] = k1
    res["K2_norm_f16"] = k2
    k8 = []                                              # W4A8 norm + int8 quantiser (layernorm_kernels.cu:350-470)
    for T, K, seed in ((4, 4096, 27), (16, 4096, 28), (33, 4096, 29), (4, 8192, 30)):
        x = acts(T, K, seed, True)
        if T >= 3:
            x[2, 5], x[2, 6] = np.float16(300.0), np.float16(-300.0)
        xd = torch.from_numpy(x).cuda()
        rq = torch.zeros(T, K, dtype=torch.int8, device="cuda")
        rsum, rs = torch.zeros(T, dtype=torch.float16, device="cuda"), torch.zeros(T, dtype=torch.float16, device="cuda")
        mod.rms_norm_general_fuse_sum_i8(rq, xd, rsum, rs, EPS, True)
        mq, msum, ms = torch.zeros_like(rq), torch.zeros_like(rsum), torch.zeros_like(rs)
        ops.rms_norm_general_fuse_sum_i8(mq, xd, msum, ms, EPS, True)
        torch.cuda.synchronize()
        oq, osc, _ = O.norm_quant_i8(x, EPS)
        rqn = rq.cpu().numpy().astype(np.int32)
        dm, do = np.abs(mq.cpu().numpy().astype(np.int32) - rqn), np.abs(oq.astype(np.int32) - rqn)
        k8.append({"T": T, "K": K,
                   "ours_vs_ref": {"values": int(dm.size), "differ": int((dm != 0).sum()), "max_abs_diff": int(dm.max()),
                                   "scale_max_ulp": int(_ulp16(ms.cpu().numpy(), rs.cpu().numpy()).max())},
                   "oracle_vs_ref": {"values": int(do.size), "differ": int((do != 0).sum()), "max_abs_diff": int(do.max()),
                                     "scale_max_ulp": int(_ulp16(osc, rs.cpu().numpy()).max())}})
    res["K1_norm_quant_i8"] = k8


def run_k4(mod, res):
    out = []
    s32 = float(1.0 / torch.tensor(32).sqrt())           # quarot_nn/hadamard.py:12-13
    s14336 = float(1.0 / torch.tensor(14336).sqrt())
    for rows, n, scale, seed in ((4 * 128, 32, s32, 31), (16 * 128, 32, s32, 32), (4 * 28, 512, s14336, 33), (16 * 28, 512, s14336, 34),
                                 (15, 256, 1.0 / 16.0, 35), (15, 2048, 2.0 ** -5.5, 36)):
        x = O.make_act(rows, n, seed, heavy_tail=True)
        xd = torch.from_numpy(x).cuda()
        ro = torch.zeros_like(xd)
        mod.faster_fast_hadamard_transform(xd, scale, ro)
        mo = torch.zeros_like(xd)
        ops.faster_fast_hadamard_transform(xd, scale, mo)
        torch.cuda.synchronize()
        oo = O.fht(x, scale)
        ron, mon = ro.cpu().numpy(), mo.cpu().numpy()
        u_m, u_o = _ulp16(mon, ron), _ulp16(oo, ron)
        out.append({"rows": rows, "n": n,
                    "ours_vs_ref": {"values": int(u_m.size), "differ": int((u_m != 0).sum()), "max_ulp": int(u_m.max())},
                    "oracle_vs_ref": {"values": int(u_o.size), "differ": int((u_o != 0).sum()), "max_ulp": int(u_o.max())}})
    res["K4_fht"] = out


def run_w4a16_triton(res):
    """The reference's own Triton statement of the W4A16 GEMM (quarot_nn/qspec_gemm.py:20-88, fp32 accumulate), AOT-compiled by
    oracle/build_ref.py; launched with the wrapper's grid (qspec_gemm.py:114-116) through the CUDA driver API."""
    import ctypes
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    meta = json.load(open(os.path.join(ref_dir, "triton_w4a16_ref.json")))
    image = open(os.path.join(ref_dir, "triton_w4a16_ref.cubin"), "rb").read()
    cu = ctypes.CDLL("libcuda.so.1")
    torch.zeros(1, device="cuda")                                        # primary context current on this thread
    mod, fn = ctypes.c_void_p(), ctypes.c_void_p()
    assert cu.cuModuleLoadData(ctypes.byref(mod), image) == 0, "cuModuleLoadData"
    assert cu.cuModuleGetFunction(ctypes.byref(fn), mod, meta["name"].encode()) == 0, "cuModuleGetFunction"
    out = []
    for M, N, K, seed in ((1, 4096, 4096, 41), (4, 4096, 4096, 42), (16, 6144, 4096, 43), (16, 4096, 14336, 44), (64, 4096, 4096, 45),
                          (13, 28672, 4096, 46)):
        wp, sW = O.make_weight(N, K, 1000 + seed)
        x = O.make_act(M, K, 2000 + seed)
        xd, wd, sd = torch.from_numpy(x).cuda(), torch.from_numpy(wp).view(torch.int8).cuda(), torch.from_numpy(sW).cuda()
        ref = torch.zeros(M, N, dtype=torch.float16, device="cuda")
        args = [ctypes.c_void_p(xd.data_ptr()), ctypes.c_void_p(wd.data_ptr()), ctypes.c_void_p(ref.data_ptr()), ctypes.c_void_p(sd.data_ptr()),
                ctypes.c_int(M), ctypes.c_int(N), ctypes.c_int(K), ctypes.c_void_p(0), ctypes.c_void_p(0)]
        argv = (ctypes.c_void_p * len(args))(*[ctypes.cast(ctypes.byref(a), ctypes.c_void_p) for a in args])
        grid = ((M + meta["block_m"] - 1) // meta["block_m"]) * ((N + meta["block_n"] - 1) // meta["block_n"])
        torch.cuda.synchronize()
        rc = cu.cuLaunchKernel(fn, grid, 1, 1, 32 * meta["num_warps"], 1, 1, meta["shared"], ctypes.c_void_p(0), argv, None)
        assert rc == 0, f"cuLaunchKernel rc={rc}"
        torch.cuda.synchronize()
        mine = torch.zeros(M, N, dtype=torch.float16, device="cuda")
        ops.gemm_w4a16(xd, wd, sd, None, mine)
        torch.cuda.synchronize()
        r = ref.float().cpu().numpy()
        orc = O.gemm_w4a16(x, wp, sW).astype(np.float32)
        den = float(np.abs(r).max())
        out.append({"M": M, "N": N, "K": K, "ref_absmax": den,
                    "ours_vs_ref_rel": float(np.abs(mine.float().cpu().numpy() - r).max() / den),
                    "oracle_vs_ref_rel": float(np.abs(orc - r).max() / den)})
    res["W4A16_triton"] = out


def main():
    res = {"device": torch.cuda.get_device_name(0), "modules": {}}
    try:
        run_w4a16_triton(res)
        res["modules"]["triton_w4a16_ref"] = "loaded"
    except Exception as e:
        res["modules"]["triton_w4a16_ref"] = f"{type(e).__name__}: {str(e)[:300]}"
    for name, fn in (("quarot_quant_ref", run_k3), ("layernorm_ops", run_k1k2), ("fast_hadamard_transform_cuda", run_k4)):
        mod, err = _load(name)
        res["modules"][name] = "loaded" if mod is not None else err
        if mod is None:
            continue
        try:
            fn(mod, res)
        except Exception as e:
            res["modules"][name] = f"run failed: {type(e).__name__}: {str(e)[:300]}"
    print("REF_KERNELS_JSON " + json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
