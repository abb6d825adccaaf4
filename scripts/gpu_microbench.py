"""Per-kernel timing on the GPU box (CUDA events, rotating weight copies > L2, CUDA-graph replay).

python scripts/gpu_microbench.py [--quick] > gpurun_out/microbench.json
Reports, per shape and M: us per call, algorithmic GB/s (SURVEY 8(d) byte counts) and TOPS.
"""
import argparse
import json
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qspec_b200 import ops  # noqa: E402

SHAPES_8B = {"qkv": (6144, 4096), "o": (4096, 4096), "gate_up": (28672, 4096), "down": (4096, 14336)}
SHAPES_70B = {"qkv70": (10240, 8192), "o70": (8192, 8192), "gate_up70": (57344, 8192), "down70": (8192, 28672)}


def time_graph(fn_list, iters=20):
    """fn_list: callables launched back to back inside one CUDA graph; returns us per callable."""
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for f in fn_list:
            f()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for f in fn_list:
                f()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e30
        tot = 0.0
        for _ in range(iters):
            e0.record(s)
            g.replay()
            e1.record(s)
            e1.synchronize()
            t = e0.elapsed_time(e1) * 1e3 / len(fn_list)
            best = min(best, t)
            tot += t
    return best, tot / iters


def bench_gemm(name, N, K, M, mode, copies_bytes=400 << 20):
    wbytes = N * K // 2
    ncopy = max(2, min(64, copies_bytes // wbytes))
    ws = [torch.randint(-128, 127, (N, K // 2), dtype=torch.int8, device="cuda") for _ in range(ncopy)]
    sW = (torch.rand(N, device="cuda") * 0.01).half()
    out = torch.empty(M, N, dtype=torch.float16, device="cuda")
    if mode == "w4a16":
        x = torch.randn(M, K, device="cuda").half()
        fns = [lambda w=w: ops.gemm_w4a16(x, w, sW, None, out) for w in ws]
        alg = N * K // 2 + 2 * N + 2 * M * K + 2 * M * N
    elif mode == "w4a8":
        a = torch.randint(-128, 127, (M, K), dtype=torch.int8, device="cuda")
        sA = (torch.rand(M, device="cuda") * 0.1).half()
        fns = [lambda w=w: ops.rowwise_scaled_linear_cutlass_s8s4(a, sA, w, sW, None, out) for w in ws]
        alg = N * K // 2 + 2 * N + M * K + 2 * M + 2 * M * N
    else:
        a = torch.randint(-128, 127, (M, K // 2), dtype=torch.int8, device="cuda")
        sA = (torch.rand(M, device="cuda") * 0.1).half()
        fns = [lambda w=w: ops.rowwise_scaled_linear_cutlass_s4s4_unified(a, sA, w, sW, None, out) for w in ws]
        alg = N * K // 2 + 2 * N + M * K // 2 + 2 * M + 2 * M * N
    reps = max(1, 64 // ncopy)
    best, avg = time_graph(fns * reps)
    return {"kernel": mode, "shape": name, "N": N, "K": K, "M": M, "us_best": round(best, 3), "us_avg": round(avg, 3),
            "alg_bytes": alg, "GBps": round(alg / avg / 1e3, 1), "TOPS": round(2.0 * M * N * K / avg / 1e6, 2)}


def bench_act(M):
    res = []
    x = torch.randn(M, 4096, device="cuda").half()
    q = torch.empty(M, 2048, dtype=torch.int8, device="cuda")
    s = torch.empty(M, dtype=torch.float16, device="cuda")
    isum = torch.empty_like(s)
    o = torch.empty_like(x)
    xm = torch.randn(M, 14336, device="cuda").half()
    gu = torch.randn(M, 28672, device="cuda").half()
    qm = torch.empty(M, 7168, dtype=torch.int8, device="cuda")
    cases = {
        "norm_quant_i4[4096]": lambda: ops.rms_norm_general_fuse_sum_i4(q, x, isum, s, 1e-5, True),
        "norm_f16[4096]": lambda: ops.rms_norm_general_fuse_sum_fp16(o, x, 1e-5),
        "sym_quant[4096]": lambda: ops.fuse_sym_quant(x, s, q, 1.0),
        "hadamard_heads+quant[32x128]": lambda: ops.hadamard_heads(x, 32, 128, 0.1767767, out_q=q, scale_out=s),
        "hadamard_k28+quant[28x512]": lambda: ops.hadamard_k28(xm, 512, 0.00835, out_q=qm, scale_out=s),
        "silu+hadamard_k28+quant[28x512]": lambda: ops.hadamard_k28(None, 512, 0.00835, gate_up=gu, out_q=qm, scale_out=s),
    }
    for name, fn in cases.items():
        best, avg = time_graph([fn] * 32)
        res.append({"kernel": name, "M": M, "us_best": round(best, 3), "us_avg": round(avg, 3)})
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--w4a8", action="store_true", help="only the W4A8 mode next to W4A4 on the 8B shapes, M in {4, 16, 64}")
    args = ap.parse_args()
    if args.w4a8:
        for name, (N, K) in SHAPES_8B.items():
            for M in (4, 16, 64):
                for mode in ("w4a4", "w4a8"):
                    print(json.dumps(bench_gemm(name, N, K, M, mode)), flush=True)
        return
    results = []
    Ms = [1, 4, 16, 64] if args.quick else [1, 4, 16, 64, 256]
    shapes = dict(SHAPES_8B)
    if not args.quick:
        shapes.update(SHAPES_70B)
    for name, (N, K) in shapes.items():
        for M in Ms:
            for mode in ("w4a16", "w4a4"):
                try:
                    r = bench_gemm(name, N, K, M, mode)
                except Exception as e:  # report and keep going
                    r = {"kernel": mode, "shape": name, "M": M, "error": str(e)[:200]}
                print(json.dumps(r), flush=True)
                results.append(r)
    for M in (4, 16, 64):
        for r in bench_act(M):
            print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
