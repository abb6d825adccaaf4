"""Grid-size sweep of the stream-K GEMM per decode shape: QSPEC_GEMM_G forces the number of CTAs (read per call)."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qspec_b200 import ops
from gpu_microbench import bench_gemm, SHAPES_8B

SHAPES = dict(SHAPES_8B)
SHAPES.update({"o_tp2": (4096, 2048), "down_tp2": (4096, 7168), "qkv_tp2": (3072, 4096), "gate_up_tp2": (14336, 4096)})
only = sys.argv[1:] or list(SHAPES)
for name in only:
    N, K = SHAPES[name]
    tiles, KB = N // 128, K // 256
    U = tiles * KB
    cands = {0}
    for q in range(1, KB + 1):
        if KB % q == 0 and U // q <= 148:
            cands.add(U // q)
    for m in range(1, 9):
        if tiles % m == 0 and tiles // m <= 148:
            cands.add(tiles // m)
    cands |= {148, 144, 128, 112, 96, 74, 64}
    for mode, M in (("w4a4", 4), ("w4a16", 16)):
        row = []
        for g in sorted(c for c in cands if c <= U):
            if g:
                os.environ["QSPEC_GEMM_G"] = str(g)
            else:
                os.environ.pop("QSPEC_GEMM_G", None)
            ops.reserve_workspaces(16, [(N, K)], torch.device("cuda")) if hasattr(ops, "reserve_workspaces") else None
            r = bench_gemm(name, N, K, M, mode)
            row.append((g, r["us_avg"]))
        os.environ.pop("QSPEC_GEMM_G", None)
        best = min(row[1:], key=lambda t: t[1]) if len(row) > 1 else row[0]
        print(f"{name:12s} {mode:6s} tiles={tiles} KB={KB} default={row[0][1]:.2f}us best G={best[0]} {best[1]:.2f}us | " +
              " ".join(f"{g}:{t:.2f}" for g, t in row[1:]), flush=True)
