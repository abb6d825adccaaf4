"""Grid-size sweep of the stream-K GEMM on the per-rank shapes of TP = 1/2/4/8 at the token counts of BASELINE configs[2]
(W4A4 M = 64, W4A16 M = 256): QSPEC_GEMM_G forces the number of CTAs (or CTA pairs), QSPEC_GEMM_PAIR the cta_group form.
Output: one line per (shape, mode, form) with the default plan's time and every forced grid's, for refitting make_plan's cost model."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qspec_b200 import ops  # noqa: F401
from gpu_microbench import bench_gemm

tps = [int(t) for t in (sys.argv[1] if len(sys.argv) > 1 else "2,4,8").split(",")]
for tp in tps:
    shapes = {"qkv": (6144 // tp, 4096), "o": (4096, 4096 // tp), "gate_up": (28672 // tp, 4096), "down": (4096, 14336 // tp)}
    for name, (N, K) in shapes.items():
        for mode, M in (("w4a4", 64), ("w4a16", 256)):
            for pair in ((0,) if M < 128 else (0, 1)):
                if pair and N % 256:
                    continue
                os.environ["QSPEC_GEMM_PAIR"] = str(pair)
                sms = 74 if pair else 148
                tiles, KB = N // (256 if pair else 128) * ((M + 127) // 128), K // 256
                U = tiles * KB
                cands = set()
                for cs in range(1, KB + 1):
                    if tiles * cs <= sms:
                        cands.add(tiles * cs)
                for m in range(2, 9):
                    if tiles % m == 0:
                        cands.add(tiles // m)
                cands |= {sms, sms - sms % 4, sms // 2}
                row = []
                os.environ.pop("QSPEC_GEMM_G", None)
                d = bench_gemm(name, N, K, M, mode)["us_avg"]
                for g in sorted(c for c in cands if 1 <= c <= min(U, sms)):
                    os.environ["QSPEC_GEMM_G"] = str(g)
                    row.append((g, bench_gemm(name, N, K, M, mode)["us_avg"]))
                os.environ.pop("QSPEC_GEMM_G", None)
                best = min(row, key=lambda t: t[1])
                print(f"tp{tp} {name:8s} {N}x{K} {mode} M={M} pair={pair} tiles={tiles} KB={KB} default={d:.2f} best G={best[0]} {best[1]:.2f} | " +
                      " ".join(f"{g}:{t:.2f}" for g, t in row), flush=True)
os.environ.pop("QSPEC_GEMM_PAIR", None)
