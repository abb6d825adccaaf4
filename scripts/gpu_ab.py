"""Same-box A/B of GEMM variants: python scripts/gpu_ab.py  (select the library with QSPEC_LIB, the kernel form with
QSPEC_GEMM_PAIR / QSPEC_GEMM_PAIR_MIN_M). Prints one line per (shape, mode): us per call at M = 4, 16, 64, 256."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_microbench import bench_gemm, SHAPES_8B, SHAPES_70B
shapes = dict(SHAPES_8B)
shapes["gate_up70"] = SHAPES_70B["gate_up70"]
tag = sys.argv[1] if len(sys.argv) > 1 else "?"
for mode in ("w4a4", "w4a16"):
    for name, (N, K) in shapes.items():
        row = {"tag": tag, "mode": mode, "shape": name}
        for M in (4, 16, 64, 256):
            row[f"M{M}"] = bench_gemm(name, N, K, M, mode)["us_avg"]
        print(json.dumps(row), flush=True)
