"""What does a kernel boundary cost in the PDL chain? Times graph-replayed chains of (GEMM, norm) pairs against chains
of GEMMs only and of norms only, for several GEMM grid sizes (QSPEC_GEMM_G)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qspec_b200 import ops
from gpu_microbench import time_graph

M, N, K = 4, 4096, 4096
dev = "cuda"
ws = [torch.randint(-128, 127, (N, K // 2), dtype=torch.int8, device=dev) for _ in range(32)]
sW = (torch.rand(N, device=dev) * 0.01).half()
a = torch.randint(-128, 127, (M, K // 2), dtype=torch.int8, device=dev)
sA = (torch.rand(M, device=dev) * 0.1).half()
out = torch.empty(M, N, dtype=torch.float16, device=dev)
x = torch.randn(M, N, device=dev).half()
q = torch.empty(M, N // 2, dtype=torch.int8, device=dev)
s = torch.empty(M, dtype=torch.float16, device=dev)
isum = torch.empty_like(s)

def gemm(w):
    return lambda: ops.rowwise_scaled_linear_cutlass_s4s4_unified(a, sA, w, sW, None, out)
def norm():
    return lambda: ops.add_rms_norm_fuse_sum_i4(q, x, out, isum, s, 1e-5)

for g in ("", "32", "64", "128"):
    if g:
        os.environ["QSPEC_GEMM_G"] = g
    else:
        os.environ.pop("QSPEC_GEMM_G", None)
    t_g = time_graph([gemm(w) for w in ws])[1]
    t_n = time_graph([norm() for _ in ws])[1]
    pair = []
    for w in ws:
        pair += [gemm(w), norm()]
    t_p = time_graph(pair)[1] * 2
    print(f"G={g or 'default'}: gemm chain {t_g:.2f} us, norm chain {t_n:.2f} us, (gemm, norm) pair {t_p:.2f} us, "
          f"pair - gemm - norm = {t_p - t_g - t_n:+.2f} us", flush=True)
