"""Timeline of one GEMM launch (CTA 0 and the last CTA): python scripts/gpu_gemm_trace.py MODE M N K
Prints per-role %globaltimer marks (ns relative to the earliest mark) recorded by the kernel."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qspec_b200 import ops, _lib  # noqa: E402

mode, M, N, K = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
lib = _lib.load()
ws = [torch.randint(-128, 127, (N, K // 2), dtype=torch.int8, device="cuda") for _ in range(6)]
sW = (torch.rand(N, device="cuda") * 0.01).half()
out = torch.empty(M, N, dtype=torch.float16, device="cuda")
if mode == "w4a16":
    x = torch.randn(M, K, device="cuda").half()
    fn = lambda w: ops.gemm_w4a16(x, w, sW, None, out)
else:
    a = torch.randint(-128, 127, (M, K // 2), dtype=torch.int8, device="cuda")
    sA = (torch.rand(M, device="cuda") * 0.1).half()
    fn = lambda w: ops.rowwise_scaled_linear_cutlass_s4s4_unified(a, sA, w, sW, None, out)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for w in ws[:4]:
    fn(w)
torch.cuda.synchronize()
trace = torch.zeros(2 * 4 * 256 + 64, dtype=torch.int64, device="cuda")
flush.zero_()
torch.cuda.synchronize()
lib.qspec_debug_set_gemm_trace(trace.data_ptr())
fn(ws[4])
fn(ws[5])           # back-to-back second launch overwrites the trace: shows the steady-state (PDL) case
torch.cuda.synchronize()
lib.qspec_debug_set_gemm_trace(None)
tc = trace.cpu()
prof = tc[2048:].tolist()
names = ["wait_w_full", "wait_a_empty", "cvt+sttm", "act_cvt", "fence_proxy", "wait_st", "arrive", "loop"]
nu = max(1, -(-(N // 128) * (K // 256) // lib.qspec_sm_count()))   # units of CTA 0 (ceil)
for w, off in (("conv warp0 (h=0)", 0), ("conv warp4 (h=1)", 8)):
    print(w, "cycles/unit:", " ".join(f"{n}={prof[off + k] / nu:.0f}" for k, n in enumerate(names)), "total/unit=%.0f" % (sum(prof[off:off + 8]) / nu))
print("mma cycles/unit: wait_d_empty=%.0f wait_a_full=%.0f issue+commit=%.0f other=%.0f" % tuple(prof[16 + k] / nu for k in range(4)))
