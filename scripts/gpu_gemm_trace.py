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
trace = torch.zeros(8 * 4096, dtype=torch.int64, device="cuda")
flush.zero_()
torch.cuda.synchronize()
lib.qspec_debug_set_gemm_trace(trace.data_ptr())
# four launches replayed from one CUDA graph (back to back, PDL edges); every launch overwrites the trace, so what is
# read back is the LAST one: the steady-state case with a predecessor still running when it starts
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        for w in (ws[4], ws[5], ws[0], ws[1]):
            fn(w)
    g.replay()
torch.cuda.synchronize()
lib.qspec_debug_set_gemm_trace(None)
tc = trace.cpu().view(8, 4096)
names = ["wait_w_full", "wait_a_empty", "cvt+sttm", "act_cvt", "fence_proxy", "wait_st", "arrive", "loop"]
nu = max(1, -(-(N // 128) * (K // 256) // lib.qspec_sm_count()))   # units of CTA 0 (ceil)
prof = tc[3, 2048:].tolist()
for w, off in (("conv warp0 (h=0)", 0), ("conv warp4 (h=1)", 8)):
    print(w, "cycles/unit:", " ".join(f"{n}={prof[off + k] / nu:.0f}" for k, n in enumerate(names)), "total/unit=%.0f" % (sum(prof[off:off + 8]) / nu))
print("mma cycles/unit: wait_d_empty=%.0f wait_a_full=%.0f (unused=%.0f) other=%.0f | umma_x4=%.0f commit=%.0f" % tuple(prof[16 + k] / nu for k in range(6)))
mk = ["start", "setup_done", "W_ring_issued", "pdl_wait_passed", "first_w_full", "convert_done", "f0_d_full", "f0_stored", "f0_reduced",
      "f1_d_full", "f1_stored", "f1_reduced", "f2_d_full", "f2_stored", "f2_reduced", "end"]
allm = tc[:4, 2048 + 64:2048 + 96]
t0 = int(allm[allm > 0].min())
for launch in range(4):
    marks = tc[launch, 2048 + 64:2048 + 96].tolist()
    for c, off in (("cta0", 0), ("ctaL", 16)):
        print(f"launch{launch} {c}:", " ".join(f"{mk[i]}={marks[off + i] - t0}" for i in range(16) if marks[off + i] > 0))

G = lib.qspec_sm_count()
for launch in range(4):
    st_ = tc[launch, 512:512 + G] - t0
    en_ = tc[launch, 1024:1024 + G] - t0
    print(f"launch{launch} all CTAs: start min/max = {int(st_.min())}/{int(st_.max())}  end min/median/max = {int(en_.min())}/{int(en_.median())}/{int(en_.max())}  argmax_end = {int(en_.argmax())}")
