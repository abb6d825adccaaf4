"""Quick 8B engine timing (random weights): ms per speculative step, eager vs graph."""
import sys, os, time, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qspec_b200.engine import QSpecEngine
from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel
from qspec_b200 import ops

B, k = 4, 3
cfg = QSpecLlamaConfig.llama3_8b(max_seq_len=1024)
t0 = time.time()
model = QSpecLlamaModel(cfg, max_batch=B, max_tokens=64, device="cuda")
torch.cuda.synchronize()
print("model built in %.1fs, mem %.2f GB" % (time.time() - t0, torch.cuda.memory_allocated() / 2**30), flush=True)
eng = QSpecEngine(model, B, k, use_cuda_graph=True)
prompts = torch.randint(0, cfg.vocab_size, (B, 128), device="cuda")
t0 = time.time(); eng.prefill(prompts); torch.cuda.synchronize(); print("prefill %.2fs" % (time.time() - t0), flush=True)
l0 = ops.launch_count()
eng.capture()
print("launches per captured step (x2: warm + capture):", ops.launch_count() - l0, flush=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(3):
    eng.step()
eng.stream.synchronize()
e0.record(eng.stream)
N = 20
for _ in range(N):
    eng.step()
e1.record(eng.stream)
e1.synchronize()
ms = e0.elapsed_time(e1) / N
m = eng.metrics()
print(json.dumps({"ms_per_step": ms, "metrics": m, "tok_per_s": m["emitted_tokens"] / (23 * ms / 1e3) if ms else None}))
