"""Speculative-step time of the Qwen2.5-14B variant of the path (random-init int4 weights, TP = 1): the same
engine and CUDA-graphed step as bench.py runs for Llama-3-8B, timed with CUDA events on the engine's stream. Not a BASELINE config:
a supplementary line for profiles/.  python scripts/gpu_qwen_step.py [--seqs 64,4] [--steps 20]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qspec_b200 import ops
from qspec_b200.engine import QSpecEngine
from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel

ap = argparse.ArgumentParser()
ap.add_argument("--seqs", default="64,4")
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--layers", type=int, default=48)
a = ap.parse_args()
dev = torch.device("cuda:0")
k, P = 3, 128
seqs = [int(s) for s in a.seqs.split(",")]
B0 = max(seqs)
cfg = QSpecLlamaConfig.qwen25_14b(num_hidden_layers=a.layers, max_seq_len=P + (a.warmup + a.steps + 6) * (k + 1) + 8)
model = QSpecLlamaModel(cfg, max_batch=B0, max_tokens=max(64, B0 * (k + 1)), device=dev)
for B in seqs:
    eng = QSpecEngine(model, B, k, use_cuda_graph=True)
    prompts = torch.randint(0, cfg.vocab_size, (B, P), generator=torch.Generator().manual_seed(0)).to(dev)
    eng.prefill(prompts)
    l0 = ops.launch_count()
    eng.capture()
    lps = (ops.launch_count() - l0) // 2
    for _ in range(a.warmup):
        eng.step()
    eng.stream.synchronize()
    c0 = eng.counters.clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(eng.stream)
    for _ in range(a.steps):
        eng.step()
    e1.record(eng.stream)
    e1.synchronize()
    ms = e0.elapsed_time(e1)
    d = (eng.counters - c0).tolist()
    print(json.dumps({"model": cfg.name, "layers": a.layers, "seqs": B, "k": k, "prompt_len": P, "ms_per_step": round(ms / a.steps, 3),
                      "tokens_per_s_at_acceptance_floor": round(d[2] / (ms / 1e3), 1), "draft_acceptance_rate": d[0] / max(1, d[1]),
                      "launches_per_step": lps, "weights_gb": round(sum(p.numel() for p in [l.qkv_proj.weight for l in model.layers]) / 1e9, 2)}), flush=True)
