"""Run-to-run determinism of the speculative step at BASELINE configs[2] sizes (64 sequences, Llama-3-8B, k = 3): the same
prompts decoded twice from the same state must emit the same tokens, under the CUDA graph and eagerly (different overlap of the
programmatic dependent launches). A difference would mean a race between kernels, not a numerics question.
python scripts/gpu_determinism.py [--layers 32] [--seqs 64] [--steps 12] [--shifts none,19,20]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qspec_b200.engine import QSpecEngine
from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=32)
ap.add_argument("--seqs", type=int, default=64)
ap.add_argument("--steps", type=int, default=12)
ap.add_argument("--shifts", default="none,19,20")
a = ap.parse_args()
dev = torch.device("cuda:0")
B, k = a.seqs, 3
cfg = QSpecLlamaConfig.llama3_8b(num_hidden_layers=a.layers, max_seq_len=512)
model = QSpecLlamaModel(cfg, max_batch=B, max_tokens=max(64, B * (k + 1)), device=dev)
prompts = torch.randint(0, cfg.vocab_size, (B, 128), generator=torch.Generator().manual_seed(0)).to(dev)


def run(graph):
    eng = QSpecEngine(model, B, k, use_cuda_graph=graph)
    eng.prefill(prompts)
    first = eng.cur_tok.clone()
    eng.capture()
    outs = []
    for _ in range(a.steps):
        eng.step()
        eng.stream.synchronize(); torch.cuda.synchronize()
        outs.append(eng.d_out_pack.clone())
    m = eng.metrics()
    return first, torch.stack(outs), m


for sh in a.shifts.split(","):
    model.make_modes_agree(None if sh == "none" else int(sh))
    runs = [run(True), run(True), run(False), run(False)]
    ref = runs[0]
    rep = {"shift": sh, "acceptance": ref[2].get("draft_acceptance_rate"), "runs": []}
    for name, r in zip(("graph#2", "eager#1", "eager#2"), runs[1:]):
        same_first = bool(torch.equal(r[0], ref[0]))
        diff = (r[1] != ref[1])
        first_step = int(diff.flatten(1).any(1).nonzero()[0]) if diff.any() else None
        rep["runs"].append({"run": name, "prefill_token_equal": same_first, "steps_equal": not bool(diff.any()),
                            "first_differing_step": first_step, "differing_seqs_at_that_step": int(diff[first_step].any(1).sum()) if first_step is not None else 0,
                            "acceptance": r[2].get("draft_acceptance_rate")})
    print(json.dumps(rep), flush=True)
