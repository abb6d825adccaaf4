"""In-loop kernel timeline of one decode step (CUPTI activity records through torch.profiler; analysis only -- numbers
taken under a profiler are never bench values). For every kernel type: launches per step, mean duration and the mean
idle gap before it (previous kernel's end -> this kernel's start), i.e. where a step's time goes between kernels.

python scripts/gpu_timeline.py [--gpus-model 8b] > gpurun_out/timeline.md
"""
import argparse, collections, json, os, re, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from qspec_b200.engine import QSpecEngine
from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="8b")
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--prompt-len", type=int, default=128)
ap.add_argument("--seqs", type=int, default=0, help="sequences in total (default: 4 per rank)")
args = ap.parse_args()
import torch.distributed as dist
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
tp_group = None
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
    tp_group = dist.group.WORLD
B, k = (args.seqs or 4 * world), 3
mk = {"8b": QSpecLlamaConfig.llama3_8b, "tiny": QSpecLlamaConfig.tiny, "qwen14b": QSpecLlamaConfig.qwen25_14b}[args.model]
cfg = mk(max_seq_len=512)
model = QSpecLlamaModel(cfg, max_batch=B, max_tokens=max(64, B * (k + 1)), device=dev, tp_group=tp_group)
eng = QSpecEngine(model, B, k, use_cuda_graph=True)
prompts = torch.randint(0, cfg.vocab_size, (B, args.prompt_len), generator=torch.Generator().manual_seed(0)).to(dev)
eng.prefill(prompts)
eng.capture()
for _ in range(5):
    eng.step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(args.steps):
        eng.step()
    torch.cuda.synchronize()
if world > 1:
    dist.barrier(); torch.cuda.synchronize()
    if local != 0:
        os._exit(0)
path = os.path.join(tempfile.gettempdir(), "qspec_trace.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "ts" in e]
ev.sort(key=lambda e: e["ts"])
if not ev:
    print("no kernel records"); sys.exit(0)
# the last full step: split at the largest gaps? simply use all records of the profiled steps
def short(n):
    n = re.sub(r"\(.*", "", n)
    n = re.sub(r"^void ", "", n)
    return n[:64]
agg = collections.OrderedDict()
prev_end = None
total_busy = 0.0
for e in ev:
    name = short(e["name"])
    g = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
    end = e["ts"] + e["dur"]
    g[0] += 1
    g[1] += e["dur"]
    total_busy += e["dur"]
    if prev_end is not None:
        g[2] += max(0.0, e["ts"] - prev_end)           # idle gap before (0 when the PDL prologue overlaps)
        g[3] += max(0.0, end - prev_end)               # critical-path increment: how much later the chain ends because of it
    prev_end = max(prev_end or 0, end)
span = ev[-1]["ts"] + ev[-1]["dur"] - ev[0]["ts"]
print(f"# in-loop timeline, {args.steps} steps, {len(ev)} device records, span {span / args.steps / 1e3:.3f} ms/step, "
      f"sum of kernel durations {total_busy / args.steps / 1e3:.3f} ms/step (a PDL dependent starts early and waits: durations overlap)\n")
print("| kernel | per step | mean duration us (incl. early start) | mean idle gap before us | mean increment of the chain us | share of step % |")
print("|---|---|---|---|---|---|")
for name, (n, d, gap, inc) in sorted(agg.items(), key=lambda kv: -kv[1][3]):
    print(f"| {name} | {n / args.steps:.0f} | {d / n:.2f} | {gap / n:.2f} | {inc / n:.2f} | {100 * inc / span:.1f} |")
# two layers in the middle of a draft pass and of the verify pass, in order
def window(title, i0, n=17):
    print(f"\n## {title} (name, start us, end us, end - previous end)")
    t0 = ev[i0]["ts"]
    pe = None
    for e in ev[i0:i0 + n]:
        end = e["ts"] + e["dur"]
        print(f"{short(e['name']):56s} {e['ts'] - t0:8.2f} {end - t0:8.2f} {'' if pe is None else f'{end - pe:7.2f}'}")
        pe = end if pe is None else max(pe, end)
per_step = len(ev) // args.steps
base = (args.steps - 1) * per_step
window("draft pass, mid", base + per_step // 8)
window("verify pass, mid", base + per_step * 7 // 8 - 20)

if world > 1:
    sys.stdout.flush()
    os._exit(0)
