#!/usr/bin/env python
"""bench.py -- QSpec Llama-3-8B speculative decode on the B200-native dual-mode int4 path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  (N>1: python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...)

A "step" is one QSpec speculative step of BASELINE.json configs[1] (Llama-3-8B architecture,
random-init int4 weights, 4 running sequences, k=3): 3 draft passes in W4A4 + 1 verify pass in
W4A16 over the same packed weights and KV cache, greedy acceptance. Prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "qspec_llama3_8b_decode_tokens_per_s"
UNIT = "tokens/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seqs", type=int, default=4, help="sequences per GPU (weak scaling: the job decodes seqs x N sequences at TP=N)")
    ap.add_argument("--strong", action="store_true", help="keep --seqs sequences in total at every N (strong scaling)")
    ap.add_argument("--k", type=int, default=3)
    ap.add_argument("--prompt-len", type=int, default=128)
    ap.add_argument("--model", default="8b", choices=["8b", "70b", "tiny"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_reference_line(args, impl_line: bool):
    """The reference's torch fake-quant / dequant path on the host cores (oracle/cpu_baseline.py)."""
    from oracle.cpu_baseline import time_spec_step_sample
    layers = {"8b": 32, "70b": 80, "tiny": 2}[args.model]
    cfg = None
    if args.model == "70b":
        cfg = dict(hidden=8192, inter=28672, heads=64, kv_heads=8)
    elif args.model == "tiny":
        cfg = dict(hidden=512, inter=28 * 64, heads=4, kv_heads=2)
    reps = max(1, min(args.steps, 2)) if impl_line else 1
    r = time_spec_step_sample(B=args.total_seqs, k=args.k, layers_total=layers, reps=reps, cfg=cfg)
    # tokens per step are not defined for the CPU sample (no lm_head); use 1 token / sequence / step,
    # the floor the GPU arm is also held to when acceptance is 0.
    tps = args.total_seqs / r["seconds_per_step"]
    return r, tps


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # N = 1 is BASELINE configs[1] (max_num_seqs=4). N > 1: tensor parallel over N GPUs with the same 4 sequences PER GPU
    # (per-GPU GEMM work fixed: N x the tokens on 1/N of every weight matrix) => weak scaling; --strong keeps 4 in total.
    args.total_seqs = args.seqs * (1 if args.strong else world)
    scaling = "strong" if args.strong or world == 1 else "weak"
    workload = (f"Llama-3-{args.model.upper()} QSpec decode, draft k={args.k}, max_num_seqs={args.total_seqs}, "
                f"random-init int4 weights, TP={world}")
    config = {"workload": workload, "seqs": args.total_seqs, "seqs_per_gpu": args.total_seqs // world, "k": args.k,
              "prompt_len": args.prompt_len, "parallelism": f"tp{world}",
              "l2": "weights streamed per pass (>= 4.5 GB in total) exceed the 126 MB L2"}

    if args.impl == "reference":
        if rank != 0:
            return
        r, tps = cpu_reference_line(args, True)
        line = {"impl": "reference", "metric": METRIC, "value": tps, "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["seconds_per_step"] * 1e3,
                "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "i8 (int4 x int4 -> int32, draft) / f16 (int4 -> f16 x f16 -> f32, verify)",
                "data": "synthetic", "config": config,
                "cpu_baseline": {"value": tps, "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
                "e2e": {"value": tps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return

    import torch
    import torch.distributed as dist
    from qspec_b200 import ops
    from qspec_b200.engine import QSpecEngine
    from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    tp_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        tp_group = dist.group.WORLD

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B, k = args.total_seqs, args.k
    max_seq = args.prompt_len + (args.steps * 2 + args.warmup + 8) * (k + 1) + 8
    mk = {"8b": QSpecLlamaConfig.llama3_8b, "70b": QSpecLlamaConfig.llama3_70b, "tiny": QSpecLlamaConfig.tiny}[args.model]
    cfg = mk(max_seq_len=max_seq)
    model = QSpecLlamaModel(cfg, max_batch=B, max_tokens=max(64, B * (k + 1)), device=dev, tp_group=tp_group)
    eng = QSpecEngine(model, B, k, use_cuda_graph=not args.no_graph)
    g = torch.Generator().manual_seed(0)
    prompts = torch.randint(0, cfg.vocab_size, (B, args.prompt_len), generator=g).to(dev)
    eng.prefill(prompts)
    l0 = ops.launch_count()
    eng.capture()
    launches_per_step = (ops.launch_count() - l0) // 2 if not args.no_graph else None   # capture() = 1 eager warm step + 1 captured step

    # ---- device-resident timed region ---------------------------------------------------------
    for _ in range(max(3, args.warmup)):
        eng.step()
    barrier()
    if args.no_graph:
        l0 = ops.launch_count(); eng.step(); eng.stream.synchronize(); launches_per_step = ops.launch_count() - l0
    c0 = eng.counters.clone()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.nvtx.range_push("timed")             # ncu --nvtx --nvtx-include "timed/" selects this region
    e0.record(eng.stream)
    for _ in range(args.steps):
        eng.step()
    e1.record(eng.stream)
    e1.synchronize()
    torch.cuda.nvtx.range_pop()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    emitted = int((eng.counters - c0)[2])
    accepted, drafted = int((eng.counters - c0)[0]), int((eng.counters - c0)[1])
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    value = emitted / (ms / 1e3)

    # ---- end-to-end through the public API with HOST buffers ------------------------------------
    host_tok = eng.cur_tok.cpu()
    for _ in range(2):
        out = eng.step_host(host_tok)
        host_tok = out[:, : k + 1].gather(1, (out[:, k + 1:] - 1)).squeeze(1).clone()
    barrier()
    t0 = time.perf_counter()
    e2e_tokens = 0
    for _ in range(args.steps):
        out = eng.step_host(host_tok)
        cnt = out[:, k + 1]
        e2e_tokens += int(cnt.sum())
        host_tok = out[:, : k + 1].gather(1, (cnt - 1).unsqueeze(1)).squeeze(1).clone()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t)
    e2e_value = e2e_tokens / e2e_s

    # ---- roofline of the dominant kernel + per-layer metric (W4A4 TOPS, W4A16 HBM GB/s) ---------
    roofline, per_layer = None, None
    if rank == 0:
        peak, peak_src = peaks()
        per_layer = {}
        sh = cfg.linear_shapes(world)

        def time_layers(fn_of_layer):
            """us per launch: the 32 layers' distinct weights back to back (> L2 for every shape),
            replayed from a CUDA graph so the number is device time, not Python launch overhead."""
            s = eng.stream
            with torch.cuda.stream(s):
                for l in model.layers[:2]:
                    fn_of_layer(l)
                s.synchronize()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=s):
                    for l in model.layers:
                        fn_of_layer(l)
                gr.replay()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.synchronize()
                a.record(s)
                for _ in range(5):
                    gr.replay()
                b.record(s)
                b.synchronize()
            return a.elapsed_time(b) * 1e3 / (5 * len(model.layers))

        Md, Mv = B, B * (k + 1)
        qd = torch.randint(-128, 127, (Md, max(cfg.hidden_size, model.inter) // 2), dtype=torch.int8, device=dev)
        sd = (torch.rand(Md, device=dev) * 0.1).half()
        xv = torch.randn(Mv, max(cfg.hidden_size, model.inter), device=dev).half()
        for name, attr in (("qkv", "qkv_proj"), ("o", "o_proj"), ("gate_up", "gate_up"), ("down", "down_proj")):
            N, K = sh[name]
            o4 = torch.empty(Md, N, dtype=torch.float16, device=dev)
            o16 = torch.empty(Mv, N, dtype=torch.float16, device=dev)
            qa = qd[:, : K // 2].contiguous()
            xa = xv[:, :K].contiguous()
            us4 = time_layers(lambda l: ops.rowwise_scaled_linear_cutlass_s4s4_unified(
                qa, sd, getattr(l, attr).weight, getattr(l, attr).weight_scales.view(-1), None, o4))
            us16 = time_layers(lambda l: ops.gemm_w4a16(xa, getattr(l, attr).weight,
                                                        getattr(l, attr).weight_scales.view(-1), None, o16))
            b4 = N * K // 2 + 2 * N + Md * K // 2 + 2 * Md + 2 * Md * N
            b16 = N * K // 2 + 2 * N + 2 * Mv * K + 2 * Mv * N
            per_layer[name] = {"N": N, "K": K,
                               "w4a4": {"M": Md, "us": round(us4, 2), "TOPS": round(2.0 * Md * N * K / us4 / 1e6, 2),
                                        "hbm_GBps": round(b4 / us4 / 1e3, 1), "frac_hbm_measured": round(b4 / us4 / 1e3 / peak, 3),
                                        "frac_8TBps": round(b4 / us4 / 1e3 / 8000.0, 3)},
                               "w4a16": {"M": Mv, "us": round(us16, 2), "hbm_GBps": round(b16 / us16 / 1e3, 1),
                                         "frac_hbm_measured": round(b16 / us16 / 1e3 / peak, 3),
                                         "frac_8TBps": round(b16 / us16 / 1e3 / 8000.0, 3)}}
        dom = per_layer["gate_up"]["w4a4"]
        N, K = sh["gate_up"]
        alg = N * K // 2 + 2 * N + Md * K // 2 + 2 * Md + 2 * Md * N
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp) and world == 1:
            traffic = json.load(open(tp)).get("gemm_w4a4_gate_up_dram_bytes")
        roofline = {"bound": "hbm", "kernel": "w4_stream_kernel<MODE_I8> = qspec_gemm_w4a4 (gate_up, draft mode: 3 of every 4 passes, largest shape)",
                    "achieved": dom["hbm_GBps"], "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                    "frac": round(dom["hbm_GBps"] / peak, 4), "algorithmic_bytes_per_launch": alg,
                    "us_per_launch": dom["us"], "traffic": traffic}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r, tps = cpu_reference_line(args, False)
        cpu_baseline = {"value": tps, "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": scaling, "vs_baseline": None, "dtype": "i8 (int4 x int4 -> int32, draft) / f16 (int4 -> f16 x f16 -> f32, verify)",
                "data": "synthetic", "config": config,
                "acceptance": {"draft_acceptance_rate": accepted / drafted if drafted else None,
                               "emitted_tokens": emitted, "tokens_per_step_per_seq": emitted / (args.steps * B)},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * 8, "d2h_bytes_per_step": B * (k + 2) * 8},
                "gpu_launches": (launches_per_step or 0) * args.steps, "launches_per_step": launches_per_step,
                "clocks": clocks, "roofline": roofline, "per_layer": per_layer, "cpu_baseline": cpu_baseline}
        print(json.dumps(line), flush=True)
    if world > 1:
        # Leave without tearing NCCL down: destroy_process_group() can block for minutes while captured CUDA graphs still
        # hold the communicator (seen at TP=2), and the line above is already out.
        sys.stdout.flush()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        os._exit(0)


if __name__ == "__main__":
    main()
