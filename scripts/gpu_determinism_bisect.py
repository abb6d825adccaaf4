"""Which kernel makes two identical forward passes differ? Per-op repeats on identical inputs, then whole passes of a 4-layer
Llama-3-8B-shaped model under feature toggles. Analysis script (prints one JSON line per case)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qspec_b200 import ops
from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel

dev = torch.device("cuda:0")
g = torch.Generator(device="cuda").manual_seed(1)


def same(fn, outs, reps=6):
    """fn() writes `outs`; returns how many of `reps` repeats differ from the first."""
    fn(); torch.cuda.synchronize()
    ref = [o.clone() for o in outs]
    bad = 0
    for _ in range(reps):
        for o in outs:
            o.zero_()
        fn(); torch.cuda.synchronize()
        if any(not torch.equal(o.view(torch.uint8), r.view(torch.uint8)) for o, r in zip(outs, ref)):
            bad += 1
    return bad


# ---- per-op repeats
for name, (N, K) in {"qkv": (6144, 4096), "o": (4096, 4096), "gate_up": (28672, 4096), "down": (4096, 14336)}.items():
    w = torch.randint(-128, 127, (N, K // 2), dtype=torch.int8, device=dev)
    sW = (torch.rand(N, device=dev) * 0.01).half()
    for M in (8, 64, 128, 256):
        x = torch.randn(M, K, device=dev).half()
        out = torch.empty(M, N, dtype=torch.float16, device=dev)
        bad16 = same(lambda: ops.gemm_w4a16(x, w, sW, None, out), [out])
        a = torch.randint(-128, 127, (M, K // 2), dtype=torch.int8, device=dev)
        sA = (torch.rand(M, device=dev) * 0.1).half()
        bad4 = same(lambda: ops.rowwise_scaled_linear_cutlass_s4s4_unified(a, sA, w, sW, None, out), [out])
        rec = {"op": "gemm", "shape": name, "M": M, "w4a16_differs": bad16, "w4a4_differs": bad4}
        if name in ("o", "down") and ops.gemm_split_factor(True, M, N, K) in range(1, 9):
            part = torch.empty(8 * M * N, dtype=torch.int32, device=dev)
            x0 = torch.randn(M, N, device=dev).half()
            xo, nrm = torch.empty_like(x0), torch.empty_like(x0)

            def split16():
                xo.copy_(x0)
                sp = ops.gemm_w4a16_split(x, w, sW, None, part)
                ops.split_add_rms_norm_fuse_sum_fp16(nrm, xo, sp, 1e-5)
            rec["w4a16_split_differs"] = same(split16, [xo, nrm])
        print(json.dumps(rec), flush=True)

# ---- whole passes under toggles
TOGGLES = [{}, {"QSPEC_GEMM_PAIR": "0"}, {"QSPEC_SPLIT_FUSED": "0"}, {"QSPEC_SEQ_ATTN": "0"}, {"QSPEC_GEMM_NP256": "0"},
           {"QSPEC_GEMM_PAIR": "0", "QSPEC_SPLIT_FUSED": "0", "QSPEC_SEQ_ATTN": "0"}]
B, P = 64, 8
for tog in TOGGLES:
    for k_, v_ in tog.items():
        os.environ[k_] = v_
    cfg = QSpecLlamaConfig.llama3_8b(num_hidden_layers=4, max_seq_len=256)
    m = QSpecLlamaModel(cfg, max_batch=B, max_tokens=256, device=dev)
    m.cache_k.zero_(); m.cache_v.zero_()
    cg = torch.Generator().manual_seed(3)
    for b in range(B):
        ids = torch.randint(0, cfg.vocab_size, (P,), generator=cg).to(dev)
        m.forward(ids, torch.arange(P, dtype=torch.int32, device=dev), torch.full((P,), b, dtype=torch.int32, device=dev), w4a4=False)
    cases = {
        "prefill T=128 one seq (verify)": (torch.randint(0, cfg.vocab_size, (128,), generator=cg).to(dev), torch.arange(128, dtype=torch.int32, device=dev),
                                           torch.zeros(128, dtype=torch.int32, device=dev), False),
        "draft T=64": (torch.randint(0, cfg.vocab_size, (64,), generator=cg).to(dev), torch.full((64,), 128, dtype=torch.int32, device=dev),
                       torch.arange(64, dtype=torch.int32, device=dev), True),
        "verify T=256": (torch.randint(0, cfg.vocab_size, (256,), generator=cg).to(dev),
                         (128 + torch.arange(4, dtype=torch.int32, device=dev)).repeat(64).contiguous(),
                         torch.arange(64, dtype=torch.int32, device=dev).repeat_interleave(4).contiguous(), False),
    }
    # (sequence 0 holds 128 positions after the first case; the others 8: positions 128.. of sequences 1..63 read zeros in between -- same every time)
    for cname, (ids, pos, seq, w4a4) in cases.items():
        ref, bad, maxd = None, 0, 0.0
        for rep in range(5):
            h = m.forward(ids, pos, seq, w4a4=w4a4).clone()
            torch.cuda.synchronize()
            if ref is None:
                ref = h
            elif not torch.equal(h, ref):
                bad += 1
                maxd = max(maxd, float((h.float() - ref.float()).abs().max()))
        print(json.dumps({"pass": cname, "toggles": tog, "repeats_differing": bad, "max_abs_diff": maxd, "finite": bool(torch.isfinite(ref.float()).all())}), flush=True)
    for k_ in tog:
        os.environ.pop(k_, None)
    del m
    torch.cuda.empty_cache()
