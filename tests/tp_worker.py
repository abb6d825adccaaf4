"""2-rank worker of tests/test_gpu_tp.py (run under torch.distributed.run, one rank per GPU).

Checks the all-reduce-free tensor-parallel path (qspec_tp_add_norm_*: peer partial sums read over NVLink inside the
norm, arrival flags between the ranks) three ways:
  1. against the same model with one NCCL all-reduce per row-parallel GEMM (RowParallelLinear.forward,
     vllm/model_executor/layers/linear.py:1143-1144): hidden states of both modes within the stated rounding bound;
  2. against the TP-aware CPU oracle (oracle/qspec_model_oracle.py, `ranks=`): hidden states within north_star's 1e-2
     relative, every rank holding the same bits;
  3. the engine's tokens with the step under a CUDA graph: fused path == NCCL path (asserted), and == the oracle-backed
     speculative loop unless the oracle's own tokens change under rounding-level jitter of attention / SiLU (the arg max
     was decided by rounding; see OracleModel.jitter).

Rounding bound of (1): per all-reduce both paths round the exact sum of two f16 partials to f16 once (NCCL adds two
halves; the fused kernel adds in fp32, exact for |a|/|b| < 2^13, then rounds), so they agree except for a double
rounding when the partials differ by more than 2^13 in magnitude (~1e-7 per element); such a 1-ulp difference then
travels through the remaining layers like any other rounding (it can flip an int4 rounding of the next quantiser).
Asserted: at least 90 % of the passes are bit-identical, and a pass that is not stays within north_star's 1e-2
relative (max|d| / max|ref|) -- the tolerance between two valid implementations of the layer.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def f16_ulp(x: torch.Tensor) -> torch.Tensor:
    """Spacing of float16 at |x| (normal range; 2^-24 below it)."""
    e = torch.floor(torch.log2(x.abs().float().clamp_min(2.0 ** -14)))
    return torch.exp2(e - 10)


def main():
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from qspec_b200.engine import QSpecEngine
    from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel
    cfg = QSpecLlamaConfig.tiny(intermediate_size=28 * 64 * world, num_attention_heads=2 * world,
                                num_key_value_heads=world, hidden_size=256 * world, num_hidden_layers=3)
    B, k = 2, 3

    def build(fused):
        os.environ["QSPEC_TP_FUSED"] = "1" if fused else "0"
        m = QSpecLlamaModel(cfg, max_batch=B, max_tokens=16, device=dev, tp_group=dist.group.WORLD, seed=4)
        assert (m.peers is not None) == fused
        return m

    mf, mn = build(True), build(False)
    if rank == 0:
        print("peer mapping:", mf.peers.how, flush=True)
    # rank 0 also holds every rank's shard (same seeds => same weights) for the TP oracle
    om = None
    if rank == 0:
        from oracle.qspec_model_oracle import OracleModel
        local = [mf] + [QSpecLlamaModel(cfg, max_batch=B, max_tokens=16, device=dev, seed=4, tp_size=world, tp_rank=r, peers=object())
                        for r in range(1, world)]
        om = OracleModel.from_gpu_shards(local, B)
        del local
    g = torch.Generator().manual_seed(21)
    T = 10
    ids = torch.randint(0, cfg.vocab_size, (T,), generator=g).to(dev)
    pos = torch.arange(T, dtype=torch.int32, device=dev) % 5
    seq = (torch.arange(T, dtype=torch.int32, device=dev) // 5)
    worst_ulps, worst_frac, worst_rel, identical, passes = 0.0, 0.0, 0.0, 0, 0
    for it in range(40):                      # many passes: epochs advance, both partial buffers are reused
        for w4a4 in (False, True):
            hf = mf.forward(ids, pos, seq, w4a4).clone()
            hn = mn.forward(ids, pos, seq, w4a4).clone()
            assert torch.isfinite(hf).all()
            d = (hf.float() - hn.float()).abs()
            ulps = (d / f16_ulp(torch.maximum(hf.abs(), hn.abs()))).max().item()
            frac = (d > 0).float().mean().item()
            worst_ulps, worst_frac = max(worst_ulps, ulps), max(worst_frac, frac)
            passes += 1
            identical += int(frac == 0.0)
            rel_fn = (d.max() / hn.float().abs().max()).item()
            assert rel_fn <= 1e-2, f"fused vs NCCL: {rel_fn:.2e} relative, {ulps:.1f} ulps, {frac:.2e} of the elements (w4a4={w4a4}, it={it})"
            if om is not None and it == 0:
                from test_gpu_tp_onegpu import assert_hidden_close

                def oracle_forward(jitter, w4a4=w4a4):
                    om.jitter, om._jitter_calls = jitter, 0
                    return om.forward(ids.cpu().numpy(), pos.cpu().numpy(), seq.cpu().numpy(), w4a4)

                rel = assert_hidden_close(hf.float().cpu().numpy(), oracle_forward, f"TP={world} w4a4={w4a4}")
                om.jitter = None
                worst_rel = max(worst_rel, rel)
                assert w4a4 or rel <= 1e-2, f"TP={world} verify mode: hidden state {rel:.3e} relative from the TP oracle"
    torch.cuda.synchronize()
    assert mf.peers.timeouts() == 0
    assert identical >= 0.9 * passes, f"only {identical} of {passes} passes are bit-identical between the fused and the NCCL reduction"
    # every rank must hold the same hidden state
    ref = hf.clone()
    dist.broadcast(ref, 0)
    assert torch.equal(ref, hf), "ranks disagree on the reduced hidden state"

    toks, prompts = [], torch.randint(0, cfg.vocab_size, (B, 9), generator=torch.Generator().manual_seed(3))
    for m in (mf, mn):
        eng = QSpecEngine(m, B, k, use_cuda_graph=True)
        eng.prefill(prompts.to(dev))
        first = eng.cur_tok.cpu().clone()
        out = []
        for _ in range(8):
            eng.step()
            torch.cuda.synchronize()
            out.append(eng.d_out_pack.cpu().clone())
        eng.metrics()                         # raises on a peer time-out
        toks.append(torch.stack(out))
    assert torch.equal(toks[0], toks[1]), "fused-reduction tokens differ from the NCCL all-reduce tokens"
    assert mf.peers.timeouts() == 0
    oracle_note = ""
    if om is not None:
        from oracle.qspec_model_oracle import spec_step
        P = prompts.shape[1]

        def oracle_run(jitter=None):
            for c in om.ck + om.cv:
                c[:] = 0
            om.jitter, om._jitter_calls = jitter, 0
            last = [om.forward(prompts[b].numpy(), np.arange(P), np.full(P, b), False)[-1] for b in range(B)]
            cur, _ = om.greedy(np.stack(last))
            first, seq_len, outs = cur.copy(), np.full(B, P), []
            for s in range(8):
                emitted, n_acc, cur, seq_len, _ = spec_step(om, cur, seq_len, k)
                outs.append([[int(t) for t in e] for e in emitted])
            return first, outs

        base = oracle_run()
        ok = np.array_equal(base[0], first.numpy())
        for s in range(8):
            for b in range(B):
                cnt = int(toks[0][s, b, k + 1])
                ok = ok and toks[0][s, b, :cnt].tolist() == base[1][s][b]
        ill = False
        if not ok:
            for js in (1, 2, 3):
                j = oracle_run((js, 1e-3))
                if not np.array_equal(j[0], base[0]) or j[1] != base[1]:
                    ill = True
                    break
        assert ok or ill, f"TP={world} tokens differ from the TP oracle, whose own tokens are stable under rounding-level jitter"
        oracle_note = f" tokens_equal_oracle={ok} decided_by_rounding={ill}"
    dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"TP_OK world={world} identical_passes={identical}/{passes} worst_fused_vs_nccl_ulps={worst_ulps:.2f} differing_frac={worst_frac:.2e} "
              f"worst_rel_vs_oracle={worst_rel:.2e} tokens_equal=True{oracle_note}", flush=True)
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
