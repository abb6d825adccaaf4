"""2-rank worker of tests/test_gpu_tp.py (run under torch.distributed.run, one rank per GPU).

Checks the all-reduce-free tensor-parallel path (qspec_tp_add_norm_*: peer partial sums read over NVLink inside the
norm) against the same model with one NCCL all-reduce per row-parallel GEMM (RowParallelLinear.forward,
layers/linear.py:1143-1144): hidden states of both modes, and the engine's tokens with the step under a CUDA graph.
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank = int(os.environ["RANK"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from qspec_b200.engine import QSpecEngine
    from qspec_b200.model import QSpecLlamaConfig, QSpecLlamaModel
    cfg = QSpecLlamaConfig.tiny(intermediate_size=28 * 128, num_hidden_layers=3)
    B, k = 2, 3

    def build(fused):
        os.environ["QSPEC_TP_FUSED"] = "1" if fused else "0"
        m = QSpecLlamaModel(cfg, max_batch=B, max_tokens=16, device=dev, tp_group=dist.group.WORLD, seed=4)
        assert (m.peers is not None) == fused
        return m

    mf, mn = build(True), build(False)
    if rank == 0:
        print("peer mapping:", mf.peers.how, flush=True)
    g = torch.Generator().manual_seed(21)
    T = 10
    ids = torch.randint(0, cfg.vocab_size, (T,), generator=g).to(dev)
    pos = torch.arange(T, dtype=torch.int32, device=dev) % 5
    seq = (torch.arange(T, dtype=torch.int32, device=dev) // 5)
    worst = 0.0
    for it in range(40):                      # many passes: epochs advance, both partial buffers are reused
        for w4a4 in (False, True):
            hf = mf.forward(ids, pos, seq, w4a4).clone()
            hn = mn.forward(ids, pos, seq, w4a4).clone()
            d = (hf.float() - hn.float()).abs().max().item()
            worst = max(worst, d)
            assert torch.isfinite(hf).all()
            # two ranks: fp32 sum + one rounding == NCCL's f16 sum up to rare double roundings
            assert d <= 2e-2, f"fused vs NCCL hidden differ by {d} (w4a4={w4a4}, it={it})"
    torch.cuda.synchronize()
    assert mf.peers.timeouts() == 0
    # every rank must hold the same hidden state
    ref = hf.clone()
    dist.broadcast(ref, 0)
    assert torch.equal(ref, hf), "ranks disagree on the reduced hidden state"

    toks = []
    for m in (mf, mn):
        eng = QSpecEngine(m, B, k, use_cuda_graph=True)
        prompts = torch.randint(0, cfg.vocab_size, (B, 9), generator=torch.Generator().manual_seed(3)).to(dev)
        eng.prefill(prompts)
        out = []
        for _ in range(8):
            eng.step()
            torch.cuda.synchronize()
            out.append(eng.d_out_pack.cpu().clone())
        toks.append(torch.stack(out))
    same = torch.equal(toks[0], toks[1])
    assert mf.peers.timeouts() == 0
    dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"TP_OK worst_hidden_diff={worst:.3e} tokens_equal={same}", flush=True)
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
