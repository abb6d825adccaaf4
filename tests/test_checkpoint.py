"""Checkpoint-format loader: key remapping, RTN quantiser, qkv / gate_up fusion order, TP shards, safetensors
round trip. Host logic on CPU; the GPU test loads a saved checkpoint into a second model and compares the speculative decode."""
import types

import numpy as np
import pytest
import torch

from oracle import qspec_oracle as O
from qspec_b200 import checkpoint as C
from qspec_b200.model import QSpecDecoderLayer, QSpecLlamaConfig


def _stub_model(cfg, seed=0, tp=1, rank=0):
    """CPU stand-in with the attributes the loader touches (the real model allocates CUDA scratch)."""
    g = torch.Generator().manual_seed(seed + 7)
    layers = [QSpecDecoderLayer(cfg, i, "cpu", tp, rank, seed) for i in range(cfg.num_hidden_layers)]
    embed = (torch.randn(cfg.vocab_size, cfg.hidden_size, generator=g) * 0.02).half()
    head = (torch.randn(cfg.vocab_size // tp, cfg.hidden_size, generator=g) * 0.02).half()
    return types.SimpleNamespace(cfg=cfg, tp_size=tp, tp_rank=rank, device=torch.device("cpu"), layers=layers, embed=embed, lm_head=head)


def test_key_remapping_follows_the_reference_and_inverts():
    ref = {"model.layers.3.self_attn.o_proj.1.weight": 1, "model.layers.3.mlp.down_proj.0.had_rem_dim": 2,
           "model.layers.3.mlp.down_proj.2.weight_scales": 3, "model.layers.3.self_attn.q_proj.weight": 4, "lm_head.weight": 5}
    m = C.remap_reference_keys(ref)
    assert set(m) == {"model.layers.3.self_attn.o_proj.weight", "model.layers.3.mlp.online_hadamard.had_rem_dim",
                      "model.layers.3.mlp.down_proj.weight_scales", "model.layers.3.self_attn.q_proj.weight", "lm_head.weight"}
    assert C.to_reference_keys(m) == ref


def test_rtn_quantiser_matches_formula_and_oracle_packing():
    g = torch.Generator().manual_seed(5)
    w = torch.randn(48, 256, generator=g) * 0.05
    w[0, 0], w[1, 1] = 10.0, -10.0                       # clamp to 7 / -8
    s = C.rtn_scales(w)
    s[0], s[1] = 0.5, 0.5
    packed, scales = C.quantize_weight_rtn(w, s)
    q = np.clip(np.rint(w.numpy() / s.float().numpy().astype(np.float16).astype(np.float32)), -8, 7).astype(np.int8)
    assert packed.dtype == torch.uint8 and tuple(packed.shape) == (48, 128) and tuple(scales.shape) == (48, 1)
    assert np.array_equal(packed.numpy(), O.pack_i4(q))
    assert int(O.unpack_i4(packed.numpy())[0, 0]) == 7 and int(O.unpack_i4(packed.numpy())[1, 1]) == -8
    with pytest.raises(ValueError):
        C.quantize_weight_rtn(w, s[:10])


def test_fusion_order_is_q_k_v_and_up_gate():
    q, k, v = torch.full((4, 2), 1), torch.full((2, 2), 2), torch.full((2, 2), 3)
    w, s = C.fuse_qkv(q, k, v, torch.ones(4, 1), 2 * torch.ones(2, 1), 3 * torch.ones(2, 1))
    assert w[:, 0].tolist() == [1, 1, 1, 1, 2, 2, 3, 3] and s.tolist() == [1, 1, 1, 1, 2, 2, 3, 3]
    gate, up = torch.full((3, 2), 7), torch.full((3, 2), 9)
    w, s = C.fuse_gate_up(gate, up, 7 * torch.ones(3), 9 * torch.ones(3))
    assert w[:, 0].tolist() == [9, 9, 9, 7, 7, 7] and s.tolist() == [9, 9, 9, 7, 7, 7]     # [up ; gate], quarot_llama.py:304-305


def test_export_load_round_trip_and_reference_key_names(tmp_path):
    cfg = QSpecLlamaConfig.tiny()
    a = _stub_model(cfg, seed=3)
    sd = C.export_state_dict(a)
    assert "model.layers.0.self_attn.o_proj.1.weight" in sd and "model.layers.1.mlp.down_proj.2.weight_scales" in sd
    assert sd["model.layers.0.self_attn.q_proj.weight"].dtype == torch.uint8
    assert tuple(sd["model.layers.0.mlp.up_proj.weight_scales"].shape) == (cfg.intermediate_size, 1)
    C.save_qspec_checkpoint(a, str(tmp_path), shards=2)
    assert sorted(p.name for p in tmp_path.iterdir()) == ["model-00001-of-00002.safetensors", "model-00002-of-00002.safetensors"]
    b = _stub_model(cfg, seed=99)
    C.load_qspec_checkpoint(b, str(tmp_path))
    for la, lb in zip(a.layers, b.layers):
        for name in ("qkv_proj", "o_proj", "gate_up", "down_proj"):
            assert torch.equal(getattr(la, name).weight.view(torch.uint8), getattr(lb, name).weight.view(torch.uint8)), name
            assert torch.equal(getattr(la, name).weight_scales.reshape(-1), getattr(lb, name).weight_scales.reshape(-1)), name
    assert torch.equal(a.embed, b.embed) and torch.equal(a.lm_head, b.lm_head)


def test_tp_shards_compute_the_same_layer_as_the_full_weights():
    """Column shards concatenate to the full output; row shards' partial products sum to it (exact in int arithmetic)."""
    cfg = QSpecLlamaConfig.tiny(num_attention_heads=4, num_key_value_heads=2)
    full = _stub_model(cfg, seed=1)
    sd = C.remap_reference_keys(C.export_state_dict(full))
    t1 = C.layer_tensors(sd, 0, cfg, 1, 0)
    shards = [C.layer_tensors(sd, 0, cfg, 2, r, rerotate=False) for r in range(2)]    # synthetic weights: plain byte slices
    D, hq, hkv, I = cfg.head_dim, 4, 2, cfg.intermediate_size
    x = torch.randint(-8, 8, (3, cfg.hidden_size), dtype=torch.int32)
    unp = lambda w: torch.from_numpy(O.unpack_i4(w.view(torch.uint8).numpy()).astype(np.int32))
    y_full = x @ unp(t1["qkv_proj.weight"]).T
    qf, kf, vf = y_full[:, :hq * D], y_full[:, hq * D:(hq + hkv) * D], y_full[:, (hq + hkv) * D:]
    ys = [x @ unp(s["qkv_proj.weight"]).T for s in shards]
    a, b = hq // 2 * D, hkv // 2 * D
    assert torch.equal(torch.cat([y[:, :a] for y in ys], 1), qf)
    assert torch.equal(torch.cat([y[:, a:a + b] for y in ys], 1), kf)
    assert torch.equal(torch.cat([y[:, a + b:] for y in ys], 1), vf)
    g_full = x @ unp(t1["gate_up.weight"]).T
    gs = [x @ unp(s["gate_up.weight"]).T for s in shards]
    n = I // 2
    assert torch.equal(torch.cat([g[:, :n] for g in gs], 1), g_full[:, :I])        # up
    assert torch.equal(torch.cat([g[:, n:] for g in gs], 1), g_full[:, I:])        # gate
    h = torch.randint(-8, 8, (3, I), dtype=torch.int32)
    d_full = h @ unp(t1["down_proj.weight"]).T
    d_parts = sum(h[:, r * n:(r + 1) * n] @ unp(shards[r]["down_proj.weight"]).T for r in range(2))
    assert torch.equal(d_parts, d_full)
    assert all(torch.equal(s["down_proj.weight_scales"], t1["down_proj.weight_scales"]) for s in shards)


def test_loader_rejects_missing_and_misshapen_tensors():
    cfg = QSpecLlamaConfig.tiny()
    a = _stub_model(cfg, seed=2)
    sd = C.export_state_dict(a)
    bad = dict(sd); del bad["model.layers.1.mlp.down_proj.2.weight_scales"]
    with pytest.raises(KeyError):
        C.load_state_dict(_stub_model(cfg, seed=4), bad)
    bad = dict(sd); bad["model.layers.0.self_attn.k_proj.weight"] = bad["model.layers.0.self_attn.k_proj.weight"][:-1]
    b = _stub_model(cfg, seed=4)
    before = b.layers[0].qkv_proj.weight.clone()
    with pytest.raises(ValueError):
        C.load_state_dict(b, bad)
    assert torch.equal(before, b.layers[0].qkv_proj.weight)                          # nothing written on failure
    with pytest.raises(ValueError):
        C.load_state_dict(_stub_model(QSpecLlamaConfig.tiny(intermediate_size=28 * 128), seed=4), sd)


@pytest.mark.gpu
def test_loaded_checkpoint_decodes_like_the_source_model(tmp_path):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from qspec_b200.engine import QSpecEngine
    from qspec_b200.model import QSpecLlamaModel
    cfg = QSpecLlamaConfig.tiny()
    B, k = 2, 3
    src = QSpecLlamaModel(cfg, max_batch=B, max_tokens=64, device="cuda", seed=5)
    C.save_qspec_checkpoint(src, str(tmp_path))
    dst = QSpecLlamaModel(cfg, max_batch=B, max_tokens=64, device="cuda", seed=77)   # different random init, then overwritten
    C.load_qspec_checkpoint(dst, str(tmp_path))
    prompts = torch.randint(0, cfg.vocab_size, (B, 12), generator=torch.Generator().manual_seed(1)).cuda()
    outs = []
    for m in (src, dst):
        eng = QSpecEngine(m, B, k, use_cuda_graph=False)
        eng.prefill(prompts)
        torch.cuda.synchronize()
        toks = []
        for _ in range(6):
            eng.step()
            torch.cuda.synchronize()
            toks.append(eng.d_out_pack.cpu().clone())
        outs.append((torch.stack(toks), eng.metrics()))
    assert torch.equal(outs[0][0], outs[1][0]) and outs[0][1] == outs[1][1]


@pytest.mark.parametrize("kind,K,head_dim", [("mlp", 28 * 64, 128), ("heads", 8 * 16, 16)])
def test_rerotation_algebra_is_exact_in_float(kind, K, head_dim):
    """sum_r W'_r (R_t x_r) == W (R x) with W'_r = (W R)[:, slice_r] R_t^T: the identity the TP re-rotation tool rests on."""
    g = torch.Generator().manual_seed(3)
    W = torch.randn(24, K, generator=g, dtype=torch.float64)
    x = torch.randn(5, K, generator=g, dtype=torch.float64)
    eye = torch.eye(K, dtype=torch.float64)
    R = C.rotate_rows(eye, kind, head_dim).T
    assert torch.allclose(R @ R.T, eye, atol=1e-12)                                  # orthogonal
    assert torch.allclose(C.rotate_rows(x, kind, head_dim), x @ R.T) and torch.allclose(C.rotate_rows(x, kind, head_dim, True), x @ R)
    y = C.rotate_rows(x, kind, head_dim) @ W.T
    w_eff = C.rotate_rows(W, kind, head_dim, transpose=True)
    for tp in (2, 4):
        k, acc = K // tp, 0
        for r in range(tp):
            wl = C.rotate_rows(w_eff[:, r * k:(r + 1) * k].contiguous(), kind, head_dim)
            acc = acc + C.rotate_rows(x[:, r * k:(r + 1) * k].contiguous(), kind, head_dim) @ wl.T
        assert torch.allclose(acc, y, atol=1e-9)


def test_rerotated_int4_shards_reproduce_the_global_layer_within_quantisation_noise():
    """down_proj of a 'real' checkpoint (weights folded with the global rotation) at TP = 2: the re-rotated, re-quantised shards with
    shard-local rotations give the global result up to a second RTN error; slicing the bytes as they are does not."""
    from qspec_b200.nn import unpack_i4
    g = torch.Generator().manual_seed(4)
    N, K, tp = 64, 28 * 64, 2
    w_fp = torch.randn(N, K, generator=g) * 0.05
    folded = C.rotate_rows(w_fp.double(), "mlp").float()                                # W_ckpt = W R^T  (row <- R row)
    packed, scales = C.quantize_weight_rtn(folded, C.rtn_scales(folded).float())
    deq = unpack_i4(packed).double() * scales.double()
    x = torch.randn(16, K, generator=g, dtype=torch.float64)
    y_ref = C.rotate_rows(x, "mlp") @ deq.T                                              # what TP = 1 computes
    shards = C.rerotate_row_parallel(packed, scales, "mlp", tp)
    k = K // tp
    y_tp, y_naive = 0, 0
    for r, (wp, sc) in enumerate(shards):
        assert wp.dtype == torch.uint8 and tuple(wp.shape) == (N, k // 2) and tuple(sc.shape) == (N, 1)
        xr = C.rotate_rows(x[:, r * k:(r + 1) * k].contiguous(), "mlp")                 # shard-local online rotation
        y_tp = y_tp + xr @ (unpack_i4(wp).double() * sc.double()).T
        y_naive = y_naive + xr @ deq[:, r * k:(r + 1) * k].T
    rel = lambda a: float((a - y_ref).norm() / y_ref.norm())
    assert rel(y_tp) < 0.2 and rel(y_naive) > 0.9
    assert C.rerotate_row_parallel(packed, scales, "mlp", 1)[0][0].data_ptr() == packed.data_ptr()
    with pytest.raises(ValueError):
        C.rerotate_row_parallel(packed, scales, "mlp", 3)
    # rank=: only that shard is computed, identical to the entry of the full list
    one = C.rerotate_row_parallel(packed, scales, "mlp", tp, rank=1)
    assert torch.equal(one[0], shards[1][0]) and torch.equal(one[1], shards[1][1])
    with pytest.raises(ValueError):
        C.rerotate_row_parallel(packed, scales, "mlp", tp, rank=2)


def test_rotation_matrices_are_the_online_rotations_of_the_oracle():
    """Ties checkpoint.rotate_rows to the rotations the kernels apply at run time (the algebra tests above build R from
    rotate_rows itself). get_had28 happens to be symmetric (checked here), so A vs A^T cannot differ for Llama-3 sizes."""
    x = O.make_act(3, 28 * 64, 77)
    ref = O.hadamard_k28(x, 64).astype(np.float64)                                      # f16-rounded oracle
    got = C.rotate_rows(torch.from_numpy(x.astype(np.float64)), "mlp").numpy()
    assert np.abs(got - ref).max() <= 2e-3 * np.abs(ref).max()
    h28 = O.had28_matrix()
    assert np.array_equal(h28, h28.T)
    xh = O.make_act(3, 8 * 128, 78)
    refh = O.hadamard_heads(xh, 8, 128).astype(np.float64)
    goth = C.rotate_rows(torch.from_numpy(xh.astype(np.float64)), "heads", 128).numpy()
    assert np.abs(goth - refh).max() <= 2e-3 * np.abs(refh).max()


def test_load_state_dict_rerotates_by_default_at_tp_gt_1():
    """A reference-keyed (i.e. real, globally rotated) checkpoint loaded into a TP shard goes through the re-rotation unless
    the caller opts out; the synthetic-weight path (reference_keys=False or rerotate=False) slices the bytes."""
    cfg = QSpecLlamaConfig.tiny(num_attention_heads=4, num_key_value_heads=2)
    full = _stub_model(cfg, seed=1)
    sd = C.export_state_dict(full)
    a, b = _stub_model(cfg, seed=5, tp=2, rank=1), _stub_model(cfg, seed=6, tp=2, rank=1)
    C.load_state_dict(a, sd)                                  # default: rerotate
    C.load_state_dict(b, sd, rerotate=False)                  # explicit opt-out: byte slices
    K2 = full.layers[0].down_proj.weight.shape[1]
    assert torch.equal(b.layers[0].down_proj.weight.view(torch.uint8), full.layers[0].down_proj.weight.view(torch.uint8)[:, K2 // 2:])
    assert not torch.equal(a.layers[0].down_proj.weight.view(torch.uint8), b.layers[0].down_proj.weight.view(torch.uint8))
    assert torch.equal(a.layers[0].qkv_proj.weight, b.layers[0].qkv_proj.weight)          # column-parallel: untouched
