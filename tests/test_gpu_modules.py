"""The drop-in boundary above the C ABI on the GPU: the module mirrors of quarot_nn wired the way
vllm/model_executor/models/quarot_llama.py wires them (attention :188-241, MLP :277-298, decoder layer :374-389), and the vLLM
linear-method plugin (8(b1): create_weights / apply + the w4a4 mode switch), each stage against the CPU oracle.
Draft mode: packed bytes, scales and f16 outputs bit-exact. Verify mode: norm bit-exact, W4A16 within 1e-2 relative."""
import numpy as np
import pytest
import torch

from oracle import qspec_oracle as O

pytestmark = pytest.mark.gpu
EPS = 1e-5


@pytest.fixture(scope="module")
def Q():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from qspec_b200 import nn
    return nn


def _linear(Q, N, K, seed):
    """Linear4bit filled the way load_state_dict fills the reference module: uint8 `weight` [N, K/2], `weight_scales` [N, 1]."""
    wp, sW = O.make_weight(N, K, seed)
    lin = Q.Linear4bit(K, N, bias=False)
    lin.weight.copy_(torch.from_numpy(wp))
    lin.weight_scales.copy_(torch.from_numpy(sW).view(-1, 1))
    return lin.cuda(), wp, sW


def _rel(a, b):
    return float(np.abs(a.astype(np.float32) - b.astype(np.float32)).max() / np.abs(b.astype(np.float32)).max())


def test_attention_side_chain_draft_and_verify(Q):
    T, H, heads, D = 5, 4096, 32, 128
    x = O.make_act(T, H, 71, heavy_tail=True)
    attn = O.make_act(T, H, 72)
    qkv_proj, wq, sq = _linear(Q, 6144, H, 1)
    o_proj, wo, so = _linear(Q, H, H, 2)
    norm, had, quant = Q.RMSNorm(H, EPS, fuse=True), Q.OnlineHadamard(heads).cuda(), Q.Quantizer()
    xd, ad = torch.from_numpy(x).cuda(), torch.from_numpy(attn).cuda()
    # ---- draft (w4a4=True), caller-provided buffers as spec_decode/draft_model_runner.py:277-320 passes them
    kw = dict(w4a4=True, quantized_buffer_qkv=torch.empty(T, H // 2, dtype=torch.int8, device="cuda"),
              scale_buffer=torch.empty(T, dtype=torch.float16, device="cuda"), input_sum_buffer=torch.empty(T, dtype=torch.float16, device="cuda"))
    p = norm(xd, **kw)
    assert type(p) is Q.PackedQuantizedTensor and p.quantized_x.data_ptr() == kw["quantized_buffer_qkv"].data_ptr()
    rq, rs, _ = O.norm_quant_i4(x, EPS)
    assert np.array_equal(p.quantized_x.cpu().numpy().view(np.uint8), rq) and np.array_equal(p.scales_x.cpu().numpy(), rs)
    C = torch.empty(T, 6144, dtype=torch.float16, device="cuda")
    y = qkv_proj(p, C, **kw)
    assert y.data_ptr() == C.data_ptr()
    assert np.array_equal(y.cpu().numpy(), O.gemm_w4a4(rq, rs, wq, sq))
    # o_proj input: transpose -> H_32 over heads -> transpose back -> quantiser (quarot_llama.py:230-238)
    t = ad.view(T, heads, D).transpose(-1, -2).reshape(T * D, heads)
    h = had(t, **kw).view(T, D, heads).transpose(-1, -2).contiguous().view(T, H)
    assert had.qspec is True                                                        # sticky flag, hadamard.py:32-33
    rh = O.hadamard_heads(attn, heads, D)
    assert np.array_equal(h.cpu().numpy(), rh)
    pq = quant(h, **kw)
    rq2, rs2 = O.sym_quant_i4(rh, 1.0)
    assert np.array_equal(pq.quantized_x.cpu().numpy().view(np.uint8), rq2) and np.array_equal(pq.scales_x.cpu().numpy(), rs2)
    out = o_proj(pq, **kw)
    assert np.array_equal(out.cpu().numpy(), O.gemm_w4a4(rq2, rs2, wo, so))
    # ---- verify (no w4a4 kwarg): same modules, same weight tensors
    yn = norm(xd)
    rn = O.norm_f16(x, EPS)
    assert np.array_equal(yn.cpu().numpy(), rn)
    assert _rel(qkv_proj(yn).cpu().numpy(), O.gemm_w4a16(rn, wq, sq)) <= 1e-2
    assert quant(h) is h                                                            # quantization.py:28-29
    assert _rel(o_proj(h).cpu().numpy(), O.gemm_w4a16(rh, wo, so)) <= 1e-2
    with pytest.raises(AssertionError):                                             # linear.py:91 asserts a PackedQuantizedTensor
        qkv_proj(yn, w4a8=True)


def test_mlp_side_chain_draft_and_verify(Q):
    T, H, I = 4, 4096, 14336
    act = O.make_act(T, I, 81, heavy_tail=True)
    down, wd, sd = _linear(Q, H, I, 3)
    had, quant = Q.OnlineHadamard(I).cuda(), Q.Quantizer()
    assert had.rem_dim == 28 and tuple(had.had_rem_dim.shape) == (28, 28)
    ad = torch.from_numpy(act).cuda()
    h = had(ad, w4a4=True)
    assert tuple(h.shape) == (T, 28, 512)                                          # hadamard.py:112
    rh = O.hadamard_k28(act, 512)
    assert np.array_equal(h.reshape(T, I).cpu().numpy(), rh)
    pq = quant(h.view(T, I), w4a4=True)
    rq, rs = O.sym_quant_i4(rh, 1.0)
    assert np.array_equal(pq.quantized_x.cpu().numpy().view(np.uint8), rq) and np.array_equal(pq.scales_x.cpu().numpy(), rs)
    assert np.array_equal(down(pq, w4a4=True).cpu().numpy(), O.gemm_w4a4(rq, rs, wd, sd))
    assert _rel(down(h.view(T, I)).cpu().numpy(), O.gemm_w4a16(rh, wd, sd)) <= 1e-2


def test_plugin_create_weights_apply_both_modes(Q):
    from qspec_b200 import plugin
    T, K, parts = 6, 4096, [4096, 1024, 1024]
    N = sum(parts)
    cfg = plugin.get_quantization_config("qspec").from_config({"input_clip_ratio": 1.0})
    layer = torch.nn.Module()
    method = cfg.get_quant_method(layer, "model.layers.0.self_attn.qkv_proj")
    method.create_weights(layer, K, parts, K, N, torch.float16, weight_loader=lambda *a, **k: None)
    assert tuple(layer.weight.shape) == (N, K // 2) and tuple(layer.weight_scales.shape) == (N, 1)
    assert layer.weight.input_dim == 1 and layer.weight.output_dim == 0 and hasattr(layer.weight, "weight_loader")
    wp, sW = O.make_weight(N, K, 5)
    layer.weight.data.copy_(torch.from_numpy(wp).view(torch.int8))
    layer.weight_scales.data.copy_(torch.from_numpy(sW).view(-1, 1))
    layer.cuda()
    method.process_weights_after_loading(layer)
    x = O.make_act(T, K, 91, heavy_tail=True)
    xd = torch.from_numpy(x).cuda()
    bias = np.random.default_rng(2).standard_normal(N).astype(np.float16)
    bd = torch.from_numpy(bias).cuda()
    # verify mode (default)
    assert plugin.get_qspec_mode() is False
    out = method.apply(layer, xd.view(2, 3, K), bd)
    assert tuple(out.shape) == (2, 3, N)
    assert _rel(out.view(T, N).cpu().numpy(), O.gemm_w4a16(x, wp, sW, bias)) <= 1e-2
    # draft mode: the caller hands either the PackedQuantizedTensor of the fused norm or plain f16 activations
    rq, rs = O.sym_quant_i4(x, 1.0)
    with plugin.qspec_mode(True):
        out4 = method.apply(layer, xd, bd)
        packed = Q.PackedQuantizedTensor(torch.from_numpy(rq).view(torch.int8).cuda(), torch.from_numpy(rs).cuda())
        out4p = method.apply(layer, packed, bd)
    ref4 = O.gemm_w4a4(rq, rs, wp, sW, bias)
    assert np.array_equal(out4.cpu().numpy(), ref4) and np.array_equal(out4p.cpu().numpy(), ref4)
    with pytest.raises(ValueError):
        method.apply(layer, packed)                                                 # packed input outside draft mode
    with pytest.raises(ValueError):
        method.create_weights(torch.nn.Module(), 100, [64], 100, 64, torch.float16)
