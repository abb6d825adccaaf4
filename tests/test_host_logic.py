"""CPU tests: C-ABI library loads and exports every declared symbol, plugin/host logic, TP sharding
with a world_size-2 gloo group (no compute kernels are called: there is no GPU here)."""
import os
import re
import socket

import numpy as np
import pytest
import torch

from oracle import qspec_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_header_symbol():
    import qspec_b200
    lib = qspec_b200._lib.load()
    hdr = open(os.path.join(ROOT, "include", "qspec_b200.h")).read()
    declared = set(re.findall(r"\b(qspec_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 19
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
        assert name in qspec_b200._lib.SIGNATURES, f"{name} has no ctypes signature"
    assert lib.qspec_abi_version() == 1
    # pure host entry points are callable without a GPU
    assert lib.qspec_gemm_w4a4_workspace_bytes(16, 4096, 4096) >= 16 * 4096 * 4
    assert lib.qspec_gemm_w4a16_workspace_bytes(16, 4096, 4096) > 0


def test_no_cpu_fallback():
    from qspec_b200 import ops
    x = torch.zeros(2, 128, dtype=torch.float16)
    with pytest.raises(ValueError, match="no CPU path"):
        ops.fuse_sym_quant(x, torch.zeros(2, dtype=torch.float16), torch.zeros(2, 64, dtype=torch.int8))
    with pytest.raises(ValueError, match="no CPU path"):
        ops.gemm_w4a16(x, torch.zeros(16, 64, dtype=torch.int8), torch.zeros(16, dtype=torch.float16), None,
                       torch.zeros(2, 16, dtype=torch.float16))


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "qspec_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f"{f} imports oracle/"


def test_pack_helpers_match_oracle():
    from qspec_b200.nn import pack_i4, unpack_i4
    q = torch.randint(-8, 8, (5, 64), dtype=torch.int8)
    assert np.array_equal(pack_i4(q).numpy(), O.pack_i4(q.numpy()))
    assert np.array_equal(unpack_i4(pack_i4(q)).numpy(), q.numpy())
    with pytest.raises(AssertionError):
        pack_i4(torch.full((1, 2), 9, dtype=torch.int8))


def test_plugin_registration_and_create_weights():
    """Modelled on tests/quantization/test_register_quantization_config.py:21-116 of the reference."""
    from qspec_b200 import plugin
    cfg_cls = plugin.get_quantization_config("qspec")
    assert cfg_cls is plugin.QSpecConfig
    with pytest.raises(ValueError):
        plugin.register_quantization_config("qspec")(plugin.QSpecConfig)      # already registered
    with pytest.raises(ValueError):
        plugin.get_quantization_config("nope")
    cfg = cfg_cls.from_config({"input_clip_ratio": 0.9})
    assert cfg.get_name() == "qspec" and cfg.get_min_capability() == 100
    method = cfg.get_quant_method(torch.nn.Module(), prefix="model.layers.0.self_attn.qkv_proj")
    assert isinstance(method, plugin.LinearMethodBase)
    layer = torch.nn.Module()
    method.create_weights(layer, 4096, [4096, 1024, 1024], 4096, 6144, torch.float16, weight_loader=lambda *a: None)
    assert layer.weight.shape == (6144, 2048) and layer.weight.dtype == torch.int8
    assert layer.weight_scales.shape == (6144, 1)
    assert layer.weight.input_dim == 1 and layer.weight.output_dim == 0 and hasattr(layer.weight, "weight_loader")
    with pytest.raises(ValueError):
        method.create_weights(torch.nn.Module(), 100, [64], 100, 64, torch.float16)
    # mode switch
    assert plugin.get_qspec_mode() is False
    with plugin.qspec_mode(True):
        assert plugin.get_qspec_mode() is True
    assert plugin.get_qspec_mode() is False


def test_module_mirrors_keep_reference_signatures():
    import inspect
    from qspec_b200 import nn
    lin = nn.Linear4bit(256, 128)
    assert lin.weight.shape == (128, 128) and lin.weight.dtype == torch.uint8          # linear.py:38-40
    assert lin.weight_scales.shape == (128, 1) and lin.bias is None
    assert list(inspect.signature(nn.Linear4bit.forward).parameters)[:3] == ["self", "x", "C"]
    assert list(inspect.signature(nn.Quantizer.forward).parameters)[:4] == ["self", "x", "scale", "out"]
    assert list(inspect.signature(nn.OnlineHadamard.forward).parameters)[:3] == ["self", "x", "out"]
    had = nn.OnlineHadamard(14336)
    assert had.rem_dim == 28 and had.had_rem_dim.shape == (28, 28)
    assert nn.OnlineHadamard(32).rem_dim == 1
    with pytest.raises(NotImplementedError):
        nn.OnlineHadamard(172 * 64)
    fl = torch.nn.Linear(64, 32, bias=False).half()
    sc = (fl.weight.abs().amax(1, keepdim=True) / 7).half()
    q = nn.Linear4bit.from_float(fl, sc)
    ref = torch.clamp((fl.weight.float() / sc.float()).round(), -8, 7).to(torch.int8)
    assert torch.equal(nn.unpack_i4(q.weight).to(torch.int8), ref)
    with pytest.raises(AssertionError):                                              # linear.py:72
        lin.forward_w4a4(torch.zeros(1, 128))


def test_sharding_helpers_reassemble():
    from qspec_b200 import parallel as P
    hq, hkv, d, hid, inter = 8, 4, 16, 128, 256
    g = torch.Generator().manual_seed(0)
    w = torch.randint(-128, 127, ((hq + 2 * hkv) * d, hid // 2), generator=g, dtype=torch.int8)
    s = torch.rand((hq + 2 * hkv) * d, 1, generator=g).half()
    parts = [P.shard_qkv(w, s, hq, hkv, d, 2, r) for r in range(2)]
    assert parts[0][0].shape == ((hq + 2 * hkv) * d // 2, hid // 2)
    q = torch.cat([p[0][: hq // 2 * d] for p in parts]);  assert torch.equal(q, w[: hq * d])
    wg = torch.randint(-128, 127, (2 * inter, hid // 2), generator=g, dtype=torch.int8)
    sg = torch.rand(2 * inter, generator=g).half()
    gp = [P.shard_gate_up(wg, sg, inter, 2, r) for r in range(2)]
    assert torch.equal(torch.cat([p[0][: inter // 2] for p in gp]), wg[:inter])      # up halves
    assert torch.equal(torch.cat([p[0][inter // 2:] for p in gp]), wg[inter:])       # gate halves
    wr = torch.randint(-128, 127, (hid, 512 // 2), generator=g, dtype=torch.int8)
    rp = [P.shard_row(wr, torch.ones(hid).half(), 2, r) for r in range(2)]
    assert torch.equal(torch.cat([p[0] for p in rp], 1), wr)
    with pytest.raises(ValueError):
        P.shard_row(wr, torch.ones(hid).half(), 3, 0)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _tp_worker(rank, world, port, ret):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        from qspec_b200 import parallel as P
        M, N, K = 5, 64, 512
        wp, sW = O.make_weight(N, K, 42)
        x = O.make_act(M, K, 43)
        # row-parallel W4A16: each rank multiplies its K-slice (CPU oracle arithmetic), all-reduce sums
        wl, sl = P.shard_row(torch.from_numpy(wp), torch.from_numpy(sW), world, rank)
        k = K // world
        xl = x[:, rank * k:(rank + 1) * k]
        part = xl.astype(np.float64) @ (O.unpack_i4(wl.numpy()).astype(np.float64) * sl.numpy().astype(np.float64)[:, None]).T
        t = torch.from_numpy(part)
        dist.all_reduce(t)
        full = x.astype(np.float64) @ (O.unpack_i4(wp).astype(np.float64) * sW.astype(np.float64)[:, None]).T
        ok1 = np.allclose(t.numpy(), full, rtol=1e-12, atol=1e-9)
        # row-parallel W4A4: exact int32 partial accumulators sum to the full accumulator
        a = O.pack_i4(np.random.default_rng(1).integers(-8, 8, (M, K), dtype=np.int8))
        al = a[:, rank * k // 2:(rank + 1) * k // 2]
        acc = torch.from_numpy(O.gemm_w4a4_acc(al, wl.numpy()).astype(np.int64))
        dist.all_reduce(acc)
        ok2 = np.array_equal(acc.numpy(), O.gemm_w4a4_acc(a, wp))
        # vocab-sharded greedy token
        V = 40
        logits = torch.from_numpy(np.random.default_rng(7).standard_normal((M, V)).astype(np.float32))
        shard = V // world
        loc = logits[:, rank * shard:(rank + 1) * shard]
        val, idx = loc.max(-1)
        tok = P.combine_sharded_argmax(val, idx, rank, shard)
        ok3 = torch.equal(tok, logits.argmax(-1))
        ret[rank] = (ok1, ok2, ok3)
    finally:
        dist.destroy_process_group()


def test_tp2_gloo_row_parallel_and_sharded_argmax():
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    with mp.get_context("spawn").Manager() as mgr:      # no fork() of the multi-threaded pytest process
        ret = mgr.dict()
        mp.spawn(_tp_worker, args=(world, port, ret), nprocs=world, join=True)
        assert all(all(v) for v in ret.values()), dict(ret)
        assert len(ret) == world


def test_plugin_binds_to_the_importable_vllm_registry():
    """The plugin classes derived from the INSTALLED vLLM's own base classes register with its registry and are found by
    its lookup, as in the reference's tests/quantization/test_register_quantization_config.py:21-116 (custom config
    registered by decorator, selected by name, get_quant_method handing out the linear method whose create_weights
    registers the parameters)."""
    try:
        from vllm.model_executor.layers.quantization import get_quantization_config as v_get
        from vllm.model_executor.layers.quantization.base_config import QuantizationConfig as VQC
        from vllm.model_executor.layers.linear import LinearMethodBase as VLMB
    except Exception as e:                                # noqa: BLE001
        pytest.skip(f"no importable vLLM: {type(e).__name__}")
    import torch
    from qspec_b200 import plugin
    cfg_cls, method_cls = plugin.bind_to_vllm()
    assert plugin.bind_to_vllm() == (cfg_cls, method_cls)                       # idempotent
    assert issubclass(cfg_cls, VQC) and issubclass(cfg_cls, plugin.QSpecConfig)
    assert issubclass(method_cls, VLMB) and issubclass(method_cls, plugin.QSpecLinearMethod)
    assert v_get("qspec") is cfg_cls                                            # vLLM's own lookup finds it
    cfg = cfg_cls.from_config({"input_clip_ratio": 0.9})
    assert cfg.get_name() == "qspec" and cfg.input_clip_ratio == 0.9 and cfg.get_min_capability() == 100
    layer = torch.nn.Module()
    method = cfg.get_quant_method(layer, "model.layers.0.self_attn.qkv_proj")
    assert isinstance(method, VLMB)
    method.create_weights(layer, 256, [128, 64, 64], 256, 256, torch.float16, weight_loader=lambda *a: None)
    assert tuple(layer.weight.shape) == (256, 128) and layer.weight.dtype == torch.int8
    assert tuple(layer.weight_scales.shape) == (256, 1) and layer.weight.input_dim == 1 and layer.weight.packed_dim == 1
    assert callable(layer.weight.weight_loader)
