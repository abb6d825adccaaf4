"""Generate tests/golden/*.npy|npz by running the REFERENCE's own pure-python helpers.

Run in the authoring container only (needs /root/reference); the outputs are committed so the
tests never read /root/reference.  What is imported from the reference (by file path, never
`import quarot`, whose __init__ pulls CUDA extensions):

* third-party/QuaRot/quarot/functional/quantization.py  -> pack_i4, unpack_i4, sym_quant
* third-party/QuaRot/quarot/functional/hadamard.py      -> get_had28, get_hadK, matmul_hadU
  (its `fast_hadamard_transform*` imports are stubbed in sys.modules; only CPU code is called)
* third-party/ao/test/test_rowwise_scaled_linear_cutlass.py:68-76 -> the `output_ref` formula,
  restated below with torch on CPU on that test's six (M,N,K) shapes.

Usage:  python tests/golden/make_golden.py
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    for stub in ("fast_hadamard_transform", "fast_hadamard_transform_cuda"):
        sys.modules.setdefault(stub, types.ModuleType(stub))
    rq = _load("ref_quantization", f"{REF}/third-party/QuaRot/quarot/functional/quantization.py")
    rh = _load("ref_hadamard", f"{REF}/third-party/QuaRot/quarot/functional/hadamard.py")

    g = torch.Generator().manual_seed(1234)

    # --- a1: pack / unpack -------------------------------------------------------------
    q = torch.randint(-8, 8, (16, 64), generator=g, dtype=torch.int8)
    packed = rq.pack_i4(q)
    assert torch.equal(rq.unpack_i4(packed).to(torch.int8), q)
    np.savez(os.path.join(HERE, "pack_i4.npz"), q=q.numpy(), packed=packed.numpy())

    # --- had28 literal and Hadamard definitions ------------------------------------------
    had28 = rh.get_had28()
    np.save(os.path.join(HERE, "had28.npy"), had28.to(torch.int8).numpy())
    hadK, K = rh.get_hadK(14336)
    assert K == 28 and torch.equal(hadK, had28)
    # matmul_hadU in float64 on CPU: the mathematical definition of both online rotations
    x32 = torch.randn(6, 32, generator=g, dtype=torch.float64)
    y32 = rh.matmul_hadU(x32)
    x14336 = torch.randn(3, 14336, generator=g, dtype=torch.float64)
    y14336 = rh.matmul_hadU(x14336)
    x512 = torch.randn(5, 512, generator=g, dtype=torch.float64)
    y512 = rh.matmul_hadU(x512)
    np.savez(os.path.join(HERE, "matmul_hadU.npz"), x32=x32.numpy(), y32=y32.numpy(),
             x512=x512.numpy(), y512=y512.numpy(),
             x14336=x14336.numpy(), y14336=y14336.numpy())

    # --- reference fake-quant sym_quant (functional/quantization.py:32-35) ----------------
    xs = torch.randn(8, 256, generator=g, dtype=torch.float32)
    sc = xs.abs().amax(dim=1, keepdim=True) / 7
    qs, _ = rq.sym_quant(xs, sc, torch.tensor(7))
    np.savez(os.path.join(HERE, "sym_quant_fake.npz"), x=xs.numpy(), scale=sc.numpy(),
             q=qs.to(torch.int8).numpy())

    # --- a3: W4A4 output_ref on the torchao test shapes -----------------------------------
    shapes = [(2, 512, 128), (3, 2048, 2048), (4, 3584, 640), (13, 8704, 8576),
              (26, 18944, 1664), (67, 6656, 1408)]
    out = {}
    for i, (M, N, K) in enumerate(shapes):
        xq = torch.randint(-8, 8, (M, K), generator=g, dtype=torch.int8)
        wq = torch.randint(-8, 8, (N, K), generator=g, dtype=torch.int8)
        sx = ((0.5 + torch.rand(M, generator=g)) * 2.0 ** -3).to(torch.float16)
        sw = ((0.5 + torch.rand(N, generator=g)) * 2.0 ** -7).to(torch.float16)
        bias = torch.rand(N, generator=g).to(torch.float16)
        # test_rowwise_scaled_linear_cutlass.py:50,60 packing
        xp = ((xq[..., 1::2] << 4) | (xq[..., 0::2] & 0xF)).to(torch.int8)
        wp = ((wq[:, 1::2] << 4) | (wq[:, 0::2] & 0xF)).to(torch.int8)
        assert torch.equal(xp.view(torch.uint8), rq.pack_i4(xq))
        # :68-76
        ref = (xq.float() @ wq.float().T) * sx.float().view(M, 1) * sw.float().view(1, N)
        ref_b = (ref + bias.float()).to(torch.float16)
        ref = ref.to(torch.float16)
        # keep files small: store packed operands for the small shapes, a column sample otherwise
        cols = torch.arange(0, N, max(1, N // 256))[:256]
        out[f"s{i}_mnk"] = np.array([M, N, K])
        out[f"s{i}_seed_xp"] = xp.numpy()
        out[f"s{i}_sx"] = sx.numpy()
        out[f"s{i}_cols"] = cols.numpy()
        out[f"s{i}_wp_cols"] = wp[cols].numpy()
        out[f"s{i}_sw_cols"] = sw[cols].numpy()
        out[f"s{i}_bias_cols"] = bias[cols].numpy()
        out[f"s{i}_ref_cols"] = ref[:, cols].numpy()
        out[f"s{i}_refb_cols"] = ref_b[:, cols].numpy()
    np.savez_compressed(os.path.join(HERE, "w4a4_output_ref.npz"), **out)

    # --- Qwen2.5-14B variant: had108 (intermediate 13824 = 108 * 128) and had40 (40 heads) ------
    had108, had40 = rh.get_had108(), rh.get_had40()
    np.save(os.path.join(HERE, "had108.npy"), had108.to(torch.int8).numpy())
    np.save(os.path.join(HERE, "had40.npy"), had40.to(torch.int8).numpy())
    hk, K = rh.get_hadK(13824)
    assert K == 108 and torch.equal(hk, had108)
    hk, K = rh.get_hadK(40)
    assert K == 40 and torch.equal(hk, had40)
    x13824 = torch.randn(2, 13824, generator=g, dtype=torch.float64)
    x40 = torch.randn(7, 40, generator=g, dtype=torch.float64)
    np.savez(os.path.join(HERE, "matmul_hadU_qwen.npz"), x13824=x13824.numpy(), y13824=rh.matmul_hadU(x13824).numpy(),
             x40=x40.numpy(), y40=rh.matmul_hadU(x40).numpy())
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
