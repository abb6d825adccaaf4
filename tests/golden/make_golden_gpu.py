"""Generate tests/golden/ref_kernels_b200.npz by running the REFERENCE's OWN CUDA kernels on a GPU.

The modules come from oracle/_ref/ (built by oracle/build_ref.py from the reference sources where they lie under
/root/reference; nothing of the reference is copied into the repo). Run on the GPU box:

    gpurun -- 'python tests/golden/make_golden_gpu.py gpurun_out/ref_kernels_b200.npz'

then copy the file into tests/golden/. The committed vectors let the CPU suite (tests/test_oracle.py) check the numpy oracle
against real outputs of K1/K2 (layernorm_kernels.cu), K3 (quant.cu), K4 (fast_hadamard_transform_cuda.cu) and of the reference's
down_proj rotation `hadK @ FHT(x)` (quarot/functional/hadamard.py:109-113, torch fp16 matmul on the GPU) without a GPU and without
/root/reference. Inputs are stored next to the outputs, so the fixtures do not depend on a random generator's version.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))

from oracle import qspec_oracle as O  # noqa: E402  (only for the seeded input generator and the had28 matrix)

import fast_hadamard_transform_cuda  # noqa: E402
import layernorm_ops  # noqa: E402
import quarot_quant_ref  # noqa: E402

EPS = 1e-5


def main(path):
    d = {}
    # K1 / K2: norm (+ int4 quant); row 1 all zero, row 2 with outliers
    x = O.make_act(12, 4096, 101, heavy_tail=True)
    x[1, :] = 0
    x[2, 5], x[2, 6] = np.float16(300.0), np.float16(-300.0)
    xd = torch.from_numpy(x).cuda()
    q = torch.zeros(12, 2048, dtype=torch.int8, device="cuda")
    isum = torch.zeros(12, dtype=torch.float16, device="cuda")
    sc = torch.zeros(12, dtype=torch.float16, device="cuda")
    layernorm_ops.rms_norm_general_fuse_sum_i4(q, xd, isum, sc, EPS, True)
    y = torch.zeros_like(xd)
    layernorm_ops.rms_norm_general_fuse_sum_fp16(y, xd, EPS)
    d.update(k12_x=x, k1_q=q.cpu().numpy().view(np.uint8), k1_scale=sc.cpu().numpy(), k1_input_sum=isum.cpu().numpy(),
             k2_y=y.cpu().numpy())
    # K3: fuse_sym_quant; zero row and +-65504
    for tag, K, seed in (("a", 4096, 102), ("b", 14336, 103)):
        x = O.make_act(6, K, seed, heavy_tail=True)
        x[1, :] = 0
        x[2, 5], x[2, 6] = np.float16(65504.0), np.float16(-65504.0)
        xd = torch.from_numpy(x).cuda()
        q = torch.zeros(6, K // 2, dtype=torch.int8, device="cuda")
        sc = torch.zeros(6, dtype=torch.float16, device="cuda")
        quarot_quant_ref.fuse_sym_quant(xd, sc, q, 1.0)
        d.update({f"k3{tag}_x": x, f"k3{tag}_q": q.cpu().numpy().view(np.uint8), f"k3{tag}_scale": sc.cpu().numpy()})
    # K4: FHT at the two QSpec sizes (+ the had28 mix of the down_proj rotation)
    s32 = float(1.0 / torch.tensor(32).sqrt())
    s14336 = float(1.0 / torch.tensor(14336).sqrt())
    x = O.make_act(2 * 128, 32, 104, heavy_tail=True)
    xd = torch.from_numpy(x).cuda()
    o = torch.zeros_like(xd)
    fast_hadamard_transform_cuda.faster_fast_hadamard_transform(xd, s32, o)
    d.update(k4a_x=x, k4a_scale=np.float32(s32), k4a_y=o.cpu().numpy())
    x = O.make_act(3, 14336, 105, heavy_tail=True)
    xd = torch.from_numpy(x).cuda().view(-1, 512)
    o = torch.zeros_like(xd)
    o = fast_hadamard_transform_cuda.faster_fast_hadamard_transform(xd, s14336, o)
    hadK = torch.from_numpy(O.had28_matrix().astype(np.float16)).cuda()
    mixed = hadK @ o.view(-1, 28, 512)                                   # hadamard.py:112
    d.update(k4b_x=x, k4b_scale=np.float32(s14336), k4b_fht=o.cpu().numpy().reshape(3, 14336),
             k5_y=mixed.cpu().numpy().reshape(3, 14336))
    torch.cuda.synchronize()
    d["device"] = np.array(torch.cuda.get_device_name(0))
    np.savez_compressed(path, **d)
    # W4A8 mode (second file, same recipe): norm + int8 quantiser and fuse_sym_quant_i8
    e = {}
    x = d["k12_x"]
    xd = torch.from_numpy(x).cuda()
    q8 = torch.zeros(12, 4096, dtype=torch.int8, device="cuda")
    isum = torch.zeros(12, dtype=torch.float16, device="cuda")
    sc = torch.zeros(12, dtype=torch.float16, device="cuda")
    layernorm_ops.rms_norm_general_fuse_sum_i8(q8, xd, isum, sc, EPS, True)
    e.update(k1i8_x=x, k1i8_q=q8.cpu().numpy(), k1i8_scale=sc.cpu().numpy())
    if hasattr(quarot_quant_ref, "fuse_sym_quant_i8"):
        x = d["k3a_x"]
        xd = torch.from_numpy(x).cuda()
        q8 = torch.zeros(6, 4096, dtype=torch.int8, device="cuda")
        sc = torch.zeros(6, dtype=torch.float16, device="cuda")
        quarot_quant_ref.fuse_sym_quant_i8(xd, sc, q8, 1.0)
        e.update(k3i8_x=x, k3i8_q=q8.cpu().numpy(), k3i8_scale=sc.cpu().numpy())
    torch.cuda.synchronize()
    np.savez_compressed(path.replace(".npz", "_w4a8.npz"), **e)
    print("wrote", path, {k: getattr(v, "shape", None) for k, v in d.items()})


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "ref_kernels_b200.npz"))
