"""TEST INFRASTRUCTURE -- builds the reference's OWN CUDA kernels for the activation side of the path into oracle/_ref/.

The reference ships no test or golden vector for K1/K2 (norm + int4 quant), K3 (fuse_sym_quant) and the QSpec FHT entry point
(SURVEY 8(c)); the only way to pin the numpy oracle for them is to run the reference kernels themselves. This recipe compiles the
reference sources WHERE THEY LIE under /root/reference (nothing is copied into the repo) as three torch extension modules, with the
flags of the reference's own setup.py files, for the GPU box (sm_100):

  layernorm_ops                 third-party/kernels/csrc/layernorm.cpp + layernorm_kernels.cu   (setup.py: -O2 --use_fast_math)
  fast_hadamard_transform_cuda  third-party/fast-hadamard-transform/csrc/fast_hadamard_transform.cpp + ..._cuda.cu (setup.py: -O3 --use_fast_math)
  quarot_quant_ref              third-party/QuaRot/quarot/kernels/quant.cu + oracle/ref_quant_binding.cpp
                                (the reference's bindings.cpp also pulls its CUTLASS GEMM and flashinfer sources, which are not on
                                this path; the binding here calls the same host launcher with the same block size, bindings.cpp:128-146)

  triton_w4a16_ref.cubin/.json  vllm/model_executor/layers/quarot_nn/qspec_gemm.py: the reference's own Triton statement of the W4A16 GEMM
                                (awq_gemm_kernel_sym_nozp_perchannel, fp32 accumulate), imported by path and AOT-compiled for sm_100 with the
                                wrapper's default tile (64 x 64 x 32, 4 warps); launched on the GPU box through the CUDA driver API. A checker
                                artefact like the .so files above: Triton is never on the product path.

Outputs only into oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot). /root/reference does not exist on the GPU
box: this runs in the build container only (`__graft_entry__.build()` calls it when the reference tree is present). Used by
tests/test_gpu_ref_kernels.py (through tests/ref_kernels_worker.py) as a checker; never imported by the product.

  python oracle/build_ref.py [--force]
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("QSPEC_REFERENCE_ROOT", "/root/reference")
TP = os.path.join(REF, "third-party")
EXT = sysconfig.get_config_var("EXT_SUFFIX")

COMMON_NVCC = ["-std=c++17", "-Xcompiler", "-fPIC", "-gencode", "arch=compute_100,code=sm_100",
               "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__", "-U__CUDA_NO_HALF2_OPERATORS__",
               "-U__CUDA_NO_BFLOAT16_OPERATORS__", "-U__CUDA_NO_BFLOAT16_CONVERSIONS__", "-U__CUDA_NO_BFLOAT162_OPERATORS__",
               "-U__CUDA_NO_BFLOAT162_CONVERSIONS__", "--expt-relaxed-constexpr", "--expt-extended-lambda", "-w"]


def _cutlass_include():
    import importlib.util
    for pkg, rel in (("flashinfer", "data/cutlass/include"), ("tilelang", "3rdparty/cutlass/include")):
        spec = importlib.util.find_spec(pkg)
        if spec and spec.submodule_search_locations:
            p = os.path.join(list(spec.submodule_search_locations)[0], rel)
            if os.path.isdir(p):
                return p
    return None


def modules():
    k = os.path.join(TP, "kernels", "csrc")
    f = os.path.join(TP, "fast-hadamard-transform", "csrc")
    q = os.path.join(TP, "QuaRot", "quarot", "kernels")
    return {
        "layernorm_ops": dict(
            sources=[os.path.join(k, "layernorm.cpp"), os.path.join(k, "layernorm_kernels.cu")],
            nvcc=["-O2", "-DENABLE_BF16", "--use_fast_math"], includes=[k]),
        "fast_hadamard_transform_cuda": dict(
            sources=[os.path.join(f, "fast_hadamard_transform.cpp"), os.path.join(f, "fast_hadamard_transform_cuda.cu")],
            nvcc=["-O3", "--use_fast_math"], includes=[f]),
        "quarot_quant_ref": dict(
            sources=[os.path.join(HERE, "ref_quant_binding.cpp"), os.path.join(q, "quant.cu")],
            nvcc=["-O3"], includes=[os.path.join(q, "include"), _cutlass_include()]),
    }


def needs_build() -> bool:
    """True when the reference tree is present and an artefact is missing or older than its sources (no torch import)."""
    if not os.path.isdir(TP):
        return False
    for name, spec in modules().items():
        so = os.path.join(OUT, name + EXT)
        if not os.path.exists(so) or any(os.path.getmtime(so) < os.path.getmtime(s) for s in spec["sources"]):
            return True
    return not (os.path.exists(os.path.join(OUT, "triton_w4a16_ref.cubin")) and os.path.exists(os.path.join(OUT, "triton_w4a16_ref.json")))


def build_ref(force: bool = False, verbose: bool = False) -> dict:
    """Returns {module name: path or 'error: ...'}; never raises for a module that does not compile (the checker is optional)."""
    if not os.path.isdir(TP):
        return {}
    import torch
    from torch.utils.cpp_extension import include_paths, library_paths
    os.makedirs(OUT, exist_ok=True)
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    inc = [*include_paths("cuda"), sysconfig.get_paths()["include"]]
    results = {}

    def one(item):
        name, spec = item
        so = os.path.join(OUT, name + EXT)
        if not force and os.path.exists(so) and all(os.path.getmtime(so) >= os.path.getmtime(s) for s in spec["sources"]):
            return name, so
        flags = [*COMMON_NVCC, *spec["nvcc"], f"-D_GLIBCXX_USE_CXX11_ABI={abi}", f"-DTORCH_EXTENSION_NAME={name}",
                 "-DTORCH_API_INCLUDE_EXTENSION_H"]
        for i in [*spec["includes"], *inc]:
            if i:
                flags += ["-I", i]
        objs = []
        for s in spec["sources"]:
            o = os.path.join(OUT, name + "__" + os.path.basename(s) + ".o")
            cmd = ["nvcc", *flags, "-x", "cu", "-c", s, "-o", o]
            if verbose:
                print(" ".join(cmd), flush=True)
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                return name, "error: " + (r.stderr or r.stdout)[-1500:]
            objs.append(o)
        link = ["nvcc", "-shared", "-o", so, *objs, *[f"-L{p}" for p in library_paths("cuda")],
                "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-cudart", "shared"]
        r = subprocess.run(link, capture_output=True, text=True)
        for o in objs:
            os.remove(o)
        if r.returncode != 0:
            return name, "error: " + (r.stderr or r.stdout)[-1500:]
        return name, so

    with ThreadPoolExecutor(max_workers=3) as ex:
        for name, res in ex.map(one, modules().items()):
            results[name] = res
    return results


def build_triton_ref(force: bool = False) -> str:
    """AOT-compiles the reference's Triton W4A16 kernel (no GPU needed) -> oracle/_ref/triton_w4a16_ref.{cubin,json}."""
    src_py = os.path.join(REF, "vllm", "model_executor", "layers", "quarot_nn", "qspec_gemm.py")
    cubin = os.path.join(OUT, "triton_w4a16_ref.cubin")
    meta = os.path.join(OUT, "triton_w4a16_ref.json")
    if not os.path.exists(src_py):
        return "absent"
    if not force and os.path.exists(cubin) and os.path.exists(meta) and os.path.getmtime(cubin) >= os.path.getmtime(src_py):
        return cubin
    try:
        import importlib.util
        import json
        import triton
        from triton.backends.compiler import GPUTarget
        from triton.compiler import ASTSource
        spec = importlib.util.spec_from_file_location("ref_qspec_gemm", src_py)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        kern = mod.awq_gemm_kernel_sym_nozp_perchannel
        # the reference wrapper (qspec_gemm.py:96-133) passes int8 weights and leaves the tile at the kernel's defaults
        sig = {"a_ptr": "*fp16", "b_ptr": "*i8", "c_ptr": "*fp16", "scales_ptr": "*fp16", "M": "i32", "N": "i32", "K": "i32",
               "BLOCK_SIZE_M": "constexpr", "BLOCK_SIZE_N": "constexpr", "BLOCK_SIZE_K": "constexpr", "SPLIT_K": "constexpr"}
        consts = {"BLOCK_SIZE_M": 64, "BLOCK_SIZE_N": 64, "BLOCK_SIZE_K": 32, "SPLIT_K": 1}
        cc = triton.compile(ASTSource(fn=kern, signature=sig, constexprs=consts), target=GPUTarget("cuda", 100, 32))
        os.makedirs(OUT, exist_ok=True)
        with open(cubin, "wb") as f:
            f.write(cc.asm["cubin"])
        md = cc.metadata
        with open(meta, "w") as f:
            json.dump({"name": md.name, "num_warps": md.num_warps, "shared": md.shared, "block_m": 64, "block_n": 64, "block_k": 32,
                       "params": ["a_ptr", "b_ptr", "c_ptr", "scales_ptr", "M:i32", "N:i32", "K:i32", "global_scratch", "profile_scratch"],
                       "triton": triton.__version__}, f)
        return cubin
    except Exception as e:  # optional checker
        return f"error: {type(e).__name__}: {str(e)[:500]}"


PY_STAGED = ("vllm/model_executor/models/quarot_llama.py",)


def stage_py_reference() -> dict:
    """Stages the reference's own model file(s) that CALL the dual-mode layer into oracle/_ref/pyref/ (git-ignored checker
    artefact, like the compiled modules above; the GPU box has no /root/reference). tests/test_gpu_dropin_reference_model.py
    executes the staged, UNMODIFIED file on top of qspec_b200.nn to show that the reference's caller runs unchanged."""
    import shutil
    out = {}
    dst_dir = os.path.join(OUT, "pyref")
    for rel in PY_STAGED:
        src = os.path.join(REF, rel)
        if not os.path.exists(src):
            out[rel] = "absent"
            continue
        os.makedirs(dst_dir, exist_ok=True)
        dst = os.path.join(dst_dir, os.path.basename(rel))
        if not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            shutil.copyfile(src, dst)
        out[rel] = dst
    return out


if __name__ == "__main__":
    print("staged python reference ->", stage_py_reference())
    print("triton_w4a16_ref ->", build_triton_ref(force="--force" in sys.argv))
    for n, r in build_ref(force="--force" in sys.argv, verbose=True).items():
        print(n, "->", r)
