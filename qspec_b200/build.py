"""In-tree build of the C-ABI CUDA library (sm_100a only): qspec_b200/_C/libqspec_b200.so.

`python -m qspec_b200.build` or `__graft_entry__.build()`. nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT_DIR = os.path.join(HERE, "_C")
LIB_PATH = os.path.join(OUT_DIR, "libqspec_b200.so")
SOURCES = ["capi.cu", "act_kernels.cu", "hadamard_kernels.cu", "gemm_w4_stream.cu", "attn_kernels.cu", "qwen_kernels.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _newest_source_mtime() -> float:
    paths = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    paths.append(os.path.join(HERE, "..", "include", "qspec_b200.h"))
    return max(os.path.getmtime(p) for p in paths)


def build_variant(suffix: str, defines, verbose: bool = False) -> str:
    """Experiment builds: libqspec_b200<suffix>.so compiled with extra -D flags (select with QSPEC_LIB=<path>)."""
    os.makedirs(OUT_DIR, exist_ok=True)
    lib_path = LIB_PATH.replace(".so", f"{suffix}.so")
    nvcc = _nvcc()

    def compile_one(src):
        obj = os.path.join(OUT_DIR, src.replace(".cu", f"{suffix}.o"))
        r = subprocess.run([nvcc, *NVCC_FLAGS, *[f"-D{d}" for d in defines], "-c", os.path.join(CSRC, src), "-o", obj],
                           capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    r = subprocess.run([nvcc, "-shared", "-o", lib_path, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "shared"],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return lib_path


def build_library(force: bool = False, verbose: bool = False, prof: bool = False, watchdog: bool = False) -> str:
    """prof=True builds libqspec_b200_prof.so with the GEMM cycle counters compiled in (scripts/gpu_gemm_trace.py;
    select it with QSPEC_LIB=<path>); watchdog=True builds libqspec_b200_wd.so whose pipeline waits trap instead of
    hanging (for protocol changes). The product library carries neither."""
    os.makedirs(OUT_DIR, exist_ok=True)
    lib_path = LIB_PATH.replace(".so", "_prof.so") if prof else LIB_PATH.replace(".so", "_wd.so") if watchdog else LIB_PATH
    if not force and os.path.exists(lib_path) and os.path.getmtime(lib_path) >= _newest_source_mtime():
        return lib_path
    nvcc = _nvcc()
    objs = []
    extra = ["-DQSPEC_GEMM_PROF"] if prof else ["-DQSPEC_GEMM_WATCHDOG"] if watchdog else []
    suffix = "_prof.o" if prof else "_wd.o" if watchdog else ".o"

    def compile_one(src):
        obj = os.path.join(OUT_DIR, src.replace(".cu", suffix))
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, "-shared", "-o", lib_path, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
           "-cudart", "shared"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return lib_path


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True, prof="--prof" in sys.argv, watchdog="--watchdog" in sys.argv))
