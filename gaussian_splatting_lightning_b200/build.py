"""In-tree build of libgsr_b200.so (hand-written CUDA for sm_100a + the C-ABI of include/gsr_b200.h).

    python -m gaussian_splatting_lightning_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU.  The shared object is written next to this file so that it
travels to the GPU box with the repository snapshot; it is git-ignored.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(CSRC, "_build")
LIB = os.path.join(HERE, "libgsr_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]

# translation unit -> extra flags.  The preprocess kernels feed integer outputs (depth bits of
# the sort keys, radii, tile rects) and therefore state every fused multiply-add explicitly.
SOURCES = {
    "api.cu": [],
    "preprocess.cu": ["-fmad=false"],
    "binning.cu": ["-fmad=false"],
    "tile_binning.cu": ["-fmad=false"],
    "radix_sort.cu": [],
    "render_forward.cu": [],
    "render_backward.cu": os.environ.get("GSR_BWD_FLAGS", "").split(),
    "preprocess_backward.cu": ["--use_fast_math"],   # gradients only (1e-4 contract): contraction, approximate div / sqrt
    "loss.cu": [],
    "optimizer.cu": ["-fmad=false"],
}
HEADERS = ["gsr_common.cuh", "gsr_math.cuh", "render_common.cuh", os.path.join("..", "..", "include", "gsr_b200.h")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    nvcc = _nvcc()
    logs = {}

    def compile_one(item):
        src, extra = item
        obj = os.path.join(BUILD, src.replace(".cu", ".o"))
        srcp = os.path.join(CSRC, src)
        if force or _stale(obj, [srcp] + hdrs):
            cmd = [nvcc, *ARCH, *COMMON, *extra, "-c", srcp, "-o", obj]
            r = subprocess.run(cmd, capture_output=True, text=True)
            logs[src] = r.stdout + r.stderr
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed on {src}:\n{' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES.items()))
    if force or _stale(LIB, objs):
        cmd = [nvcc, *ARCH, "-shared", "-o", LIB, *objs, "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if logs:
        with open(os.path.join(BUILD, "ptxas.log"), "w") as f:
            for k, v in logs.items():
                f.write(f"==== {k}\n{v}\n")
        if verbose:
            for k, v in logs.items():
                print(f"==== {k}\n{v}")
    return LIB


EXT = os.path.join(HERE, "_C.so")


def build_torch_ext(force: bool = False) -> str:
    """`_C`: the thin torch C++ extension over the C-ABI (csrc/torch_ext.cpp), compiled with g++ against the torch
    headers of this interpreter and linked to libgsr_b200.so next to it (rpath $ORIGIN).  In-tree, like the library."""
    import sysconfig

    import torch
    from torch.utils import cpp_extension as ce
    src = os.path.join(CSRC, "torch_ext.cpp")
    deps = [src, os.path.join(CSRC, "..", "..", "include", "gsr_b200.h"), LIB, os.path.abspath(__file__)]
    if not force and not _stale(EXT, deps):
        return EXT
    tlib = ce.library_paths()[0]
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", src, "-o", EXT,
           "-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           *[f"-I{i}" for i in ce.include_paths()], f"-I{sysconfig.get_paths()['include']}", f"-I{cuda_inc}",
           f"-L{tlib}", "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-ltorch_python", "-lc10", "-lc10_cuda",
           f"-L{HERE}", "-lgsr_b200", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tlib}",
           f"-L{os.path.join(os.path.dirname(cuda_inc), 'lib64')}", "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"g++ failed on torch_ext.cpp:\n{' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
    return EXT


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(p)
    print(build_torch_ext(force="--force" in sys.argv))
