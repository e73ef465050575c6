"""Recipe for oracle/_ref: compiles the reference's OWN fused-SSIM CUDA extension from the sources where they lie
(/root/reference/submodules/fused-ssim/{ssim.cu, ext.cpp}, a torch C++ extension of two files) for sm_100a.

TEST INFRASTRUCTURE ONLY.  Nothing is copied into the repository: the only output is
oracle/_ref/fused_ssim_cuda.so (git-ignored, travels to the GPU box with the snapshot).  The GPU tests load it to pin
gaussian_splatting_lightning_b200.training.fused_ssim against the reference kernel itself and bench.py times it as
the reference arm of that kernel.  The rasterizer's own CUDA source is an un-vendored submodule
and cannot be built this way.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "fused_ssim_cuda.so")
SRC_DIR = "/root/reference/submodules/fused-ssim"


def available() -> bool:
    return os.path.exists(OUT)


def build(force: bool = False) -> str | None:
    """Returns the path of the built module, or None when the reference sources are not on this machine."""
    srcs = [os.path.join(SRC_DIR, "ssim.cu"), os.path.join(SRC_DIR, "ext.cpp")]
    if not all(os.path.exists(s) for s in srcs):
        return OUT if os.path.exists(OUT) else None
    if os.path.exists(OUT) and not force and all(os.path.getmtime(OUT) >= os.path.getmtime(s) for s in srcs):
        return OUT
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(OUT_DIR, exist_ok=True)
    inc = [f"-I{p}" for p in ce.include_paths("cuda")] + [f"-I{sysconfig.get_paths()['include']}", f"-I{SRC_DIR}"]
    abi = f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"
    common = ["-O3", "-std=c++17", "-DTORCH_EXTENSION_NAME=fused_ssim_cuda", "-DTORCH_API_INCLUDE_EXTENSION_H", abi]
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    for s in srcs:
        o = os.path.join(OUT_DIR, os.path.basename(s) + ".o")
        cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", *common, "-Xcompiler", "-fPIC", *inc, "-x", "cu",
               "-c", s, "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"reference fused-ssim does not compile here:\n{' '.join(cmd)}\n{r.stderr[-3000:]}")
        objs.append(o)
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = [nvcc, "-shared", "-o", OUT, *objs, f"-L{libdir}", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda",
           "-ltorch", "-ltorch_python", "-lcudart", f"-Xlinker=-rpath,{libdir}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link of the reference fused-ssim failed:\n{r.stderr[-3000:]}")
    for o in objs:
        os.remove(o)
    return OUT


def load():
    """Import oracle/_ref/fused_ssim_cuda.so (exports fusedssim, fusedssim_backward)."""
    import importlib.util
    import torch  # noqa: F401  (the module links against libtorch)
    spec = importlib.util.spec_from_file_location("fused_ssim_cuda", OUT)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
