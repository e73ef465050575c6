"""CPU check of the sub-tile culling predicate the compositing kernels use (render_common.cuh is
__host__ __device__ for exactly this): never rejects a splat that reaches alpha >= 1/255 somewhere
in the 8x4 rectangle, and is tight (few false accepts)."""
import os
import subprocess

from conftest import ROOT


def test_subtile_cull_is_conservative_and_tight(tmp_path):
    src = os.path.join(ROOT, "tests", "native", "cull_harness.cpp")
    exe = str(tmp_path / "cull_harness")
    inc = os.path.join(ROOT, "gaussian_splatting_lightning_b200", "csrc")
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-I", inc, "-I", cuda_inc, "-x", "c++", src,
                    "-o", exe], check=True)
    r = subprocess.run([exe, "2000000"], capture_output=True, text=True)
    n, false_reject, false_accept, accepts = map(int, r.stdout.split())
    assert r.returncode == 0 and false_reject == 0, r.stdout
    assert accepts > 0.02 * n                      # the sample does exercise the accept side
    assert false_accept <= 0.02 * accepts, r.stdout  # tight: rectangle-exact up to the FP32 margin
