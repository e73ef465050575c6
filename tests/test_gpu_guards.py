"""Memory-safety checks of the CUDA path without compute-sanitizer (the tool is closed on this GPU pool: "runs under it
have left GPUs needing a reset ... find a bad access with bounds checks and asserts of your own, small cases, and a
comparison with the CPU reference" -- profiles/sanitizer_r2.txt).  What memcheck / initcheck would look for is tested
directly through the C-ABI (the ctypes binding, which lets the test own every buffer):

  * out-of-bounds WRITES: every buffer a kernel writes (the three scratch buffers, the outputs, the backward scratch,
    the gradients) sits between two 4 KiB guard zones filled with a canary; the canaries must be intact after a forward
    + backward, at image sizes that are not multiples of the tile and Gaussian counts that are not multiples of any
    block size;
  * reads of UNINITIALISED scratch: the same frame with every scratch buffer pre-filled with 0xFF bytes (NaN / -1
    patterns) and with zeros must give bit-identical images, radii, inverse depth and equal gradients (the stale-shared-
    memory bug of commit b416666 was of this kind; its own regression test is test_gpu_full_size.py).
"""
import ctypes as C

import numpy as np
import pytest
import torch

import gpu_utils as U
from gaussian_splatting_lightning_b200 import _native as N
from gaussian_splatting_lightning_b200 import rasterizer as RZ

pytestmark = pytest.mark.gpu
GUARD = 4096
CANARY = 0xA5


class Guarded:
    """A device buffer between two canary zones."""

    def __init__(self, nbytes, fill=None, device=U.DEV):
        self.nbytes = int(nbytes)
        self.raw = torch.full((self.nbytes + 2 * GUARD,), CANARY, dtype=torch.uint8, device=device)
        self.inner = self.raw[GUARD:GUARD + self.nbytes]
        if fill is not None:
            self.inner.fill_(fill)

    def ptr(self):
        return self.inner.data_ptr()

    def intact(self):
        return bool((self.raw[:GUARD] == CANARY).all()) and bool((self.raw[GUARD + self.nbytes:] == CANARY).all())

    def view(self, dtype, shape):
        return self.inner.view(dtype).view(shape)


class GuardedAllocator(N.Allocator):
    def __init__(self, device, fill):
        super().__init__(device)
        self.fill, self.guarded, self.every = fill, {}, []   # every: also the buffers a retry replaced

    def _alloc(self, _user, which, nbytes):
        try:
            g = Guarded(max(int(nbytes), 1), self.fill, self.device)
            self.guarded[int(which)] = g
            self.every.append(g)
            self.buffers[int(which)] = g.inner
            return g.ptr()
        except BaseException as e:
            self.error = e
            return None


def _frame(cam, arrays, fill, hint=0):
    """forward + backward through the raw C-ABI with every written buffer guarded; returns results + guard status."""
    rs = U.settings_from_camera(cam)
    d = U.to_dev(arrays)
    fr = RZ._Frame(rs, d["means3D"], d.get("shs"), None, d["opacities"], d["scales"], d["rotations"], None)
    P, M, H, W = fr.P, fr.M, fr.H, fr.W
    L = N.lib()
    guards = dict(color=Guarded(3 * H * W * 4, fill), invdepth=Guarded(H * W * 4, fill), radii=Guarded(max(P, 1) * 4, fill))
    alloc = GuardedAllocator(fr.device, fill)
    R, cap = C.c_int64(0), C.c_int64(0)
    N.check(L.gsr_forward(C.byref(fr.settings), C.byref(fr.gaussians), guards["color"].ptr(), guards["radii"].ptr(),
                          guards["invdepth"].ptr(), alloc.cfunc, None, int(hint), C.byref(R), C.byref(cap),
                          N.stream_ptr(fr.device)), "gsr_forward")
    assert alloc.error is None
    torch.cuda.synchronize()
    shapes = dict(dL_dmeans2D=(P, 3), dL_dcolors=None, dL_dopacity=(P, 1), dL_dmeans3D=(P, 3), dL_dcov3D=None,
                  dL_dsh=(P, M, 3), dL_dscales=(P, 3), dL_drotations=(P, 4))
    gg = {n: (Guarded(int(np.prod(s)) * 4, fill) if s else None) for n, s in shapes.items()}
    scratch = Guarded(L.gsr_backward_scratch_bytes(P), fill)
    g = torch.Generator().manual_seed(1)
    dC = torch.randn(3, H, W, generator=g).to(fr.device)
    dD = torch.randn(1, H, W, generator=g).to(fr.device)
    grads = N.GsrGrads(*[(gg[n].ptr() if gg[n] is not None else None) for n in RZ._GRAD_NAMES])
    N.check(L.gsr_backward(C.byref(fr.settings), C.byref(fr.gaussians), guards["radii"].ptr(), dC.data_ptr(), dD.data_ptr(),
                           alloc.guarded[N.GSR_BUF_GEOM].ptr(), alloc.guarded[N.GSR_BUF_BINNING].ptr(),
                           alloc.guarded[N.GSR_BUF_IMG].ptr(), int(R.value), scratch.ptr(), C.byref(grads), 0, None,
                           N.stream_ptr(fr.device)), "gsr_backward")
    torch.cuda.synchronize()
    all_guards = list(guards.values()) + alloc.every + [x for x in gg.values() if x is not None] + [scratch]
    out = dict(R=int(R.value), color=guards["color"].view(torch.float32, (3, H, W)).clone(),
               invdepth=guards["invdepth"].view(torch.float32, (H, W)).clone(),
               radii=guards["radii"].view(torch.int32, (-1,))[:P].clone(),
               grads={n: gg[n].view(torch.float32, shapes[n]).clone() for n in gg if gg[n] is not None})
    return out, [x.intact() for x in all_guards]


CASES = [(7001, 3, 203, 117), (1, 0, 17, 9), (33, 2, 16, 16), (12_345, 1, 321, 200)]


@pytest.mark.parametrize("P,deg,W,H", CASES)
def test_no_write_outside_any_buffer_and_no_read_of_uninitialised_scratch(P, deg, W, H):
    cam, arrays = U.synthetic_case(P, deg, W, H, seed=P, bg=(0.1, 0.2, 0.3))
    a, ok_a = _frame(cam, arrays, fill=0xFF)     # scratch starts as NaN / 0xFFFFFFFF patterns
    b, ok_b = _frame(cam, arrays, fill=0x00)
    assert all(ok_a) and all(ok_b), "a kernel wrote outside one of its buffers"
    assert a["R"] == b["R"]
    assert torch.equal(a["color"], b["color"]) and torch.equal(a["invdepth"], b["invdepth"]) and torch.equal(a["radii"], b["radii"])
    assert bool(torch.isfinite(a["color"]).all())
    for n in a["grads"]:
        ga, gb = a["grads"][n], b["grads"][n]
        assert bool(torch.isfinite(ga).all()), n
        s = max(1.0, float(gb.abs().max()))
        assert float((ga - gb).abs().max()) / s <= 1e-5, n      # equal up to the order of the atomics


def test_guards_hold_on_the_binning_hint_retry_path():
    """A too-small hint makes gsr_forward run binning + compositing on a truncated list first (device-side clamp to the
    capacity): nothing may be written past the small buffer."""
    cam, arrays = U.synthetic_case(9000, 1, 200, 120, seed=3)
    base, ok = _frame(cam, arrays, fill=0xFF, hint=0)
    assert all(ok)
    for hint in (1, base["R"] // 3, base["R"] - 1):
        o, ok = _frame(cam, arrays, fill=0xFF, hint=hint)
        assert all(ok), hint
        assert o["R"] == base["R"] and torch.equal(o["color"], base["color"])
