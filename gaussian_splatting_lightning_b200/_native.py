"""ctypes binding of libgsr_b200.so -- the ONLY FFI layer between Python and the CUDA kernels.

Every function here passes raw device pointers (tensor.data_ptr()) and sizes through the
C-ABI declared in include/gsr_b200.h.  torch is used for device memory and streams only.
There is no CPU fallback: if the shared object is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgsr_b200.so")

GSR_BUF_GEOM, GSR_BUF_BINNING, GSR_BUF_IMG = 0, 1, 2

ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_int, C.c_size_t)


class GsrSettings(C.Structure):
    _fields_ = [("image_height", C.c_int32), ("image_width", C.c_int32), ("tanfovx", C.c_float),
                ("tanfovy", C.c_float), ("scale_modifier", C.c_float), ("sh_degree", C.c_int32),
                ("prefiltered", C.c_int32), ("debug", C.c_int32), ("antialiasing", C.c_int32),
                ("bg", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
                ("campos", C.c_void_p)]


class GsrGaussians(C.Structure):
    _fields_ = [("P", C.c_int32), ("M", C.c_int32), ("means3D", C.c_void_p), ("opacities", C.c_void_p),
                ("shs", C.c_void_p), ("colors_precomp", C.c_void_p), ("scales", C.c_void_p),
                ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p)]


class GsrGrads(C.Structure):
    _fields_ = [("dL_dmeans2D", C.c_void_p), ("dL_dcolors", C.c_void_p), ("dL_dopacity", C.c_void_p),
                ("dL_dmeans3D", C.c_void_p), ("dL_dcov3D", C.c_void_p), ("dL_dsh", C.c_void_p),
                ("dL_dscales", C.c_void_p), ("dL_drotations", C.c_void_p)]


class GsrDensifyStats(C.Structure):
    _fields_ = [("max_radii2D", C.c_void_p), ("xyz_grad_accum", C.c_void_p), ("xyz_grad_count", C.c_void_p)]


class GsrAdamGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("n", C.c_int64), ("lr", C.c_double)]


GSR_GRADS_ACCUMULATE = 1
GSR_BWD_RENDER_ONLY = 2


class GsrViewState(C.Structure):
    _fields_ = [("settings", C.POINTER(GsrSettings)), ("radii", C.c_void_p), ("geom", C.c_void_p), ("scratch", C.c_void_p)]

GSR_ADAM_MAX_GROUPS = 8
GSR_ADAM_ZERO_GRAD = 1

_EXPORT_FIELDS = ["depths", "means2D", "cov3D", "conic_opacity", "rgb", "clamped", "tiles_touched",
                  "point_offsets", "keys_unsorted", "vals_unsorted", "keys_sorted", "vals_sorted", "ranges",
                  "final_T", "n_contrib"]


class GsrStateExport(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _EXPORT_FIELDS]


# every symbol include/gsr_b200.h declares (tests check that the library exports them all)
EXPORTED_SYMBOLS = [
    "gsr_version", "gsr_last_error", "gsr_geom_bytes", "gsr_img_bytes", "gsr_binning_bytes",
    "gsr_backward_scratch_bytes", "gsr_forward", "gsr_backward", "gsr_preprocess_backward_views", "gsr_preprocess_backward_views_range", "gsr_mark_visible", "gsr_export_state",
    "gsr_sort_workspace_bytes", "gsr_sort_pairs", "gsr_launch_count", "gsr_set_binning", "gsr_get_binning", "gsr_set_priority_lanes", "gsr_get_priority_lanes", "gsr_profile_enable", "gsr_profile_reset",
    "gsr_profile_collect", "gsr_profile_stage_count", "gsr_profile_stage_name", "gsr_profile_stage_ms",
    "gsr_profile_stage_calls", "gsr_l1_loss_scratch_bytes", "gsr_l1_loss", "gsr_l1_loss_noalloc", "gsr_adam_step", "gsr_densify_fold", "gsr_fused_ssim_scratch_bytes",
    "gsr_fused_ssim", "gsr_photometric_loss",
]

_lib = None


class GsrError(RuntimeError):
    pass


def lib():
    """Load libgsr_b200.so; fail loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA extension has not been built. "
            "Run `python -m gaussian_splatting_lightning_b200.build` (needs nvcc). "
            "There is no CPU or PyTorch fallback for the rasterizer.")
    L = C.CDLL(LIB_PATH)
    L.gsr_version.restype = C.c_char_p
    L.gsr_last_error.restype = C.c_char_p
    L.gsr_launch_count.restype = C.c_int64
    for name in ("gsr_geom_bytes", "gsr_img_bytes", "gsr_binning_bytes", "gsr_backward_scratch_bytes",
                 "gsr_sort_workspace_bytes"):
        getattr(L, name).restype = C.c_size_t
    L.gsr_geom_bytes.argtypes = [C.c_int32]
    L.gsr_img_bytes.argtypes = [C.c_int32, C.c_int32]
    L.gsr_binning_bytes.argtypes = [C.c_int64, C.c_int32, C.c_int32]
    L.gsr_backward_scratch_bytes.argtypes = [C.c_int32]
    L.gsr_sort_workspace_bytes.argtypes = [C.c_int64]
    L.gsr_forward.restype = C.c_int
    L.gsr_forward.argtypes = [C.POINTER(GsrSettings), C.POINTER(GsrGaussians), C.c_void_p, C.c_void_p, C.c_void_p,
                              ALLOC_FN, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                              C.c_void_p]
    L.gsr_backward.restype = C.c_int
    L.gsr_backward.argtypes = [C.POINTER(GsrSettings), C.POINTER(GsrGaussians), C.c_void_p, C.c_void_p, C.c_void_p,
                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(GsrGrads),
                               C.c_int32, C.POINTER(GsrDensifyStats), C.c_void_p]
    L.gsr_preprocess_backward_views.restype = C.c_int
    L.gsr_preprocess_backward_views.argtypes = [C.POINTER(GsrViewState), C.c_int32, C.POINTER(GsrGaussians),
                                                C.POINTER(GsrGrads), C.c_int32, C.POINTER(GsrDensifyStats), C.c_void_p]
    L.gsr_preprocess_backward_views_range.restype = C.c_int
    L.gsr_preprocess_backward_views_range.argtypes = [C.POINTER(GsrViewState), C.c_int32, C.POINTER(GsrGaussians),
                                                      C.POINTER(GsrGrads), C.c_int32, C.POINTER(GsrDensifyStats),
                                                      C.c_int32, C.c_int32, C.c_void_p]
    L.gsr_mark_visible.restype = C.c_int
    L.gsr_mark_visible.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gsr_export_state.restype = C.c_int
    L.gsr_export_state.argtypes = [C.POINTER(GsrSettings), C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_int64, C.c_int64, C.POINTER(GsrStateExport), C.c_void_p]
    L.gsr_sort_pairs.restype = C.c_int
    L.gsr_sort_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                 C.c_int32, C.c_void_p, C.c_void_p]
    L.gsr_l1_loss_scratch_bytes.restype = C.c_size_t
    L.gsr_l1_loss_scratch_bytes.argtypes = []
    L.gsr_l1_loss.restype = C.c_int
    L.gsr_l1_loss.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gsr_l1_loss_noalloc.restype = C.c_int
    L.gsr_l1_loss_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gsr_fused_ssim_scratch_bytes.restype = C.c_size_t
    L.gsr_fused_ssim_scratch_bytes.argtypes = [C.c_int32, C.c_int32, C.c_int32]
    L.gsr_fused_ssim.restype = C.c_int
    L.gsr_fused_ssim.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_float,
                                 C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gsr_photometric_loss.restype = C.c_int
    L.gsr_photometric_loss.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gsr_adam_step.restype = C.c_int
    L.gsr_adam_step.argtypes = [C.POINTER(GsrAdamGroup), C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double,
                                C.c_int32, C.c_void_p]
    L.gsr_densify_fold.restype = C.c_int
    L.gsr_densify_fold.argtypes = [C.c_int32] + [C.c_void_p] * 7
    L.gsr_profile_stage_name.restype = C.c_char_p
    L.gsr_profile_stage_ms.restype = C.c_double
    L.gsr_profile_stage_calls.restype = C.c_int64
    L.gsr_profile_reset.restype = None
    _lib = L
    return L


EXT_PATH = os.path.join(_HERE, "_C.so")
_ext = None
_ext_tried = False


def ext():
    """`_C`, the thin torch C++ extension over the C-ABI (csrc/torch_ext.cpp) -- the FFI layer of the rasterizer calls.
    GSR_FFI=ctypes selects the ctypes binding of the same C-ABI instead (kept as the second binding; the tests run the
    parity cases through both).  Returns None when ctypes is selected or `_C.so` has not been built."""
    global _ext, _ext_tried
    if os.environ.get("GSR_FFI", "") == "ctypes":
        return None
    if _ext_tried:
        return _ext
    _ext_tried = True
    lib()   # fail loudly first if the CUDA library itself is missing
    if os.path.exists(EXT_PATH):
        import importlib.util
        spec = importlib.util.spec_from_file_location(__package__ + "._C", EXT_PATH)
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        import sys
        sys.modules[__package__ + "._C"] = m
        _ext = m
    else:
        import warnings
        warnings.warn(f"{EXT_PATH} has not been built (python -m gaussian_splatting_lightning_b200.build); "
                      "using the ctypes binding of libgsr_b200.so")
    return _ext


def set_binning(mode: str) -> None:
    """'tile' (default: tile-segmented binning) or 'onesweep' (global radix sort); process-wide, at run time."""
    check(lib().gsr_set_binning({"tile": 0, "onesweep": 1}[mode]), "gsr_set_binning")


def get_binning() -> str:
    return ("tile", "onesweep")[int(lib().gsr_get_binning())]


def set_priority_lanes(on: bool) -> None:
    """Run the stages before compositing on a high-priority stream per caller stream (include/gsr_b200.h)."""
    check(lib().gsr_set_priority_lanes(int(bool(on))), "gsr_set_priority_lanes")


def get_priority_lanes() -> bool:
    return bool(lib().gsr_get_priority_lanes())


def profile_enable(on: bool) -> None:
    lib().gsr_profile_enable(int(on))


def profile_reset() -> None:
    lib().gsr_profile_reset()


def profile_read() -> dict:
    """{stage: (total_ms, calls)} after waiting for the recorded events."""
    L = lib()
    check(L.gsr_profile_collect(), "gsr_profile_collect")
    return {L.gsr_profile_stage_name(i).decode(): (float(L.gsr_profile_stage_ms(i)), int(L.gsr_profile_stage_calls(i)))
            for i in range(L.gsr_profile_stage_count())}


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().gsr_last_error().decode("utf-8", "replace")
        raise GsrError(f"{what} failed (code {rc}): {msg}")


def version() -> str:
    return lib().gsr_version().decode()


def launch_count() -> int:
    return int(lib().gsr_launch_count())


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def f32c(t: Optional[torch.Tensor], device=None) -> Optional[torch.Tensor]:
    """contiguous float32 (on `device` if given); None and empty tensors become None."""
    if t is None or t.numel() == 0:
        return None
    if device is not None and t.device != device:
        t = t.to(device)
    if t.dtype != torch.float32:
        t = t.float()
    t = t.contiguous()
    if t.data_ptr() % 16:
        t = t.clone(memory_format=torch.contiguous_format)
    return t


def stream_ptr(device) -> int:
    return int(torch.cuda.current_stream(device).cuda_stream)


class Allocator:
    """The resize callback of gsr_forward (notes/rasterizer_note.h:27-40): torch owns the memory."""

    def __init__(self, device):
        self.device = device
        self.buffers: Dict[int, torch.Tensor] = {}
        self.error: Optional[BaseException] = None
        self.cfunc = ALLOC_FN(self._alloc)

    def _alloc(self, _user, which, nbytes):
        try:
            t = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=self.device)
            self.buffers[int(which)] = t
            return t.data_ptr()
        except BaseException as e:  # never let an exception cross the C boundary
            self.error = e
            return None
