"""ctypes front end of the CPU oracle (oracle/gsr_oracle.cpp).

TEST INFRASTRUCTURE ONLY.  Importable from tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py; the product package
(gaussian_splatting_lightning_b200) must never import this module.

Everything is numpy in / numpy out so the oracle carries no torch dependency.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libgsr_oracle.so")
_lib = None

TILE = 16


def build(force: bool = False) -> str:
    """Compile oracle/gsr_oracle.cpp -> oracle/libgsr_oracle.so (g++, OpenMP)."""
    src = os.path.join(_HERE, "gsr_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libgsr_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.gsro_scan.restype = C.c_int64
        _lib.gsro_num_threads.restype = C.c_int
    return _lib


def num_threads() -> int:
    return int(lib().gsro_num_threads())


def set_num_threads(n: int) -> None:
    lib().gsro_set_num_threads(C.c_int(n))


def _f32(a, shape=None):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float32))
    if shape is not None:
        a = a.reshape(shape)
    return a


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


@dataclass
class Camera:
    """One view: the 13 GaussianRasterizationSettings fields the kernels read."""
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    viewmatrix: np.ndarray  # (4,4) row-vector convention, translation in the last row
    projmatrix: np.ndarray  # (4,4) view @ projection, same convention
    campos: np.ndarray      # (3,)
    bg: np.ndarray = field(default_factory=lambda: np.zeros(3, np.float32))
    scale_modifier: float = 1.0
    sh_degree: int = 3
    antialiasing: bool = False

    def __post_init__(self):
        self.viewmatrix = _f32(self.viewmatrix, (16,))
        self.projmatrix = _f32(self.projmatrix, (16,))
        self.campos = _f32(self.campos, (3,))
        self.bg = _f32(self.bg, (3,))
        self.tanfovx = float(self.tanfovx)
        self.tanfovy = float(self.tanfovy)

    @property
    def grid(self):
        return ((self.image_width + TILE - 1) // TILE, (self.image_height + TILE - 1) // TILE)


def mark_visible(means3D, viewmatrix) -> np.ndarray:
    means3D = _f32(means3D)
    P = means3D.shape[0]
    out = np.zeros(P, np.uint8)
    lib().gsro_mark_visible(C.c_int(P), _p(means3D), _p(_f32(viewmatrix, (16,))), _p(out))
    return out.astype(bool)


def forward(cam: Camera, means3D, scales, rotations, opacities, shs=None, colors_precomp=None,
            cov3D_precomp=None, stop_after: Optional[str] = None) -> dict:
    """Full forward with every intermediate exposed (SURVEY.md Appendix C layout).

    stop_after in {None, "preprocess", "sort"} cuts the pipeline early.
    """
    L = lib()
    means3D = _f32(means3D)
    P = means3D.shape[0]
    opacities = _f32(opacities, (P,))
    scales = None if scales is None else _f32(scales, (P, 3))
    rotations = None if rotations is None else _f32(rotations, (P, 4))
    cov3D_precomp = None if cov3D_precomp is None else _f32(cov3D_precomp, (P, 6))
    colors_precomp = None if colors_precomp is None else _f32(colors_precomp, (P, 3))
    if shs is not None:
        shs = _f32(shs)
        M = shs.shape[1]
        assert shs.shape == (P, M, 3)
        assert M >= (cam.sh_degree + 1) ** 2
    else:
        M = 0
    W, H = cam.image_width, cam.image_height
    gx, gy = cam.grid

    st = dict(P=P, M=M, cam=cam)
    st["inputs"] = dict(means3D=means3D, scales=scales, rotations=rotations, opacities=opacities,
                        shs=shs, colors_precomp=colors_precomp, cov3D_precomp=cov3D_precomp)
    depths = np.zeros(P, np.float32)
    radii = np.zeros(P, np.int32)
    means2D = np.zeros((P, 2), np.float32)
    cov3D = np.zeros((P, 6), np.float32)
    conic_opacity = np.zeros((P, 4), np.float32)
    rgb = np.zeros((P, 3), np.float32)
    clamped = np.zeros((P, 3), np.uint8)
    tiles_touched = np.zeros(P, np.uint32)
    rects = np.zeros((P, 4), np.int32)
    L.gsro_preprocess(C.c_int(P), C.c_int(M), C.c_int(cam.sh_degree), _p(means3D), _p(scales),
                      _p(rotations), _p(opacities), _p(shs), _p(colors_precomp), _p(cov3D_precomp),
                      _p(cam.viewmatrix), _p(cam.projmatrix), _p(cam.campos), C.c_int(W), C.c_int(H),
                      C.c_float(cam.tanfovx), C.c_float(cam.tanfovy), C.c_float(cam.scale_modifier),
                      C.c_int(int(cam.antialiasing)), _p(depths), _p(radii), _p(means2D), _p(cov3D),
                      _p(conic_opacity), _p(rgb), _p(clamped), _p(tiles_touched), _p(rects))
    st.update(depths=depths, radii=radii, means2D=means2D, cov3D=cov3D, conic_opacity=conic_opacity,
              rgb=rgb, clamped=clamped, tiles_touched=tiles_touched, rects=rects)
    if stop_after == "preprocess":
        return st

    offsets = np.zeros(P, np.uint32)
    R = int(L.gsro_scan(C.c_int(P), _p(tiles_touched), _p(offsets)))
    keys_unsorted = np.zeros(R, np.uint64)
    vals_unsorted = np.zeros(R, np.uint32)
    L.gsro_duplicate(C.c_int(P), C.c_int(W), _p(depths), _p(rects), _p(tiles_touched), _p(offsets),
                     _p(keys_unsorted), _p(vals_unsorted))
    keys = np.zeros(R, np.uint64)
    vals = np.zeros(R, np.uint32)
    L.gsro_sort_pairs(C.c_int64(R), _p(keys_unsorted), _p(vals_unsorted), _p(keys), _p(vals))
    ranges = np.zeros((gx * gy, 2), np.uint32)
    L.gsro_tile_ranges(C.c_int64(R), _p(keys), C.c_int(gx * gy), _p(ranges))
    st.update(point_offsets=offsets, num_rendered=R, keys_unsorted=keys_unsorted,
              vals_unsorted=vals_unsorted, keys=keys, point_list=vals, ranges=ranges)
    if stop_after == "sort":
        return st

    color = np.zeros((3, H, W), np.float32)
    invdepth = np.zeros((1, H, W), np.float32)
    final_T = np.zeros((H, W), np.float32)
    n_contrib = np.zeros((H, W), np.uint32)
    L.gsro_render_forward(C.c_int(W), C.c_int(H), _p(ranges), _p(vals), _p(means2D),
                          _p(conic_opacity), _p(rgb), _p(depths), _p(cam.bg), _p(color),
                          _p(invdepth), _p(final_T), _p(n_contrib))
    st.update(color=color, invdepth=invdepth, final_T=final_T, n_contrib=n_contrib)
    return st


def backward(st: dict, dL_dcolor, dL_dinvdepth=None) -> dict:
    """Analytic backward from a forward() state."""
    L = lib()
    cam: Camera = st["cam"]
    P, M = st["P"], st["M"]
    W, H = cam.image_width, cam.image_height
    inp = st["inputs"]
    dL_dcolor = _f32(dL_dcolor, (3, H, W))
    dL_dinvdepth = None if dL_dinvdepth is None else _f32(dL_dinvdepth, (H, W))

    dL_dmean2D = np.zeros((P, 3), np.float32)
    dL_dconic = np.zeros((P, 3), np.float32)
    dL_dopacity = np.zeros(P, np.float32)
    dL_dcolors = np.zeros((P, 3), np.float32)
    dL_dinvz = np.zeros(P, np.float32)
    L.gsro_render_backward(C.c_int(P), C.c_int(W), C.c_int(H), _p(st["ranges"]), _p(st["point_list"]),
                           _p(st["means2D"]), _p(st["conic_opacity"]), _p(st["rgb"]), _p(st["depths"]),
                           _p(cam.bg), _p(st["final_T"]), _p(st["n_contrib"]), _p(dL_dcolor),
                           _p(dL_dinvdepth), _p(dL_dmean2D), _p(dL_dconic), _p(dL_dopacity),
                           _p(dL_dcolors), _p(dL_dinvz))
    g = dict(dL_dmeans2D=dL_dmean2D, dL_dconic=dL_dconic.copy(), dL_dcolors=dL_dcolors,
             dL_dinvz=dL_dinvz, dL_dopacity_render=dL_dopacity.copy())

    dL_dmeans3D = np.zeros((P, 3), np.float32)
    dL_dcov3D = np.zeros((P, 6), np.float32)
    dL_dsh = np.zeros((P, M, 3), np.float32) if M > 0 else None
    dL_dscales = np.zeros((P, 3), np.float32)
    dL_drot = np.zeros((P, 4), np.float32)
    L.gsro_preprocess_backward(
        C.c_int(P), C.c_int(M), C.c_int(cam.sh_degree), _p(inp["means3D"]), _p(inp["scales"]),
        _p(inp["rotations"]), _p(inp["opacities"]), _p(inp["shs"]), _p(inp["colors_precomp"]),
        _p(inp["cov3D_precomp"]), _p(cam.viewmatrix), _p(cam.projmatrix), _p(cam.campos),
        C.c_int(W), C.c_int(H), C.c_float(cam.tanfovx), C.c_float(cam.tanfovy),
        C.c_float(cam.scale_modifier), C.c_int(int(cam.antialiasing)), _p(st["radii"]),
        _p(st["cov3D"]), _p(st["clamped"]), _p(dL_dmean2D), _p(dL_dconic), _p(dL_dopacity),
        _p(dL_dcolors), _p(dL_dinvz), _p(dL_dmeans3D), _p(dL_dcov3D), _p(dL_dsh), _p(dL_dscales),
        _p(dL_drot))
    g.update(dL_dmeans3D=dL_dmeans3D, dL_dcov3D=dL_dcov3D, dL_dsh=dL_dsh, dL_dscales=dL_dscales,
             dL_drotations=dL_drot, dL_dopacity=dL_dopacity.reshape(P, 1))
    return g
