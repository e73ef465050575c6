"""PLY checkpoints of the Gaussian set: the on-disk format on either side of the render path.

Mirrors GaussianModel.save_ply / load_model_ply (gs_lightning/modules/gaussian_model.py:112-171), which go through the
`plyfile` package: one `vertex` element of float32 properties, binary little-endian, in the order
    x y z  nx ny nz  f_dc_0..2  f_rest_0..3(M-1)-1  opacity  scale_0..2  rot_0..3          (gaussian_model.py:166-167)
with f_dc / f_rest written CHANNEL-major (`.transpose(1, 2)` before flattening, :156-157).

load_ply reproduces what the reference's loader returns, quirks included (gaussian_model.py:112-140): property
names are collected by prefix and sorted LEXICOGRAPHICALLY (so f_rest_10 sorts before f_rest_2), stacked in that order
and reshaped to (N, -1, 3) without undoing the channel-major transpose.  `canonical=True` returns instead the
tensors save_ply was given (numeric order, transpose undone) -- what a round trip should give.

The whole vertex table is read with one np.fromfile into pinned memory and moved with ONE host->device copy; the
de-interleave into the per-parameter tensors happens on the device.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2",
              "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}


def property_order(num_rest: int) -> List[str]:
    """Property names in file order for `num_rest` = 3 (M - 1) higher-order SH values (gaussian_model.py:146-167)."""
    return (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] +
            [f"f_rest_{i}" for i in range(num_rest)] + ["opacity"] + [f"scale_{i}" for i in range(3)] +
            [f"rot_{i}" for i in range(4)])


def save_ply(ply_path: str, xyz: torch.Tensor, features_dc: torch.Tensor, features_rest: torch.Tensor,
             opacity: torch.Tensor, scaling: torch.Tensor, rotation: torch.Tensor) -> bool:
    """GaussianModel.save_ply: xyz (N,3), features_dc (N,1,3), features_rest (N,M-1,3), opacity (N,1),
    scaling (N,3), rotation (N,4) -- the raw (pre-activation) parameters."""
    n = int(xyz.shape[0])
    cols = [xyz.detach().reshape(n, 3), torch.zeros(n, 3, dtype=xyz.dtype, device=xyz.device),
            features_dc.detach().transpose(1, 2).reshape(n, -1), features_rest.detach().transpose(1, 2).reshape(n, -1),
            opacity.detach().reshape(n, 1), scaling.detach().reshape(n, -1), rotation.detach().reshape(n, -1)]
    table = torch.cat([c.float() for c in cols], dim=1).contiguous()        # interleave on the device, one D2H copy
    names = property_order(int(cols[3].shape[1]))
    assert table.shape[1] == len(names)
    header = "ply\nformat binary_little_endian 1.0\n" + f"element vertex {n}\n" + \
             "".join(f"property float {p}\n" for p in names) + "end_header\n"
    with open(ply_path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(table.cpu().numpy().astype("<f4", copy=False).tobytes())
    return True


def read_vertex_table(ply_path: str) -> Tuple[np.ndarray, List[str]]:
    """(structured array of the `vertex` element, property names in file order) of a binary little-endian or
    ASCII PLY whose vertex element holds scalar properties only."""
    with open(ply_path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{ply_path}: not a PLY file")
        fmt, elements, current = None, [], None
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{ply_path}: header without end_header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                current = dict(name=tok[1], count=int(tok[2]), props=[])
                elements.append(current)
            elif tok[0] == "property":
                if tok[1] == "list":
                    if current["name"] == "vertex":
                        raise ValueError(f"{ply_path}: list property in the vertex element")
                    current["props"].append(("list", tok[-1]))
                else:
                    if tok[1] not in _PLY_TYPES:
                        raise ValueError(f"{ply_path}: unknown property type {tok[1]!r}")
                    current["props"].append((_PLY_TYPES[tok[1]], tok[2]))
            elif tok[0] == "end_header":
                break
        if not elements or elements[0]["name"] != "vertex":
            raise ValueError(f"{ply_path}: the first element must be `vertex`")
        v = elements[0]
        names = [p[1] for p in v["props"]]
        dtype = np.dtype([(nm, ty) for ty, nm in v["props"]])
        if fmt == "binary_little_endian":
            data = np.fromfile(f, dtype=dtype, count=v["count"])
        elif fmt == "ascii":
            flat = np.loadtxt(f, dtype=np.float64, max_rows=v["count"], ndmin=2)
            data = np.empty(v["count"], dtype=dtype)
            for i, nm in enumerate(names):
                data[nm] = flat[:, i]
        else:
            raise ValueError(f"{ply_path}: unsupported PLY format {fmt!r} (binary_little_endian and ascii are)")
        if data.shape[0] != v["count"]:
            raise ValueError(f"{ply_path}: expected {v['count']} vertices, file holds {data.shape[0]}")
    return data, names


def _numeric_suffix(name: str) -> int:
    return int(name.rsplit("_", 1)[1])


def load_ply(ply_path: str, device: Optional[torch.device] = None, canonical: bool = False) -> Dict[str, torch.Tensor]:
    """GaussianModel.load_model_ply: dict with xyz (N,3), features_dc (N,1,3), features_rest (N,M-1,3),
    scaling (N,3), rotation (N,4), opacity (N,1) float32 tensors on `device`, and active_sh_degree
    (= int(sqrt(features_rest.shape[-1] + 1)) exactly as gaussian_model.py:132 computes it)."""
    data, names = read_vertex_table(ply_path)
    n = data.shape[0]
    uniform = all(data.dtype[nm] == np.dtype("<f4") for nm in names)
    if uniform:   # the format save_ply writes: the table is one (N, K) float32 matrix -> one H2D copy
        host = torch.from_numpy(data.view("<f4").reshape(n, len(names)))
    else:
        host = torch.from_numpy(np.stack([data[nm].astype(np.float32) for nm in names], axis=1))
    if device is not None and torch.device(device).type == "cuda":
        host = host.pin_memory()
        table = host.to(device, non_blocking=True)
    else:
        table = host
    col = {nm: i for i, nm in enumerate(names)}

    def take(prefix: str) -> torch.Tensor:
        group = [nm for nm in names if nm.startswith(prefix)]
        group = sorted(group, key=_numeric_suffix) if (canonical and prefix != "opacity") else sorted(group)
        if not group:
            return table.new_zeros((n, 0))
        return table[:, [col[nm] for nm in group]]

    xyz = table[:, [col["x"], col["y"], col["z"]]].contiguous()
    f_dc, f_rest = take("f_dc"), take("f_rest")
    if canonical:   # undo the channel-major transpose of save_ply
        features_dc = f_dc.reshape(n, 3, 1).transpose(1, 2).contiguous()
        features_rest = f_rest.reshape(n, 3, -1).transpose(1, 2).contiguous()
        degree = int(round(np.sqrt(features_rest.shape[1] + 1))) - 1
    else:
        features_dc = f_dc.reshape(n, 1, 3).contiguous()
        features_rest = f_rest.reshape(n, -1, 3).contiguous()
        degree = int(np.sqrt(features_rest.shape[-1] + 1))
    return dict(xyz=xyz, features_dc=features_dc, features_rest=features_rest,
                scaling=take("scale").reshape(n, 3).contiguous(), rotation=take("rot").reshape(n, 4).contiguous(),
                opacity=take("opacity").reshape(n, 1).contiguous(), active_sh_degree=degree)
