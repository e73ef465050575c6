"""SURVEY.md 8(f) rows 3-4 on the CPU: PLY checkpoints (gaussian_model.py:112-171) and the camera contract
(colmap_dataset.py:77-109, utils/camera.py:4-41, golden values produced by the reference's own function:
tests/golden/make_camera_golden.py)."""
import os

import numpy as np
import pytest
import torch

from gaussian_splatting_lightning_b200 import cameras, io

GOLD = os.path.join(os.path.dirname(__file__), "golden", "cameras.npz")


def _model(n=57, M=16, seed=0):
    g = torch.Generator().manual_seed(seed)
    return dict(xyz=torch.randn(n, 3, generator=g), features_dc=torch.randn(n, 1, 3, generator=g),
                features_rest=torch.randn(n, M - 1, 3, generator=g), opacity=torch.randn(n, 1, generator=g),
                scaling=torch.randn(n, 3, generator=g), rotation=torch.randn(n, 4, generator=g))


def _reference_load(vertices, names):
    """GaussianModel.load_model_ply / load_vertices_properties (gaussian_model.py:112-140) over a structured array
    standing in for plyfile's PlyElement (vertices[name] -> column)."""
    def props(prefix):
        group = sorted(p for p in names if p.startswith(prefix))
        return torch.Tensor(np.stack([vertices[p] for p in group], -1))
    xyz = torch.Tensor(np.stack([vertices["x"], vertices["y"], vertices["z"]], -1))
    N = len(xyz)
    return dict(xyz=xyz, features_dc=props("f_dc").reshape(N, 1, 3), features_rest=props("f_rest").reshape(N, -1, 3),
                scaling=props("scale").reshape(N, 3), rotation=props("rot").reshape(N, 4),
                opacity=props("opacity").reshape(N, 1))


@pytest.mark.parametrize("M", [1, 4, 16])
def test_ply_file_layout_and_reference_loader_semantics(tmp_path, M):
    m = _model(M=M)
    path = str(tmp_path / "pc.ply")
    assert io.save_ply(path, **m)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().split("\n")
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 57"]
    names = [l.split()[2] for l in lines if l.startswith("property")]
    assert all(l.startswith("property float ") for l in lines[3:] if l)
    assert names == io.property_order(3 * (M - 1)) and len(body) == 57 * 4 * len(names)
    table = np.frombuffer(body, "<f4").reshape(57, -1)
    np.testing.assert_array_equal(table[:, 0:3], m["xyz"].numpy())
    assert not table[:, 3:6].any()                                                     # normals are zeros
    np.testing.assert_array_equal(table[:, 6:9], m["features_dc"].transpose(1, 2).reshape(57, -1).numpy())
    np.testing.assert_array_equal(table[:, 9:9 + 3 * (M - 1)], m["features_rest"].transpose(1, 2).reshape(57, -1).numpy())
    # the loader reproduces the reference's (lexicographic, untransposed) reading of that file
    vertices, vnames = io.read_vertex_table(path)
    got = io.load_ply(path)
    if M == 1:   # no f_rest properties: the reference's loader fails (np.stack of nothing); ours returns (N, 0, 3)
        with pytest.raises(ValueError):
            _reference_load(vertices, vnames)
        assert got["features_rest"].shape == (57, 0, 3) and torch.equal(got["xyz"], m["xyz"])
        return
    want = _reference_load(vertices, vnames)
    for k, w in want.items():
        assert got[k].shape == w.shape and torch.equal(got[k], w), k
    assert got["active_sh_degree"] == int(np.sqrt(want["features_rest"].shape[-1] + 1))
    # canonical=True is the true inverse of save_ply
    back = io.load_ply(path, canonical=True)
    for k, w in m.items():
        assert torch.equal(back[k], w), k
    assert back["active_sh_degree"] == int(round(np.sqrt(M))) - 1


def test_ply_ascii_mixed_types_and_errors(tmp_path):
    p = tmp_path / "a.ply"
    p.write_text("ply\nformat ascii 1.0\ncomment x\nelement vertex 2\nproperty float x\nproperty float y\nproperty float z\n"
                 "property double opacity\nproperty float scale_0\nproperty float scale_1\nproperty float scale_2\n"
                 "property float rot_0\nproperty float rot_1\nproperty float rot_2\nproperty float rot_3\n"
                 "property float f_dc_0\nproperty float f_dc_1\nproperty float f_dc_2\nend_header\n"
                 "1 2 3 0.5 1 1 1 1 0 0 0 .1 .2 .3\n4 5 6 0.25 2 2 2 0 1 0 0 .4 .5 .6\n")
    d = io.load_ply(str(p))
    assert d["xyz"].tolist() == [[1, 2, 3], [4, 5, 6]] and d["opacity"].tolist() == [[0.5], [0.25]]
    assert d["features_rest"].shape == (2, 0, 3) and d["features_dc"].shape == (2, 1, 3)
    bad = tmp_path / "b.ply"
    bad.write_bytes(b"plx\n")
    with pytest.raises(ValueError):
        io.load_ply(str(bad))
    trunc = tmp_path / "c.ply"
    trunc.write_bytes(b"ply\nformat binary_little_endian 1.0\nelement vertex 3\nproperty float x\nend_header\n" + b"\0" * 8)
    with pytest.raises(ValueError):
        io.read_vertex_table(str(trunc))


def test_camera_contract_matches_reference_golden():
    gold = np.load(GOLD)
    for i in range(6):
        v = gold[f"in_{i}"]
        c2w, (fx, fy, W, H, zn, zf) = v[:12].reshape(3, 4), v[12:]
        np.testing.assert_array_equal(cameras.get_projection_matrix(fx, fy, int(W), int(H), zn, zf), gold[f"proj_{i}"])
        cam = cameras.camera_from_colmap(c2w, fx, fy, int(W), int(H), zn, zf)
        np.testing.assert_array_equal(cam["viewmatrix"].numpy(), gold[f"view_{i}"])
        np.testing.assert_array_equal(cam["projmatrix"].numpy(), gold[f"full_{i}"])
        np.testing.assert_array_equal(cam["campos"].numpy(), gold[f"campos_{i}"])
        assert [cam["tanfovx"], cam["tanfovy"]] == gold[f"tan_{i}"].tolist()


def test_settings_from_collated_batch():
    from torch.utils.data import default_collate
    gold = np.load(GOLD)
    items = []
    for i in range(4):
        v = gold[f"in_{i}"]
        cam = cameras.camera_from_colmap(v[:12].reshape(3, 4), v[12], v[13], 640, 480)
        items.append(dict(image=torch.zeros(3, 480, 640), background=torch.tensor([0.0, 0.0, 0.0]), **cam))
    one = cameras.settings_from_item(items[2], sh_degree=3)
    assert (one.image_height, one.image_width, one.sh_degree, one.scale_modifier) == (480, 640, 3, 1.0)
    assert one.viewmatrix is items[2]["viewmatrix"] and one.prefiltered is False and one.antialiasing is False
    sets = cameras.settings_from_batch(default_collate(items), sh_degree=3)
    assert len(sets) == 4
    for s, it in zip(sets, items):
        assert torch.equal(s.viewmatrix, it["viewmatrix"]) and torch.equal(s.projmatrix, it["projmatrix"])
        assert torch.equal(s.campos, it["campos"]) and s.tanfovx == it["tanfovx"] and s.tanfovy == it["tanfovy"]
        assert (s.image_height, s.image_width) == (480, 640) and s.bg.tolist() == [0, 0, 0]


def test_save_ply_bytes_spelled_out(tmp_path):
    """f3 pin without `plyfile` (absent from the image): the exact bytes GaussianModel.save_ply writes
    (gs_lightning/modules/gaussian_model.py:150-171) for two Gaussians of SH degree 1, spelled out by hand from the
    format: plyfile's header for a structured array of 'f4' fields (`property float <name>`, binary_little_endian 1.0,
    one element `vertex`), property order x y z nx ny nz f_dc_* f_rest_* opacity scale_* rot_*, normals zero, and the
    features written CHANNEL-major (`features.transpose(1, 2)` then flatten, gaussian_model.py:157-158)."""
    import struct
    import torch
    from gaussian_splatting_lightning_b200 import io
    xyz = torch.tensor([[1.0, 2.0, 3.0], [-4.0, 5.5, 0.25]])
    f_dc = torch.tensor([[[0.1, 0.2, 0.3]], [[0.4, 0.5, 0.6]]])                       # (N, 1, 3): [coef][channel]
    f_rest = torch.arange(18, dtype=torch.float32).reshape(2, 3, 3) + 10.0             # (N, 3, 3): [coef][channel]
    opacity = torch.tensor([[-1.5], [2.5]])
    scaling = torch.tensor([[-3.0, -2.0, -1.0], [0.5, 0.25, 0.125]])
    rotation = torch.tensor([[1.0, 0.0, 0.0, 0.0], [0.5, -0.5, 0.5, -0.5]])
    path = str(tmp_path / "two.ply")
    io.save_ply(path, xyz, f_dc, f_rest, opacity, scaling, rotation)
    names = (["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + [f"f_rest_{i}" for i in range(9)]
             + ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"])
    header = ("ply\nformat binary_little_endian 1.0\nelement vertex 2\n"
              + "".join(f"property float {n}\n" for n in names) + "end_header\n").encode("ascii")
    f32 = lambda *v: struct.pack("<%df" % len(v), *v)
    row0 = (f32(1.0, 2.0, 3.0) + f32(0.0, 0.0, 0.0) + f32(0.1, 0.2, 0.3)
            # f_rest of Gaussian 0 is [[10,11,12],[13,14,15],[16,17,18]] ([coef][channel]); channel-major: R of the three
            # coefficients, then G, then B
            + f32(10.0, 13.0, 16.0, 11.0, 14.0, 17.0, 12.0, 15.0, 18.0)
            + f32(-1.5) + f32(-3.0, -2.0, -1.0) + f32(1.0, 0.0, 0.0, 0.0))
    row1 = (f32(-4.0, 5.5, 0.25) + f32(0.0, 0.0, 0.0) + f32(0.4, 0.5, 0.6)
            + f32(19.0, 22.0, 25.0, 20.0, 23.0, 26.0, 21.0, 24.0, 27.0)
            + f32(2.5) + f32(0.5, 0.25, 0.125) + f32(0.5, -0.5, 0.5, -0.5))
    assert open(path, "rb").read() == header + row0 + row1
    # and the reference loader's reading of those bytes (lexicographic property order, no transpose back:
    # gaussian_model.py:112-147) is what load_ply returns by default
    ck = io.load_ply(path)
    assert ck["xyz"].tolist() == xyz.tolist() and ck["active_sh_degree"] == int(np.sqrt(3 + 1))
    assert ck["features_rest"].reshape(2, -1)[0].tolist() == [10.0, 13.0, 16.0, 11.0, 14.0, 17.0, 12.0, 15.0, 18.0]
