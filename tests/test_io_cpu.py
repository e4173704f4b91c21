"""Mesh / texture readers of the renderer classes (render_py_multi.py:66-82 reads
`<model_dir>/<class>/textured.obj` + `texture_map.png` through glumpy's objload): pure host code."""
import numpy as np

import helpers  # noqa: F401  (path setup)


def _write_obj(path, with_uv=True, with_n=True):
    lines = ["# unit square as one quad + a triangle fan on top",
             "v 0 0 0", "v 1 0 0", "v 1 1 0", "v 0 1 0", "v 0.5 1.5 0.25"]
    if with_uv:
        lines += ["vt 0 0", "vt 1 0", "vt 1 1", "vt 0 1", "vt 0.5 0.9"]
    if with_n:
        lines += ["vn 0 0 1", "vn 0 0.6 0.8"]
    if with_uv and with_n:
        lines += ["f 1/1/1 2/2/1 3/3/1 4/4/1", "f 4/4/2 3/3/2 -1/-1/-1"]
    elif with_uv:
        lines += ["f 1/1 2/2 3/3 4/4", "f 4/4 3/3 5/5"]
    else:
        lines += ["f 1 2 3 4", "f 4 3 5"]
    path.write_text("\n".join(lines) + "\n")


def test_load_obj_indexing_and_fan_triangulation(tmp_path):
    from mx_deepim_b200.render import load_obj

    p = tmp_path / "a.obj"
    _write_obj(p)
    m = load_obj(str(p))
    # quad -> 2 triangles (fan), + 1 triangle; vertices de-duplicated per (v, vt, vn) triple:
    # 4 with normal 1, then v4/v3 again with normal 2, + the apex (negative indices)
    assert m["triangles"].shape == (3, 3) and m["triangles"].dtype == np.int32
    assert m["vertices"].shape == (7, 3) and m["uv"].shape == (7, 2) and m["normals"].shape == (7, 3)
    assert m["triangles"].tolist()[:2] == [[0, 1, 2], [0, 2, 3]]
    apex = m["triangles"][2, 2]
    assert np.allclose(m["vertices"][apex], [0.5, 1.5, 0.25]) and np.allclose(m["uv"][apex], [0.5, 0.9])
    assert np.allclose(m["normals"][apex], [0, 0.6, 0.8])


def test_load_obj_rescale_and_missing_attributes(tmp_path):
    from mx_deepim_b200.render import load_obj

    p = tmp_path / "b.obj"
    _write_obj(p, with_uv=False, with_n=False)
    m = load_obj(str(p), rescale=True)  # objload(rescale=True): centred, largest extent 1
    assert m["uv"] is None
    ext = m["vertices"].max(0) - m["vertices"].min(0)
    assert abs(ext.max() - 1.0) < 1e-6 and np.allclose(m["vertices"].max(0) + m["vertices"].min(0), 0, atol=1e-6)
    n = m["normals"]  # computed when the file has none
    assert n.shape == m["vertices"].shape and np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-5)


def test_load_texture_is_rgb_top_row_first(tmp_path):
    import cv2

    from mx_deepim_b200.render import load_texture

    img = np.zeros((4, 6, 3), np.uint8)
    img[0, :, 2] = 255   # top row red (BGR order for cv2.imwrite)
    img[3, :, 0] = 200   # bottom row blue
    cv2.imwrite(str(tmp_path / "t.png"), img)
    t = load_texture(str(tmp_path / "t.png"))
    assert t.shape == (4, 6, 3) and t.dtype == np.uint8 and t.flags["C_CONTIGUOUS"]
    assert t[0, 0].tolist() == [255, 0, 0] and t[3, 0].tolist() == [0, 0, 200]


def test_dataset_readers_vs_reference_loaders(tmp_path):
    """mx_deepim_b200/dataio.py readers (storage dtypes, no decode on the host) against the
    outputs of the reference's own loaders on the same files
    (tests/golden/make_golden_dataio.py -> dataio_reference.npz)."""
    import os

    import torch

    from mx_deepim_b200 import dataio

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "dataio_reference.npz"))
    pairdb = helpers.write_pair_files(tmp_path, 41)
    assert np.array_equal(dataio.load_object_points(str(tmp_path / "model" / "points.xyz")), g["points"])
    recs = [dataio.read_pair(r) for r in pairdb]
    assert np.array_equal(np.stack([r["pose_observed"] for r in recs]), g["poses"].astype(np.float32))
    batch = dataio.stack_pairs(recs)
    assert batch["image_observed"].dtype == torch.uint8 and batch["image_observed"].shape == (2, 480, 640, 3)
    assert batch["depth_rendered"].dtype == torch.int16 and batch["mask_gt_observed"].dtype == torch.uint8
    assert batch["mask_idx"].tolist() == [1, 2]
    S = 4
    # the raw values are what the reference decodes: redo its float formulas on the host here
    raw = batch["depth_rendered"].numpy().view(np.uint16)
    assert np.array_equal((raw.astype(np.float32) / 1000)[:, None, ::S, ::S], g["depth_rendered"])
    raw = batch["depth_observed"].numpy().view(np.uint16).astype(np.float32)
    lab = batch["mask_gt_observed"].numpy()
    raw *= lab == np.array([1, 2]).reshape(2, 1, 1)  # image.py:244
    assert np.array_equal((raw / 1000)[:, None, ::S, ::S], g["depth_observed"])
    im = batch["image_observed"].numpy()
    means = np.array([102.9801, 115.9465, 122.7717])
    ref = np.stack([im[:, :, :, 2 - i] - means[2 - i] for i in range(3)], axis=1)  # image.py:717-720
    assert np.array_equal(ref[..., ::S, ::S].astype(np.float32), g["image_observed"])
    import pytest

    with pytest.raises(AssertionError):
        dataio.read_depth(str(tmp_path / "missing.png"))
    with pytest.raises(RuntimeError):
        dataio.decode_batch(batch, "cpu")  # no CPU fallback
