"""CUDA rasteriser vs the CPU restatement of the GL pipeline at BASELINE config 5's mesh density:
20 480-triangle meshes (synth.make_mesh(level=5)), where a triangle is a few pixels at working
distance and micro-triangle culling (raster.cu: triangles whose pixel box is <= 3x3 and that
contain no pixel centre never reach the tile lists) fires on ~40 % of the records.

Poses cover the working range (z 0.6-1.2 m), close-ups (z 0.3-0.6 m: triangles grow, the object
leaves the image), the far end (z 2-5 m: EVERY triangle is sub-pixel, so the picture is made of
whatever survives the micro-triangle test) and z ~ zNear (the near plane cuts the mesh).

Bars (north_star): depth within 1e-5 relative, >= 99.9 % per-pixel mask agreement; the winning
triangle must agree on >= 99.9 % of the pixels.  All three shading modes: textured (a2), Lambert /
ModelNet (a3), lit LINEMOD (f3).
Reference: lib/render_glumpy/render_py_multi.py:112-147,
lib/render_glumpy/render_py_light_modelnet_multi.py:149-231, render_py_light.py:136-213."""
import numpy as np
import pytest

import helpers
from oracle import raster_np

pytestmark = pytest.mark.gpu
K = helpers.K_LM


@pytest.fixture(scope="module")
def dense_meshes():
    from mx_deepim_b200 import synth

    rng = np.random.default_rng(20181804 + 5)  # bench.py's cfg5 mesh generator
    meshes = [synth.make_mesh(rng, level=5) for _ in range(3)]
    assert all(len(m["triangles"]) == 20480 for m in meshes)
    return meshes


def _pose_set():
    """(R, t) x 10: working range, close-ups, far, near-plane straddling, half off-screen."""
    from mx_deepim_b200 import synth

    rng = np.random.default_rng(515)
    ts = [(0.0, 0.0, 0.9), (0.05, -0.03, 0.65), (-0.1, 0.06, 1.2),      # working range
          (0.0, 0.0, 0.45), (0.03, 0.02, 0.32), (-0.02, 0.01, 0.55),    # close-ups
          (0.1, -0.1, 2.5), (0.0, 0.0, 5.0),                            # far: sub-pixel triangles only
          (0.01, 0.0, 0.27),                                            # cut by the near plane
          (0.33, 0.2, 0.8)]                                             # partly outside the image
    return [(synth.random_rotation(rng), np.asarray(t, np.float64)) for t in ts]


def _check(o_depth, o_tri, depth, tri, min_px):
    mo, mg = o_depth > 0, depth > 0
    assert (mo == mg).mean() >= 0.999, ("mask", (mo == mg).mean())
    both = mo & mg
    assert both.sum() >= min_px, both.sum()
    rel = np.abs(depth[both] - o_depth[both]) / o_depth[both]
    assert rel.max() <= 1e-5, ("depth", rel.max())
    assert (tri == o_tri).mean() >= 0.999, ("tri", (tri == o_tri).mean())
    return bool(np.array_equal(mo, mg) and np.array_equal(tri, o_tri))


def test_dense_mesh_textured_matches_oracle(cuda_device, dense_meshes):
    import torch

    from mx_deepim_b200.render import Render_Py

    r = Render_Py(None, ["a", "b", "c"], K, meshes=dense_meshes, max_batch=16)
    poses = _pose_set()
    P = np.zeros((len(poses), 3, 4), np.float32)
    for i, (R, t) in enumerate(poses):
        P[i, :, :3], P[i, :, 3] = R, t
    cls = (np.arange(len(poses)) % 3).astype(np.int32)
    out = r.render_batch(torch.from_numpy(cls).to(cuda_device), torch.from_numpy(P).to(cuda_device),
                         want=("depth", "bgr", "tri_id", "mask"))
    torch.cuda.synchronize()
    r.dev.check_status()
    depth, bgr, tri = out["depth"].cpu().numpy(), out["bgr"].cpu().numpy(), out["tri_id"].cpu().numpy()
    exact = 0
    for i in range(len(poses)):
        # the oracle sees the float32 pose the device got
        o_bgr, o_depth, o_tri = raster_np.render_py(dense_meshes[cls[i]], P[i, :, :3], P[i, :, 3], K,
                                                    r_type="mat", want_tri=True)
        z = float(P[i, 2, 3])
        exact += _check(o_depth, o_tri, depth[i, 0], tri[i], min_px=20 if z > 2 else 500)
        both = (o_depth > 0) & (depth[i, 0] > 0)
        same = (np.abs(bgr[i][both] - o_bgr[both]).max(axis=-1) == 0).mean()
        assert same >= 0.999, ("colour", i, same)
        assert np.array_equal(out["mask"][i, 0].cpu().numpy(), (depth[i, 0] > 0.2).astype(np.float32))
    assert exact >= len(poses) - 1  # coverage and winners are bit-identical (one pose of slack)
    # micro-triangle culling did fire: far fewer records than triangles on the far poses
    st = r.dev.stats(len(poses))
    assert 0 < st["records"] < 2 * 20480 * len(poses)


def test_dense_mesh_single_render_entry(cuda_device, dense_meshes):
    """The drop-in Render_Py.render(cls, r, t) entry (one pair, host arrays) on the same meshes."""
    from mx_deepim_b200.render import Render_Py

    r = Render_Py(None, ["a", "b", "c"], K, meshes=dense_meshes, max_batch=1)
    for i, (R, t) in enumerate(_pose_set()[:4]):
        R32, t32 = R.astype(np.float32), t.astype(np.float32)
        bgr, depth = r.render(i % 3, R32, t32, r_type="mat")
        o_bgr, o_depth = raster_np.render_py(dense_meshes[i % 3], R32, t32, K, r_type="mat")
        mo, mg = o_depth > 0, depth > 0
        assert (mo == mg).mean() >= 0.999
        both = mo & mg
        assert (np.abs(depth[both] - o_depth[both]) / o_depth[both]).max() <= 1e-5
        assert (np.abs(bgr[both] - o_bgr[both]).max(axis=-1) == 0).mean() >= 0.999


@pytest.mark.parametrize("mode", ["lambert", "lit_linemod"])
def test_dense_mesh_lit_modes_match_oracle(cuda_device, dense_meshes, mode):
    import torch

    from mx_deepim_b200 import _lib
    from mx_deepim_b200.render import DeviceRenderer

    meshes = []
    for m in dense_meshes:
        m = dict(m)
        if mode == "lambert":
            m["texture"] = np.full((8, 8, 3), 180, np.uint8)  # gray_texture.png stand-in
        meshes.append(m)
    dev = DeviceRenderer(meshes, K, max_batch=8, device=cuda_device)
    poses = _pose_set()[:7]
    P = np.zeros((len(poses), 3, 4), np.float32)
    for i, (R, t) in enumerate(poses):
        P[i, :, :3], P[i, :, 3] = R, t
    cls = (np.arange(len(poses)) % 3).astype(np.int32)
    t_ = P[:, :, 3]
    lp = (np.array([0, 1, 1], np.float32) * 0.5 + np.stack([t_[:, 0], -t_[:, 1], -t_[:, 2]], 1)).astype(np.float32)
    li = np.tile(np.array([[0.95, 1.0, 1.05]], np.float32), (len(poses), 1))
    ratio = 0.7 if mode == "lambert" else 0.4
    out = dev.render_batch(torch.from_numpy(cls).to(cuda_device), torch.from_numpy(P).to(cuda_device),
                           mode=_lib.RC_RENDER_LAMBERT if mode == "lambert" else _lib.RC_RENDER_LIT_LINEMOD,
                           light_pos=torch.from_numpy(lp).to(cuda_device),
                           light_int=torch.from_numpy(li).to(cuda_device), brightness_ratio=ratio,
                           want=("depth", "bgr"))
    torch.cuda.synchronize()
    dev.check_status()
    depth, bgr = out["depth"].cpu().numpy(), out["bgr"].cpu().numpy()
    fn = raster_np.render_py_light_modelnet if mode == "lambert" else raster_np.render_py_light
    for i in range(len(poses)):
        o_bgr, o_depth = fn(meshes[cls[i]], P[i, :, :3], P[i, :, 3], lp[i], li[i], K, r_type="mat",
                            brightness_ratio=ratio)
        mo, mg = o_depth > 0, depth[i, 0] > 0
        assert (mo == mg).mean() >= 0.999
        both = mo & mg
        assert both.sum() > (100 if P[i, 2, 3] > 2 else 400)
        assert (np.abs(depth[i, 0][both] - o_depth[both]) / o_depth[both]).max() <= 1e-5
        # 8-bit colour: within one level on >= 99.9 % of the covered pixels
        d = np.abs(np.round(bgr[i][both]).astype(np.int32) - o_bgr[both].astype(np.int32)).max(axis=-1)
        assert (d <= 1).mean() >= 0.999, (mode, i, (d <= 1).mean())
