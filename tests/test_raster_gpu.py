"""CUDA rasteriser vs the CPU restatement of the GL pipeline (oracle/raster_oracle.c).

Bars (north_star): depth within 1e-5 relative, >= 99.9 % per-pixel mask agreement."""
import numpy as np
import pytest

import helpers
from oracle import raster_np

pytestmark = pytest.mark.gpu
K = helpers.K_LM


def _meshes():
    from mx_deepim_b200 import synth

    rng = np.random.default_rng(77)
    return [synth.make_mesh(rng, level=3, kind=0), synth.make_mesh(rng, level=4, kind=1),
            synth.make_mesh(rng, level=4, kind=2, texture_size=128)]


def _poses(n, seed, zrange=(0.6, 1.2)):
    from mx_deepim_b200 import synth

    rng = np.random.default_rng(seed)
    out = np.zeros((n, 3, 4), np.float32)
    for i in range(n):
        _, pt = synth.random_pose_pair(rng)
        out[i] = pt
    return out


def _compare(o_bgr, o_depth, bgr, depth, colour_tol=0.0):
    mo, mg = o_depth > 0, depth > 0
    agree = (mo == mg).mean()
    assert agree >= 0.999, agree
    both = mo & mg
    assert both.sum() > 1000
    rel = np.abs(depth[both] - o_depth[both]) / o_depth[both]
    assert rel.max() <= 1e-5, rel.max()
    same = (np.abs(bgr[both] - o_bgr[both]).max(axis=-1) <= colour_tol).mean()
    assert same >= 0.999, same
    return agree


def test_render_py_matches_oracle(cuda_device):
    from mx_deepim_b200.render import Render_Py

    meshes = _meshes()
    r = Render_Py(None, ["a", "b", "c"], K, meshes=meshes, max_batch=8)
    poses = _poses(6, 5)
    exact = []
    for i in range(6):
        m = i % 3
        bgr, depth = r.render(m, poses[i, :, :3], poses[i, :, 3], r_type="mat")
        assert bgr.shape == (480, 640, 3) and bgr.dtype == np.float32 and depth.shape == (480, 640)
        o_bgr, o_depth = raster_np.render_py(meshes[m], poses[i, :, :3], poses[i, :, 3], K, r_type="mat")
        exact.append(_compare(o_bgr, o_depth, bgr, depth))
        assert bgr[depth == 0].max() == 0
    assert min(exact) == 1.0  # coverage is bit-exact on these cases
    # quaternion entry, K override
    from oracle import pose_np

    q = pose_np.mat2quat(poses[0, :, :3])
    K2 = K.copy()
    K2[0, 2] += 7.5
    bgr, depth = r.render(0, q, poses[0, :, 3], r_type="quat", K=K2)
    o_bgr, o_depth = raster_np.render_py(meshes[0], q, poses[0, :, 3], K2, r_type="quat")
    _compare(o_bgr, o_depth, bgr, depth)


def test_render_batch_fused_post(cuda_device):
    """Batched fast path: mixed meshes in one call; image = RGB - means (CHW), mask = depth > 0.2
    (batch_updater_py_multi.py:279-320); tri ids against the oracle."""
    import torch

    from mx_deepim_b200.render import Render_Py

    meshes = _meshes()
    r = Render_Py(None, ["a", "b", "c"], K, meshes=meshes, max_batch=8)
    poses = _poses(7, 9)
    cls = np.array([0, 1, 2, 2, 1, 0, 1], np.int32)
    means = helpers.PIXEL_MEANS[[2, 1, 0]]
    out = r.render_batch(torch.from_numpy(cls).to(cuda_device), torch.from_numpy(poses).to(cuda_device),
                         pixel_means=means, want=("depth", "image", "mask", "bgr", "tri_id"))
    torch.cuda.synchronize()
    assert int(r.dev.status.item()) == 0
    depth = out["depth"].cpu().numpy()
    image = out["image"].cpu().numpy()
    mask = out["mask"].cpu().numpy()
    bgr = out["bgr"].cpu().numpy()
    tri = out["tri_id"].cpu().numpy()
    for i in range(7):
        o_bgr, o_depth, o_tri = raster_np.render_py(meshes[cls[i]], poses[i, :, :3], poses[i, :, 3], K,
                                                    r_type="mat", want_tri=True)
        _compare(o_bgr, o_depth, bgr[i], depth[i, 0])
        assert (tri[i] == o_tri).mean() >= 0.999
        ref_img = bgr[i][:, :, [2, 1, 0]].transpose(2, 0, 1) - means.reshape(3, 1, 1)
        assert np.array_equal(image[i], ref_img.astype(np.float32))
        assert np.array_equal(mask[i, 0], (depth[i, 0] > 0.2).astype(np.float32))
    st = r.dev.stats(7)
    # micro-triangles go straight to the z-winner blocks: neither record nor tile-list entry
    assert st["records"] > 0 and st["tile_entries"] >= st["records"] - st["direct"] >= 0


def test_near_plane_clipping_and_offscreen(cuda_device):
    from mx_deepim_b200.render import Render_Py

    meshes = _meshes()
    r = Render_Py(None, ["a", "b", "c"], K, meshes=meshes, max_batch=2)
    rng = np.random.default_rng(3)
    from mx_deepim_b200 import synth

    R = synth.random_rotation(rng)
    # object straddling the near plane (z = 0.25) and partly outside the image
    for t in ([0.02, 0.01, 0.27], [0.25, -0.15, 0.45], [0.0, 0.0, 0.1]):
        bgr, depth = r.render(1, R, np.array(t), r_type="mat")
        o_bgr, o_depth = raster_np.render_py(meshes[1], R, np.array(t), K, r_type="mat")
        assert ((depth > 0) == (o_depth > 0)).mean() >= 0.999
        both = (depth > 0) & (o_depth > 0)
        if both.any():
            assert (np.abs(depth[both] - o_depth[both]) / o_depth[both]).max() <= 1e-5
            assert depth[both].min() >= 0.25 * (1 - 1e-6)
    # fully behind the camera: empty frame
    bgr, depth = r.render(1, R, np.array([0.0, 0.0, -1.0]), r_type="mat")
    assert not depth.any() and not bgr.any()


def test_plane_depth_property(cuda_device):
    """A fronto-parallel quad renders to constant depth z (SURVEY.md section 4 property test)."""
    from mx_deepim_b200.render import Render_Py

    quad = {"vertices": np.array([[-0.1, -0.1, 0], [0.1, -0.1, 0], [0.1, 0.1, 0], [-0.1, 0.1, 0]], np.float32),
            "uv": np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32), "normals": None,
            "triangles": np.array([[0, 1, 2], [0, 2, 3]], np.int32),
            "texture": np.full((4, 4, 3), 200, np.uint8)}
    r = Render_Py(None, ["q"], K, meshes=[quad], max_batch=1)
    bgr, depth = r.render(0, np.eye(3), np.array([0.0, 0.0, 0.8]), r_type="mat")
    fg = depth > 0
    assert abs(fg.sum() - (0.2 * K[0, 0] / 0.8) * (0.2 * K[1, 1] / 0.8)) < 600
    assert np.abs(depth[fg] - 0.8).max() < 1e-6
    assert np.all(bgr[fg] == 200.0)


def test_lambert_modelnet(cuda_device):
    from mx_deepim_b200.render import Render_Py_Light_ModelNet_Multi

    meshes = _meshes()
    for m in meshes:
        m["texture"] = np.full((8, 8, 3), 180, np.uint8)  # gray_texture.png stand-in
    r = Render_Py_Light_ModelNet_Multi(None, None, K, meshes=meshes, max_batch=2)
    poses = _poses(3, 21)
    for i in range(3):
        t = poses[i, :, 3]
        light_position = np.array([0, 1, 1]) * 0.5 + np.array([t[0], -t[1], -t[2]])  # :233-251
        light_intensity = np.array([0.95, 1.0, 1.05])
        bgr, depth = r.render(i, poses[i, :, :3], t, light_position, light_intensity, r_type="mat")
        assert bgr.dtype == np.uint8
        o_bgr, o_depth = raster_np.render_py_light_modelnet(meshes[i], poses[i, :, :3], t, light_position,
                                                            light_intensity, K, r_type="mat")
        _compare(o_bgr.astype(np.float32), o_depth, bgr.astype(np.float32), depth, colour_tol=1.0)
        assert bgr[depth > 0].std() > 1.0  # shading varies over the surface


def test_linemod_light_and_single_model_renderers(cuda_device):
    """The remaining lib/render_glumpy classes, imported by the reference's module names."""
    from mx_deepim_b200.render_glumpy.render_py import Render_Py as Render_Py_1
    from mx_deepim_b200.render_glumpy.render_py_depth import Render_Py as Render_Py_D
    from mx_deepim_b200.render_glumpy.render_py_light import Render_Py_Light
    from mx_deepim_b200.render_glumpy.render_py_light_modelnet import Render_Py_Light_ModelNet
    from mx_deepim_b200.render_glumpy.render_py_light_multi_program import Render_Py_Light_MultiProgram
    from mx_deepim_b200.render_glumpy.render_py_multi import Render_Py as Render_Py_M  # noqa: F401

    meshes = _meshes()
    poses = _poses(3, 33)
    R, t = poses[0, :, :3], poses[0, :, 3]
    # single-model textured + its uint8 "depth" twin
    r1 = Render_Py_1(None, K, mesh=meshes[2], max_batch=1)
    bgr, depth = r1.render(R, t, r_type="mat")
    o_bgr, o_depth = raster_np.render_py(meshes[2], R, t, K, r_type="mat")
    _compare(o_bgr, o_depth, bgr, depth)
    rd = Render_Py_D(None, K, mesh=meshes[2], max_batch=1)
    bgr8, depth8 = rd.render(R, t, r_type="mat")
    assert bgr8.dtype == np.uint8 and np.array_equal(depth8, depth)
    assert np.array_equal(bgr8, np.round(bgr).astype(np.uint8))
    # LINEMOD light shader: t * ((1-b) + b * brightness * I), three brightness programs
    lp = np.array([0.3, -0.4, -0.2])
    li = np.array([1.3, 0.9, 1.1])
    rl = Render_Py_Light(None, K, mesh=meshes[2], max_batch=1)
    outs = []
    for k, ratio in enumerate([0.4, 0.3, 0.2]):
        bgr, depth = rl.render(R, t, lp, li, brightness_k=k, r_type="mat")
        assert bgr.dtype == np.uint8
        o_bgr, o_depth = raster_np.render_py_light(meshes[2], R, t, lp, li, K, r_type="mat",
                                                   brightness_ratio=ratio)
        _compare(o_bgr.astype(np.float32), o_depth, bgr.astype(np.float32), depth, colour_tol=1.0)
        outs.append(bgr.astype(np.float32)[depth > 0].mean())
    assert len(set(np.round(outs, 3))) == 3  # the three programs shade differently
    # ... and it is NOT the ModelNet formula
    m_bgr, _ = raster_np.render_py_light_modelnet(meshes[2], R, t, lp, li, K, r_type="mat",
                                                  brightness_ratio=0.4)
    bgr, depth = rl.render(R, t, lp, li, brightness_k=0, r_type="mat")
    assert np.abs(bgr.astype(np.float32) - m_bgr.astype(np.float32))[depth > 0].max() > 2
    # multi-class by name
    names = ["ape", "can", "cat"]
    rm = Render_Py_Light_MultiProgram(names, None, K, meshes=meshes, max_batch=1)
    for i, n in enumerate(names):
        Ri, ti = poses[i, :, :3], poses[i, :, 3]
        bgr, depth = rm.render(Ri, ti, lp, li, n, brightness_k=1, r_type="mat")
        o_bgr, o_depth = raster_np.render_py_light(meshes[i], Ri, ti, lp, li, K, r_type="mat",
                                                   brightness_ratio=0.3)
        _compare(o_bgr.astype(np.float32), o_depth, bgr.astype(np.float32), depth, colour_tol=1.0)
    # single-model ModelNet
    rn = Render_Py_Light_ModelNet(None, None, K, mesh=meshes[0], max_batch=1)
    bgr, depth = rn.render(R, t, lp, li, r_type="mat")
    o_bgr, o_depth = raster_np.render_py_light_modelnet(meshes[0], R, t, lp, li, K, r_type="mat")
    _compare(o_bgr.astype(np.float32), o_depth, bgr.astype(np.float32), depth, colour_tol=1.0)


def test_chained_render_equals_serialised_full_size(cuda_device):
    """BASELINE-sized property test (cfg 3 batch, cfg 5 mesh density): the production issue order
    (setup ... cover chained behind the background pass by programmatic dependent launches) gives
    bit-identical frames to the serialised order that event marks force, back to back over
    changing poses, so does the incremental background (only the previous frame's tiles are
    re-cleared), and every output plane is consistent with the depth plane."""
    import torch

    from mx_deepim_b200 import _lib, synth
    from mx_deepim_b200.render import DeviceRenderer

    rng = np.random.default_rng(5)
    meshes = [synth.make_mesh(rng, level=4), synth.make_mesh(rng, level=5),
              synth.make_mesh(rng, level=3, kind=2)]
    B = 64
    ren = DeviceRenderer(meshes, synth.K_LINEMOD, synth.W, synth.H, synth.ZNEAR, synth.ZFAR,
                         max_batch=B, device=cuda_device)
    cls = torch.from_numpy(rng.integers(0, len(meshes), size=B).astype(np.int32)).to(cuda_device)
    means = (1.0, 2.0, 3.0)
    want = ("depth", "image", "mask", "tri_id")
    chained, serial, incr = {}, {}, {}
    for rep, zr in enumerate(((0.6, 1.2), (0.3, 0.6), (0.6, 1.2), (0.8, 1.0))):
        poses = _poses(B, 900 + rep)
        poses[:, 2, 3] = rng.uniform(zr[0], zr[1], size=B)
        pd = torch.from_numpy(poses).to(cuda_device)
        ren.render_batch(cls, pd, pixel_means=means, want=want, out=chained)
        marks = [_lib.Event() for _ in range(_lib.RC_RENDER_MARKS)]
        ren.render_batch(cls, pd, pixel_means=means, want=want, out=serial, marks=marks)
        # incremental background: `incr` still holds this renderer's previous frame only if the
        # previous call wrote to it -- the first call below follows a call on other buffers (full
        # background), the second one follows itself (only the dirtied tiles are re-cleared)
        ren.render_batch(cls, pd, pixel_means=means, want=want, out=incr, incremental=True)
        if rep > 0:
            pd2 = torch.from_numpy(_poses(B, 950 + rep)).to(cuda_device)
            ren.render_batch(cls, pd2, pixel_means=means, want=want, out=incr, incremental=True)
            ren.render_batch(cls, pd, pixel_means=means, want=want, out=incr, incremental=True)
        torch.cuda.synchronize()
        assert int(ren.status.item()) == 0
        for k in want:
            assert torch.equal(chained[k], serial[k]), (rep, k)
            assert torch.equal(incr[k], serial[k]), (rep, k, "incremental")
        depth, tri = chained["depth"], chained["tri_id"]
        assert torch.equal(chained["mask"], (depth > 0.2).float())
        assert torch.equal(tri[:, None] >= 0, depth > 0)  # a winner everywhere a depth was written
        bg = (depth == 0).expand(-1, 3, -1, -1)
        img = chained["image"]
        for c in range(3):
            assert bool((img[:, c][bg[:, c]] == -means[c]).all())
        fg = depth[depth > 0]
        assert float(fg.min()) >= synth.ZNEAR * (1 - 1e-6) and float(fg.max()) <= synth.ZFAR
        assert int((depth > 0).sum()) > 1000 * B / 4


@pytest.mark.parametrize("width,height", [(200, 150), (332, 250), (1024, 64)])
def test_render_other_frame_sizes(cuda_device, width, height):
    """Frames whose size is not a multiple of the 16x16 raster tile (partial tiles at the right
    and bottom border) or of the 128-px row chunks of the background pass, against the oracle."""
    import torch

    from mx_deepim_b200 import synth
    from mx_deepim_b200.render import DeviceRenderer

    rng = np.random.default_rng(width)
    mesh = synth.make_mesh(rng, level=3, kind=1)
    Ks = np.array([[250.0, 0, width / 2 - 3.3], [0, 251.0, height / 2 + 1.7], [0, 0, 1]], np.float32)
    ren = DeviceRenderer([mesh], Ks, width, height, 0.25, 6.0, max_batch=3, device=cuda_device)
    poses = np.zeros((3, 3, 4), np.float32)
    for b in range(3):
        poses[b, :, :3] = synth.random_rotation(rng)
        # the third one hangs over the right / bottom border
        poses[b, :, 3] = [(0.0, 0.0, 0.7), (-0.05, 0.02, 0.5), (0.26 * width / 250, 0.2 * height / 251, 0.9)][b]
    cls = torch.zeros(3, dtype=torch.int32, device=cuda_device)
    out = {}
    for incremental in (False, True, True):
        ren.render_batch(cls, torch.from_numpy(poses).to(cuda_device), want=("depth", "bgr", "mask", "tri_id"),
                         out=out, incremental=incremental)
    torch.cuda.synchronize()
    assert int(ren.status.item()) == 0
    depth, bgr = out["depth"].cpu().numpy(), out["bgr"].cpu().numpy()
    for b in range(3):
        o_bgr, o_depth = raster_np.render_py(mesh, poses[b, :, :3], poses[b, :, 3], Ks, r_type="mat",
                                             width=width, height=height)
        assert ((depth[b, 0] > 0) == (o_depth > 0)).mean() >= 0.999
        both = (depth[b, 0] > 0) & (o_depth > 0)
        if both.any():
            assert (np.abs(depth[b, 0][both] - o_depth[both]) / o_depth[both]).max() <= 1e-5
            assert (np.abs(bgr[b][both] - o_bgr[both]).max(axis=-1) == 0).mean() >= 0.999
    assert (depth[0] > 0).sum() > 200


def test_render_py_from_model_dir(cuda_device, tmp_path):
    """The reference's constructor path (render_py_multi.py:49-82): `<model_dir>/<class>/
    textured.obj` + `texture_map.png` read from disk, rendered, and compared with the oracle's
    render of the same files' contents."""
    import cv2

    from mx_deepim_b200 import synth
    from mx_deepim_b200.render import Render_Py, load_obj, load_texture

    rng = np.random.default_rng(41)
    classes = ["ape", "can"]
    for i, cls in enumerate(classes):
        m = synth.make_mesh(rng, level=2, kind=i, texture_size=64)
        d = tmp_path / cls
        d.mkdir()
        with open(d / "textured.obj", "w") as fh:
            for v in m["vertices"]:
                fh.write("v %.9g %.9g %.9g\n" % tuple(v))
            for t in m["uv"]:
                fh.write("vt %.9g %.9g\n" % tuple(t))
            for n in m["normals"]:
                fh.write("vn %.9g %.9g %.9g\n" % tuple(n))
            for a, b, c in m["triangles"] + 1:
                fh.write("f %d/%d/%d %d/%d/%d %d/%d/%d\n" % (a, a, a, b, b, b, c, c, c))
        cv2.imwrite(str(d / "texture_map.png"), np.ascontiguousarray(m["texture"][:, :, ::-1]))
    r = Render_Py(str(tmp_path), classes, K, max_batch=2)
    poses = _poses(2, 43)
    for i, cls in enumerate(classes):
        mesh = load_obj(str(tmp_path / cls / "textured.obj"))
        mesh["texture"] = load_texture(str(tmp_path / cls / "texture_map.png"))
        bgr, depth = r.render(i, poses[i, :, :3], poses[i, :, 3], r_type="mat")
        o_bgr, o_depth = raster_np.render_py(mesh, poses[i, :, :3], poses[i, :, 3], K, r_type="mat")
        _compare(o_bgr, o_depth, bgr, depth)
        assert (depth > 0).sum() > 500
