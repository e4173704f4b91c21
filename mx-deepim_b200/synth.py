"""Synthetic LINEMOD-shaped inputs for benchmarks and smoke runs (no dataset, no network).

Recipe of SURVEY.md section 8(d): LINEMOD intrinsics
(/root/reference/experiments/deepim/cfgs/*LM*.yaml:36-45), 640x480, target
poses with t.z in [0.6, 1.2] m and the projected centre inside the image,
source poses = target o noise with Euler sigma 15 deg (reject > 45 deg),
sigma_xy 0.01 m, sigma_z 0.05 m
(/root/reference/toolkit/LM6d_1_gen_rendered_pose.py:59,115-117), observed
depth quantised to 1 mm (u16 PNG / 1000, /root/reference/lib/utils/image.py:200-205).

Two geometry sources: procedural triangle meshes (for the rasteriser path) and
analytic ellipsoids (closed form depth, used where no renderer is involved).
"""
from __future__ import annotations

import numpy as np

from . import pose as P

H, W = 480, 640
K_LINEMOD = np.array([[572.4114, 0.0, 325.2611], [0.0, 573.57043, 242.04899], [0.0, 0.0, 1.0]],
                     dtype=np.float32)
PIXEL_MEANS = np.array([123.68, 116.779, 103.939], dtype=np.float32)
ZNEAR, ZFAR = 0.25, 6.0
NUM_3D_SAMPLE = 3000
LW_PM = 0.1
NORMALIZE_3D_POINT = 0.1


def random_rotation(rng):
    q = rng.normal(size=4)
    return P.quat2mat(q / np.linalg.norm(q))


def noise_rotation(rng, sigma_deg=15.0, max_deg=45.0):
    while True:
        ang = np.deg2rad(rng.normal(0, sigma_deg, size=3))
        if np.all(np.abs(ang) < np.deg2rad(max_deg)):
            break
    cx, cy, cz = np.cos(ang)
    sx, sy, sz = np.sin(ang)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def random_pose_pair(rng, K=K_LINEMOD, margin=60, height=H, width=W):
    """(pose_src, pose_tgt) float32 [3,4]."""
    z = rng.uniform(0.6, 1.2)
    u = rng.uniform(margin, width - margin)
    v = rng.uniform(margin, height - margin)
    t = np.array([(u - K[0, 2]) / K[0, 0] * z, (v - K[1, 2]) / K[1, 1] * z, z])
    R = random_rotation(rng)
    pose_tgt = np.concatenate([R, t[:, None]], axis=1)
    dt = rng.normal(0, [0.01, 0.01, 0.05])
    pose_src = np.concatenate([noise_rotation(rng) @ R, (t + dt)[:, None]], axis=1)
    return pose_src.astype(np.float32), pose_tgt.astype(np.float32)


def ellipsoid_depth(pose, radii, K=K_LINEMOD, height=H, width=W):
    """Analytic depth map (metres, 0 = background) of an ellipsoid at pose [R|t]."""
    R = pose[:, :3].astype(np.float64)
    t = pose[:, 3].astype(np.float64)
    Kinv = np.linalg.inv(K.astype(np.float64))
    u, v = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    rays = np.stack([u, v, np.ones_like(u)], axis=-1) @ Kinv.T
    S = np.diag(1.0 / np.asarray(radii, dtype=np.float64))
    d = rays @ (S @ R.T).T
    o = -(S @ R.T @ t)
    a = np.sum(d * d, axis=-1)
    b = 2 * np.sum(d * o, axis=-1)
    c = np.dot(o, o) - 1.0
    disc = b * b - 4 * a * c
    hit = disc > 0
    s = np.where(hit, (-b - np.sqrt(np.where(hit, disc, 0))) / (2 * a), 0.0)
    return np.where(hit & (s > 0), s, 0.0).astype(np.float32)


def sample_points(rng, radii, num_points=NUM_3D_SAMPLE, num_keep=None):
    """Model point cloud + weights, zero padded (lib/utils/image.py:559-589)."""
    num_keep = num_points if num_keep is None else min(num_keep, num_points)
    d = rng.normal(size=(num_keep, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    pts = np.zeros((3, num_points), np.float32)
    wts = np.zeros((3, num_points), np.float32)
    pts[:, :num_keep] = (d * np.asarray(radii)).T
    wts[:, :num_keep] = 1
    return pts, wts


def make_compare_batch(batch, seed, K=K_LINEMOD, num_points=NUM_3D_SAMPLE, occlusion=False):
    """Host (numpy, float32) inputs of one zoom + flow + loss step for `batch` pairs.

    Keys follow the reference's blob names (batch_updater_py_multi.py:100-125):
      image_observed, image_rendered [B,3,H,W] (mean subtracted, RGB order)
      depth_gt_observed, depth_rendered [B,1,H,W]; mask_observed, mask_gt_observed [B,1,H,W]
      src_pose, tgt_pose [B,3,4]; KT [B,3,4]; Kinv [3,3]
      point_cloud_model, point_cloud_weights, point_cloud_observed [B,3,N]
      rot_est [B,4], trans_est [B,3]  (a stand-in network output: damped GT update + noise)
      trans_gt [B,3]
    """
    rng = np.random.default_rng(seed)
    pm = PIXEL_MEANS[[2, 1, 0]].reshape(1, 3, 1, 1)
    out = {
        "depth_rendered": np.zeros((batch, 1, H, W), np.float32),
        "depth_gt_observed": np.zeros((batch, 1, H, W), np.float32),
        "src_pose": np.zeros((batch, 3, 4), np.float32),
        "tgt_pose": np.zeros((batch, 3, 4), np.float32),
        "KT": np.zeros((batch, 3, 4), np.float32),
        "point_cloud_model": np.zeros((batch, 3, num_points), np.float32),
        "point_cloud_weights": np.zeros((batch, 3, num_points), np.float32),
        "point_cloud_observed": np.zeros((batch, 3, num_points), np.float32),
        "rot_est": np.zeros((batch, 4), np.float32),
        "trans_est": np.zeros((batch, 3), np.float32),
        "trans_gt": np.zeros((batch, 3), np.float32),
    }
    label = np.zeros((batch, 1, H, W), np.float32)
    for b in range(batch):
        ps, pt = random_pose_pair(rng, K)
        radii = rng.uniform(0.04, 0.12, size=3)
        out["depth_rendered"][b, 0] = ellipsoid_depth(ps, radii, K)
        d = np.round(ellipsoid_depth(pt, radii, K).astype(np.float64) * 1000.0) / 1000.0
        label[b, 0] = d > 0  # object label mask: an occluder only ever removes pixels from it
        if occlusion:
            x0, y0 = rng.integers(0, W - 80), rng.integers(0, H - 80)
            d[y0:y0 + 80, x0:x0 + 80] = 0.4
            label[b, 0, y0:y0 + 80, x0:x0 + 80] = 0
        out["depth_gt_observed"][b, 0] = d.astype(np.float32)
        out["src_pose"][b], out["tgt_pose"][b] = ps, pt
        out["KT"][b] = P.make_KT(K, ps, pt)
        pts, wts = sample_points(rng, radii, num_points, num_keep=int(rng.integers(2000, num_points + 1)))
        out["point_cloud_model"][b] = pts
        out["point_cloud_weights"][b] = wts
        out["point_cloud_observed"][b] = (pt[:, :3].astype(np.float64) @ pts.astype(np.float64) +
                                          pt[:, 3:].astype(np.float64))
        q, t = P.calc_RT_delta(ps, pt, np.zeros(3), np.ones(3), "CAMERA")
        out["trans_gt"][b] = t
        qe = q + rng.normal(0, 0.02, size=4)
        out["rot_est"][b] = qe / np.linalg.norm(qe)
        out["trans_est"][b] = t + rng.normal(0, 0.01, size=3)
    out["Kinv"] = P.make_Kinv(K)
    dr, do = out["depth_rendered"], out["depth_gt_observed"]
    out["mask_gt_observed"] = label
    mo = out["mask_gt_observed"].copy()
    for b in range(batch):
        y0, x0 = rng.integers(0, H - 40), rng.integers(0, W - 40)
        mo[b, 0, y0:y0 + 40, x0:x0 + 40] = 0
    out["mask_observed"] = mo
    tex = rng.uniform(0, 255, size=(batch, 3, H, W)).astype(np.float32)
    out["image_rendered"] = (tex * (dr > 0) - pm).astype(np.float32)
    bg = rng.uniform(1, 255, size=(batch, 3, H, W)).astype(np.float32)
    out["image_observed"] = (np.where(label > 0, tex[:, ::-1], bg) - pm).astype(np.float32)
    return out


# ----------------------------------------------------------------- procedural meshes
def icosphere(level):
    """Unit icosphere: 20 * 4**level triangles (level 4 -> 5120 tris, level 5 -> 20480)."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t),
         (0, 1, -t), (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    verts = [np.array(p, dtype=np.float64) / np.linalg.norm(p) for p in v]
    faces = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4),
             (11, 10, 2), (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8),
             (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    for _ in range(level):
        cache = {}

        def mid(a, b):
            key = (a, b) if a < b else (b, a)
            if key not in cache:
                m = verts[a] + verts[b]
                verts.append(m / np.linalg.norm(m))
                cache[key] = len(verts) - 1
            return cache[key]

        nf = []
        for a, b, c in faces:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        faces = nf
    return np.array(verts), np.array(faces, dtype=np.int32)


def vertex_normals(vertices, triangles):
    """Area-weighted per-vertex normals."""
    v = np.asarray(vertices, np.float64)
    f = np.asarray(triangles)
    fn = np.cross(v[f[:, 1]] - v[f[:, 0]], v[f[:, 2]] - v[f[:, 0]])
    n = np.zeros_like(v)
    for k in range(3):
        np.add.at(n, f[:, k], fn)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    return n / np.where(ln > 0, ln, 1.0)


def procedural_texture(rng, size=512):
    """[size,size,3] uint8 RGB: smooth colour field + checker so texel errors are visible."""
    y, x = np.mgrid[0:size, 0:size].astype(np.float64) / size
    ph = rng.uniform(0, 2 * np.pi, size=6)
    img = np.stack([0.5 + 0.35 * np.sin(2 * np.pi * (3 * x + 2 * y) + ph[0]),
                    0.5 + 0.35 * np.sin(2 * np.pi * (1 * x - 4 * y) + ph[1]),
                    0.5 + 0.35 * np.sin(2 * np.pi * (5 * x + 1 * y) + ph[2])], axis=-1)
    checker = (((x * 16).astype(int) + (y * 16).astype(int)) % 2)[..., None] * 0.12
    return np.clip((img + checker) * 255, 1, 255).astype(np.uint8)


def make_mesh(rng, level=4, diameter=None, kind=None, texture_size=512):
    """Closed procedural mesh (displaced icosphere / superquadric / squashed blob) with
    spherical UVs, vertex normals and an RGB texture.  Diameter 0.1-0.3 m (LINEMOD-like)."""
    sphere, tris = icosphere(level)
    kind = kind if kind is not None else int(rng.integers(0, 3))
    x, y, z = sphere.T
    if kind == 0:    # low-frequency displacement noise
        a = rng.uniform(1.5, 4.0, size=3)
        p = rng.uniform(0, 2 * np.pi, size=3)
        r = 1 + 0.18 * np.sin(a[0] * x + p[0]) * np.sin(a[1] * y + p[1]) + 0.1 * np.sin(a[2] * z + p[2])
        verts = sphere * r[:, None]
    elif kind == 1:  # superquadric
        e = rng.uniform(0.5, 1.6, size=2)
        verts = np.sign(sphere) * np.abs(sphere) ** e[0]
        verts[:, 2] = np.sign(z) * np.abs(z) ** e[1]
    else:            # anisotropic blob
        verts = sphere * rng.uniform(0.45, 1.0, size=3)
    diameter = diameter if diameter is not None else rng.uniform(0.1, 0.3)
    extent = np.linalg.norm(verts.max(0) - verts.min(0))
    verts = (verts - verts.mean(0)) * (diameter / extent)
    u = 0.5 + np.arctan2(sphere[:, 1], sphere[:, 0]) / (2 * np.pi)
    v = 0.5 + np.arcsin(np.clip(sphere[:, 2], -1, 1)) / np.pi
    return {"vertices": verts.astype(np.float32), "uv": np.stack([u, v], 1).astype(np.float32),
            "normals": vertex_normals(verts, tris).astype(np.float32),
            "triangles": tris.astype(np.int32), "texture": procedural_texture(rng, texture_size)}


def mesh_points(mesh, rng, num_points=NUM_3D_SAMPLE):
    """point_cloud_model / weights from mesh vertices (lib/utils/image.py:559-589)."""
    v = np.asarray(mesh["vertices"], np.float64)
    keep = min(len(v), num_points)
    idx = rng.permutation(len(v))[:keep]
    pts = np.zeros((3, num_points), np.float32)
    wts = np.zeros((3, num_points), np.float32)
    pts[:, :keep] = v[idx].T
    wts[:, :keep] = 1
    return pts, wts
