"""Seeded synthetic inputs shared by the tests (LINEMOD-shaped, no dataset needed)."""
from __future__ import annotations

import os

import numpy as np

H, W = 480, 640
# LINEMOD intrinsics, /root/reference/experiments/deepim/cfgs/*LM*.yaml:36-45
K_LM = np.array([[572.4114, 0.0, 325.2611], [0.0, 573.57043, 242.04899], [0.0, 0.0, 1.0]],
                dtype=np.float32)
PIXEL_MEANS = np.array([123.68, 116.779, 103.939], dtype=np.float32)  # cfgs/*.yaml:56-59


def random_rotation(rng):
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)],
    ])


def small_rotation(rng, sigma_deg=15.0, max_deg=45.0):
    """Pose noise like toolkit/LM6d_1_gen_rendered_pose.py:59,115-117."""
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


def random_pose_pair(rng, K=K_LM, margin=60):
    """(pose_src, pose_tgt) float32 [3,4]; projected centre at least `margin` px inside."""
    z = rng.uniform(0.6, 1.2)
    u = rng.uniform(margin, W - margin)
    v = rng.uniform(margin, H - margin)
    t = np.array([(u - K[0, 2]) / K[0, 0] * z, (v - K[1, 2]) / K[1, 1] * z, z])
    R = random_rotation(rng)
    pose_tgt = np.concatenate([R, t[:, None]], axis=1)
    dR = small_rotation(rng)
    dt = rng.normal(0, [0.01, 0.01, 0.05])
    pose_src = np.concatenate([dR @ R, (t + dt)[:, None]], axis=1)
    return pose_src.astype(np.float32), pose_tgt.astype(np.float32)


def ellipsoid_depth(pose, radii, K=K_LM, height=H, width=W):
    """Analytic depth map (metres, 0 = background) of an ellipsoid with semi-axes `radii`
    (object frame) at pose [R|t]: nearest intersection of each pixel ray with the surface."""
    R = pose[:, :3].astype(np.float64)
    t = pose[:, 3].astype(np.float64)
    Kinv = np.linalg.inv(K.astype(np.float64))
    u, v = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    rays = np.stack([u, v, np.ones_like(u)], axis=-1) @ Kinv.T  # camera-frame direction, z = 1
    S = np.diag(1.0 / np.asarray(radii, dtype=np.float64))
    d = rays @ (S @ R.T).T  # direction in the scaled object frame
    o = -(S @ R.T @ t)      # camera origin in the scaled object frame
    a = np.sum(d * d, axis=-1)
    b = 2 * np.sum(d * o, axis=-1)
    c = np.dot(o, o) - 1.0
    disc = b * b - 4 * a * c
    hit = disc > 0
    s = np.where(hit, (-b - np.sqrt(np.where(hit, disc, 0))) / (2 * a), 0.0)
    depth = np.where(hit & (s > 0), s, 0.0)  # ray has z = 1, so the parameter is the depth
    return depth.astype(np.float32)


def make_flow_inputs(batch, seed, occlude=False, K=K_LM):
    """depth_src (rendered at src pose), depth_tgt (observed at tgt pose, 1 mm quantised),
    KT, Kinv as the batch updater builds them (batch_updater_py_multi.py:306-345)."""
    from oracle import pose_np

    rng = np.random.default_rng(seed)
    depth_src = np.zeros((batch, 1, H, W), np.float32)
    depth_tgt = np.zeros((batch, 1, H, W), np.float32)
    KT = np.zeros((batch, 3, 4), np.float64)
    poses_src = np.zeros((batch, 3, 4), np.float32)
    poses_tgt = np.zeros((batch, 3, 4), np.float32)
    for b in range(batch):
        ps, pt = random_pose_pair(rng, K)
        radii = rng.uniform(0.04, 0.12, size=3)
        depth_src[b, 0] = ellipsoid_depth(ps, radii, K)
        d = ellipsoid_depth(pt, radii, K)
        d = np.round(d * 1000.0) / 1000.0  # u16 PNG depth in mm (lib/utils/image.py:200-205)
        if occlude:
            x0, y0 = rng.integers(0, W - 80), rng.integers(0, H - 80)
            d[y0:y0 + 80, x0:x0 + 80] = 0.4
        depth_tgt[b, 0] = d.astype(np.float32)
        KT[b] = pose_np.make_KT(K, ps, pt)
        poses_src[b], poses_tgt[b] = ps, pt
    Kinv = np.array(np.linalg.inv(np.matrix(K)))  # batch_updater_py_multi.py:70
    return dict(depth_src=depth_src, depth_tgt=depth_tgt, KT=KT.astype(np.float32),
                Kinv=Kinv.astype(np.float32), pose_src=poses_src, pose_tgt=poses_tgt)


def make_zoom_inputs(batch, seed, K=K_LM, empty_rendered=()):
    """Everything the zoom plugins consume for `batch` pairs: masks, mean-subtracted images,
    depths, flow + weights (from the flow oracle) and src poses."""
    from oracle import flow_np

    rng = np.random.default_rng(seed)
    f = make_flow_inputs(batch, seed, K=K)
    depth_r, depth_o = f["depth_src"], f["depth_tgt"]
    for b in empty_rendered:
        depth_r[b] = 0
    mask_rendered = depth_r.copy()  # the graph feeds rendered depth as the mask (zoom_mask.py:40)
    mask_gt_observed = (depth_o > 0).astype(np.float32)
    mask_observed = mask_gt_observed.copy()
    for b in range(batch):  # the estimated observed mask is a perturbed GT mask
        y0, x0 = rng.integers(0, H - 40), rng.integers(0, W - 40)
        mask_observed[b, 0, y0:y0 + 40, x0:x0 + 40] = 0
    pm = PIXEL_MEANS[[2, 1, 0]].reshape(1, 3, 1, 1)  # batch_updater_py_multi.py:24-25
    tex = rng.uniform(0, 255, size=(batch, 3, H, W)).astype(np.float32)
    image_rendered = (tex * (depth_r > 0) - pm).astype(np.float32)
    bg = rng.uniform(1, 255, size=(batch, 3, H, W)).astype(np.float32)
    image_observed = (np.where(depth_o > 0, tex[:, ::-1], bg) - pm).astype(np.float32)
    flow, valid = flow_np.flow_kernel_cpu(depth_r, depth_o, f["KT"], f["Kinv"])
    flow_weights = np.tile(valid, [1, 2, 1, 1])  # batch_updater_py_multi.py:352
    return dict(mask_observed=mask_observed, mask_gt_observed=mask_gt_observed,
                mask_rendered=mask_rendered, image_observed=image_observed,
                image_rendered=image_rendered, depth_observed=depth_o, depth_rendered=depth_r,
                flow=flow, flow_weights=flow_weights, src_pose=f["pose_src"],
                tgt_pose=f["pose_tgt"], KT=f["KT"], Kinv=f["Kinv"])


def make_points_inputs(batch, seed, num_points=3000, num_keep=None):
    """point_cloud_model / weights / observed as lib/utils/image.py:559-596 builds them, random
    unit quaternion + translation deltas near the GT update, src/tgt poses."""
    from oracle import pose_np

    rng = np.random.default_rng(seed)
    num_keep = num_points if num_keep is None else num_keep
    pts = np.zeros((batch, 3, num_points), np.float32)
    wts = np.zeros((batch, 3, num_points), np.float32)
    obs = np.zeros((batch, 3, num_points), np.float32)
    rot = np.zeros((batch, 4), np.float32)
    trans = np.zeros((batch, 3), np.float32)
    poses_src = np.zeros((batch, 3, 4), np.float32)
    poses_tgt = np.zeros((batch, 3, 4), np.float32)
    for b in range(batch):
        ps, pt = random_pose_pair(rng)
        model = rng.uniform(-0.1, 0.1, size=(num_keep, 3))
        pts[b, :, :num_keep] = model.T
        wts[b, :, :num_keep] = 1
        obs[b] = (pt[:, :3].astype(np.float64) @ pts[b].astype(np.float64) +
                  pt[:, 3:].astype(np.float64))
        q, t = pose_np.calc_RT_delta(ps, pt, np.zeros(3), np.ones(3), "CAMERA", "QUAT")
        q = q + rng.normal(0, 0.02, size=4)
        q /= np.linalg.norm(q)
        rot[b] = q
        trans[b] = t + rng.normal(0, 0.01, size=3)
        poses_src[b], poses_tgt[b] = ps, pt
    return dict(points=pts, weights=wts, points_obs=obs, rotation=rot, translation=trans,
                pose_src=poses_src, pose_tgt=poses_tgt)


def make_eval_inputs(seed, batch=4, num_points=1500):
    """Seeded inputs of the evaluation reductions: pose pairs, zero-padded model points with a
    per-pair valid count, and three depth maps (test scene, GT render, estimate render)."""
    rng = np.random.default_rng(seed)
    pose_est = np.zeros((batch, 3, 4), np.float32)
    pose_gt = np.zeros((batch, 3, 4), np.float32)
    points = np.zeros((batch, 3, num_points), np.float32)
    n_valid = np.zeros((batch,), np.int32)
    for b in range(batch):
        pose_est[b], pose_gt[b] = random_pose_pair(rng)
        n = num_points if b == 0 else int(rng.integers(num_points // 3, num_points))
        n_valid[b] = n
        points[b, :, :n] = (rng.normal(size=(3, n)) * np.array([[0.04], [0.06], [0.03]])).astype(np.float32)
    radii = rng.uniform(0.05, 0.1, size=3)
    d_gt = ellipsoid_depth(pose_gt[0], radii)
    d_est = ellipsoid_depth(pose_est[0], radii)
    d_test = d_gt.copy()
    d_test[d_test > 0] += rng.normal(0, 0.004, size=int((d_test > 0).sum())).astype(np.float32)
    d_test[200:260, 280:360] = 0.45  # an occluder in front of part of the frame
    d_test[:, :40] = 0
    return dict(pose_est=pose_est, pose_gt=pose_gt, points=points, n_valid=n_valid,
                d_test=d_test.astype(np.float32), d_gt=d_gt, d_est=d_est)


def make_updater_inputs(batch, seed, K=K_LM):
    """A batch for one pass of batchUpdaterPyMulti.forward: two procedural meshes, src / tgt
    poses, the observed depth (GL-restatement render at the target pose, 1 mm quantised) and
    'network predictions' = the ground-truth update damped by 0.5 plus noise."""
    from mx_deepim_b200 import synth
    from oracle import pose_np, raster_np

    rng = np.random.default_rng(seed)
    meshes = [synth.make_mesh(rng, level=3, kind=0, texture_size=128),
              synth.make_mesh(rng, level=3, kind=2, texture_size=128)]
    cls = rng.integers(0, len(meshes), size=batch).astype(np.int32)
    src = np.zeros((batch, 3, 4), np.float32)
    tgt = np.zeros((batch, 3, 4), np.float32)
    rot = np.zeros((batch, 4), np.float32)
    trans = np.zeros((batch, 3), np.float32)
    depth = np.zeros((batch, 1, H, W), np.float32)
    for b in range(batch):
        src[b], tgt[b] = random_pose_pair(rng, K)
        q, t = pose_np.calc_RT_delta(src[b], tgt[b], np.zeros(3), np.ones(3), "CAMERA", "QUAT")
        q = np.array([1.0, 0, 0, 0]) + 0.5 * (q - np.array([1.0, 0, 0, 0])) + rng.normal(0, 0.01, 4)
        rot[b] = q / np.linalg.norm(q)
        trans[b] = 0.5 * t + rng.normal(0, 0.005, size=3)
        d = raster_np.render_py(meshes[cls[b]], tgt[b, :, :3], tgt[b, :, 3], K, r_type="mat")[1]
        depth[b, 0] = np.round(d.astype(np.float64) * 1000.0) / 1000.0
    return dict(meshes=meshes, class_index=cls, src_pose=src, tgt_pose=tgt, rot_est=rot,
                trans_est=trans, depth_gt_observed=depth)


def make_prep_inputs(seed):
    """One frame for the loader / test-time update golden vectors: BGR uint8 image, model points,
    a pose, a depth map and the rendered / observed masks of a test-time batch."""
    rng = np.random.default_rng(seed)
    bgr = rng.integers(0, 256, size=(H, W, 3), dtype=np.uint8)
    points_obj = np.round(rng.normal(size=(1234, 3)) * 0.05, 6)
    pose = random_pose_pair(rng)[0]
    depth = np.zeros((H, W), np.float32)
    depth[100:300, 200:420] = rng.uniform(0.5, 1.2, size=(200, 220)).astype(np.float32)
    mask_rendered = np.zeros((H, W), np.float32)
    mask_rendered[120:310, 180:400] = (rng.random((190, 220)) > 0.3)
    mask_rendered[120, 180] = mask_rendered[309, 399] = 1
    mask_observed = np.zeros((H, W), np.float32)
    mask_observed[90:280, 220:450] = (rng.random((190, 230)) > 0.5)
    mask_observed[90, 220] = mask_observed[279, 449] = 1
    return dict(bgr=bgr, pixel_means=np.array([102.9801, 115.9465, 122.7717]), points_obj=points_obj,
                pose=pose, depth=depth, mask_rendered=mask_rendered, mask_observed=mask_observed)


def write_pair_files(root, seed, n_pairs=2):
    """Write `n_pairs` frames in the dataset's on-disk formats under `root` (8-bit BGR colour
    PNGs, 16-bit depth PNGs in mm, 8-bit label PNGs, -pose.txt with one header row, and one
    points.xyz) and return the pairdb records the reference's loaders take."""
    import cv2

    rng = np.random.default_rng(seed)
    root = str(root)
    os.makedirs(os.path.join(root, "model"), exist_ok=True)
    np.savetxt(os.path.join(root, "model", "points.xyz"),
               np.round(rng.normal(size=(500, 3)) * 0.05, 6))
    recs = []
    for i in range(n_pairs):
        p = lambda s: os.path.join(root, "%06d-%s" % (i, s))  # noqa: E731
        label = np.zeros((H, W), np.uint8)
        label[100 + 10 * i:300, 150:400 + 20 * i] = 1 + i
        label[20:60, 30:90] = 7  # another instance
        for name in ("color_obs", "color_rend"):
            cv2.imwrite(p(name + ".png"), rng.integers(0, 256, size=(H, W, 3), dtype=np.uint8))
        for name in ("depth_obs", "depth_rend"):
            d = rng.integers(400, 3000, size=(H, W)).astype(np.uint16)
            d[rng.random((H, W)) < 0.5] = 0
            cv2.imwrite(p(name + ".png"), d)
        cv2.imwrite(p("label.png"), label)
        pose = random_pose_pair(rng)[0]
        with open(p("pose.txt"), "w") as f:
            f.write("%d\n" % (1 + i))
            np.savetxt(f, pose)
        recs.append({"image_observed": p("color_obs.png"), "image_rendered": p("color_rend.png"),
                     "depth_observed": p("depth_obs.png"), "depth_gt_observed": p("depth_obs.png"),
                     "depth_rendered": p("depth_rend.png"), "mask_gt_observed": p("label.png"),
                     "mask_idx": 1 + i, "pose_observed": p("pose.txt"), "img_flipped": False})
    return recs


def make_pose_mode_inputs(seed, batch=12):
    """float32 blobs of the updater's pose block: src / tgt poses and unnormalised quaternion +
    translation 'predictions'."""
    rng = np.random.default_rng(seed)
    src = np.zeros((batch, 3, 4), np.float32)
    tgt = np.zeros((batch, 3, 4), np.float32)
    for b in range(batch):
        src[b], tgt[b] = random_pose_pair(rng)
    rot = rng.normal(size=(batch, 4)).astype(np.float32)
    rot[:, 0] = np.abs(rot[:, 0]) + 2.0
    trans = rng.normal(0, 0.05, size=(batch, 3)).astype(np.float32)
    return dict(src_pose=src, tgt_pose=tgt, rot_est=rot, trans_est=trans)
