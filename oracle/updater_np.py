"""oracle/updater_np.py -- TEST INFRASTRUCTURE ONLY (never on the product path).

numpy restatement of the host driver between refinement iterations
  /root/reference/lib/pair_matching/batch_updater_py_multi.py:100-387 (forward)
for the LINEMOD branch: pose update -> render -> image post -> new labels -> KT -> gpu_flow ->
flow_weights tile -> mask_rendered.  Rendering uses oracle/raster_np (GL restatement), flow uses
oracle/flow_np (flow_kernel restatement).
"""
from __future__ import annotations

import numpy as np

from . import flow_np, pose_np, raster_np


def _quat2mat_f32(q):
    """RT_transform.quat2mat (:393-443) on np.float32 scalars under numpy-1.x promotion: Nq is
    accumulated in float32, `2.0 / Nq` and everything after it is float64."""
    w, x, y, z = [np.float32(v) for v in q]
    Nq = np.float32(np.float32(np.float32(np.float32(w * w) + np.float32(x * x)) + np.float32(y * y)) +
                    np.float32(z * z))
    if Nq < pose_np._FLOAT_EPS:
        return np.eye(3)
    s = 2.0 / np.float64(Nq)
    w, x, y, z = np.float64(w), np.float64(x), np.float64(y), np.float64(z)
    X, Y, Z = x * s, y * s, z * s
    return np.array([[1.0 - (y * Y + z * Z), x * Y - w * Z, x * Z + w * Y],
                     [x * Y + w * Z, 1.0 - (x * X + z * Z), y * Z - w * X],
                     [x * Z - w * Y, y * Z + w * X, 1.0 - (x * X + y * Y)]])


def pose_update(src_pose, rot_est, trans_est, tgt_pose, K, T_means, T_stds, rot_coord="CAMERA"):
    """batch_updater_py_multi.py:211-312 without the render call.  float32 inputs as the
    `.asnumpy()` blobs are; returns float64 arrays as the reference holds them before nd.array."""
    B = src_pose.shape[0]
    refined = np.zeros((B, 3, 4))
    rot_res = np.zeros((B, 4))
    trans_res = np.zeros((B, 3))
    KT = np.zeros((B, 3, 4))
    for b in range(B):
        pre_pose = np.squeeze(src_pose[b])
        r = np.squeeze(rot_est[b]).astype(np.float32)
        quat = (r / np.float32(np.sqrt(np.sum(r * r, dtype=np.float32)))).astype(np.float32)
        Rd = _quat2mat_f32(quat)
        pose = np.zeros((3, 4))
        pose[:, :3] = pose_np.R_transform(pre_pose[:, :3], Rd, rot_coord)
        pose[:, 3] = pose_np.T_transform(pre_pose[:, 3].astype(np.float32),
                                         np.squeeze(trans_est[b]).astype(np.float32),
                                         np.asarray(T_means, np.float32),
                                         np.asarray(T_stds, np.float32), rot_coord)
        refined[b] = pose
        rr, tt = pose_np.calc_RT_delta(pose, np.squeeze(tgt_pose[b]), T_means, T_stds, rot_coord, "QUAT")
        rot_res[b], trans_res[b] = rr, tt
        KT[b] = pose_np.make_KT(K, pose, np.squeeze(tgt_pose[b]))
    return refined, rot_res, trans_res, KT


def updater_forward(meshes, class_index, src_pose, tgt_pose, rot_est, trans_est, depth_gt_observed,
                    K, pixel_means_bgr_order, T_means, T_stds, rot_coord="CAMERA", zNear=0.25,
                    zFar=6.0, flow_threads=0):
    """The update_package of batch_updater_py_multi.py:322-356 (LINEMOD branch)."""
    B = src_pose.shape[0]
    H, W = depth_gt_observed.shape[2:]
    refined, rot_res, trans_res, KT = pose_update(src_pose, rot_est, trans_est, tgt_pose, K, T_means,
                                                  T_stds, rot_coord)
    image = np.zeros((B, 3, H, W))
    depth = np.zeros((B, 1, H, W))
    pm = np.asarray(pixel_means_bgr_order, np.float32).reshape(3, 1, 1)
    for b in range(B):
        bgr, d = raster_np.render_py(meshes[int(class_index[b])], refined[b, :, :3], refined[b, :, 3], K,
                                     r_type="mat", width=W, height=H, zNear=zNear, zFar=zFar)
        img = bgr[:, :, [2, 1, 0]].transpose([2, 0, 1]).astype(np.float32)
        image[b] = img - pm
        depth[b, 0] = d
    mask = np.zeros(depth.shape)
    mask[depth > 0.2] = 1
    Kinv = np.array(np.linalg.inv(np.asarray(K, np.float64)))
    flow, valid = flow_np.flow_kernel_cpu(depth.astype(np.float32), depth_gt_observed.astype(np.float32),
                                          KT.astype(np.float32), Kinv.astype(np.float32),
                                          num_threads=flow_threads)
    return {"image_rendered": image.astype(np.float32), "depth_rendered": depth.astype(np.float32),
            "src_pose": refined.astype(np.float32), "rot": rot_res.astype(np.float32),
            "trans": trans_res.astype(np.float32), "flow": flow,
            "flow_weights": np.tile(valid, [1, 2, 1, 1]), "mask_rendered": mask.astype(np.float32),
            "KT": KT.astype(np.float32)}


def update_mask_observed(policy, mask_observed, mask_rendered):
    """The mask_observed branch of update_data_batch,
    /root/reference/lib/pair_matching/data_pair.py:101-134, for ONE 2-D mask pair (the
    statements are lifted; only the MXNet `nd.array` wrap is dropped).  `mask_observed` is the
    package's update_data, `mask_rendered` is package["mask_rendered"]."""
    update_data = mask_observed
    if policy == "box_rendered":
        update_data = np.copy(mask_rendered)
        out = np.zeros(update_data.shape)
        x_max = np.max(update_data, 0)
        y_max = np.max(update_data, 1)
        nz_x = np.nonzero(x_max)[0]
        nz_y = np.nonzero(y_max)[0]
        x_start = np.min(nz_x)
        x_end = np.max(nz_x)
        y_start = np.min(nz_y)
        y_end = np.max(nz_y)
        out[y_start:y_end, x_start:x_end] = 1.0  # rectangle
    elif policy == "box_observed":
        out = np.zeros(update_data.shape)
        x_max = np.max(update_data, 0)
        y_max = np.max(update_data, 1)
        nz_x = np.nonzero(x_max)[0]
        nz_y = np.nonzero(y_max)[0]
        if len(nz_x) == 0 or len(nz_y) == 0:
            raise Exception("no point valid in mask_observed")
        x_start = np.min(nz_x)
        x_end = np.max(nz_x)
        y_start = np.min(nz_y)
        y_end = np.max(nz_y)
        out[y_start:y_end, x_start:x_end] = 1.0  # rectangle
    else:
        out = update_data
    return out
