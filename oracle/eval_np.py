"""oracle/eval_np.py -- TEST INFRASTRUCTURE ONLY (never on the product path).

numpy restatement of the evaluation reductions that follow the render-and-compare path:
  /root/reference/lib/utils/pose_error.py:20-147   transform_pts_Rt(_2d), arp_2d, add, adi, re, te
  /root/reference/lib/utils/visibility.py:7-33     estimate_visib_mask(_gt/_est)
  /root/reference/deepim/core/tester.py:719-736    calc_EPE_one_pair
Pinned by tests/golden/eval_reference.npz (outputs of the reference's own pose_error.py and
visibility.py, imported unmodified by tests/golden/make_golden_eval.py).
"""
from __future__ import annotations

import numpy as np
from scipy import spatial
from scipy.linalg import logm


def transform_pts_Rt(pts, R, t):  # pose_error.py:20-32
    return (R.dot(pts.T) + t.reshape((3, 1))).T


def transform_pts_Rt_2d(pts, R, t, K):  # :35-52
    pts_t = R.dot(pts.T) + t.reshape((3, 1))
    pts_c_t = K.dot(pts_t)
    out = np.zeros((pts.shape[0], 2))
    out[:, 0] = pts_c_t[0, :] / pts_c_t[2, :]
    out[:, 1] = pts_c_t[1, :] / pts_c_t[2, :]
    return out


def arp_2d(R_est, t_est, R_gt, t_gt, pts, K):  # :55-68
    return np.linalg.norm(transform_pts_Rt_2d(pts, R_est, t_est, K) -
                          transform_pts_Rt_2d(pts, R_gt, t_gt, K), axis=1).mean()


def add(R_est, t_est, R_gt, t_gt, pts):  # :71-86
    return np.linalg.norm(transform_pts_Rt(pts, R_est, t_est) - transform_pts_Rt(pts, R_gt, t_gt),
                          axis=1).mean()


def adi(R_est, t_est, R_gt, t_gt, pts):  # :89-109
    pts_est = transform_pts_Rt(pts, R_est, t_est)
    pts_gt = transform_pts_Rt(pts, R_gt, t_gt)
    nn_dists, _ = spatial.cKDTree(pts_est).query(pts_gt, k=1)
    return nn_dists.mean()


def re(R_est, R_gt):  # :126-131
    temp = logm(np.dot(np.transpose(R_est), R_gt))
    return np.linalg.norm(temp, "fro") / np.sqrt(2) / np.pi * 180


def te(t_est, t_gt):  # :134-145
    return np.linalg.norm(t_gt - t_est)


def estimate_visib_mask(d_test, d_model, delta):  # visibility.py:7-22
    mask_valid = np.logical_and(d_test > 0, d_model > 0)
    d_diff = d_model.astype(np.float32) - d_test.astype(np.float32)
    return np.logical_and(d_diff <= delta, mask_valid)


def estimate_visib_mask_est(d_test, d_est, visib_gt, delta):  # :30-33
    visib_est = estimate_visib_mask(d_test, d_est, delta)
    return np.logical_or(visib_est, np.logical_and(visib_gt, d_est > 0))


def calc_EPE_one_pair(flow_pred, flow_gt, visible, bg):  # tester.py:719-736, one flow type
    x_diff = flow_gt[:, :, 0] - flow_pred[:, :, 0]
    y_diff = flow_gt[:, :, 1] - flow_pred[:, :, 1]
    point_diff = np.sqrt(np.square(x_diff) + np.square(y_diff))
    return {"epe_all": point_diff.sum(), "num_all": point_diff.size,
            "epe_viz": point_diff[visible == 1].sum(), "num_viz": visible.sum(),
            "epe_vizbg": point_diff[np.logical_or(visible, bg)].sum(),
            "num_vizbg": np.logical_or(visible, bg).sum()}
