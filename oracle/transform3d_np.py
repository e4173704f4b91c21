"""oracle/transform3d_np.py -- TEST INFRASTRUCTURE ONLY (never on the product path).

numpy restatement of
  /root/reference/deepim/operator_py/transform3d.py:42-118   forward
  /root/reference/deepim/operator_py/transform3d.py:120-191  backward
  /root/reference/deepim/operator_py/transform3d.py:193-225  T_transform_backward
  /root/reference/deepim/operator_py/transform3d.py:227-327  quat2mat_forward / _backward
  /root/reference/deepim/symbols/deepIM_flownet.py:460-499   point-matching loss graph
  /root/reference/deepim/core/metric.py:108-120              PointMatchingLossMetric

MXNet NDArrays are float32; `.asnumpy()` scalars are np.float32 and follow the
reference's numpy (1.x) promotion: np.float32 with a Python scalar -> float64.

Pinned by the reference's own self-test recipe (transform3d.py:365-638: seed 1,
B = 8, N = 3000, forward < 1e-4 against RT_transform + matmul, central
differences with step 1e-2 and threshold 5e-3), re-hosted on synthetic points
in tests/test_oracle_transform3d.py.
"""
from __future__ import annotations

import numpy as np

from . import pose_np

f32 = np.float32
f64 = np.float64


def quat2mat_forward(q):
    """transform3d.py:227-254.  q float32 [4] -> float32 [3,3]."""
    w, x, y, z = [f32(v) for v in q]
    Nq = f32(f32(f32(w * w) + f32(x * x)) + f32(y * y)) + f32(z * z)
    Nq = f32(Nq)
    if not (-1e-2 < f64(Nq) - 1 < 1e-2):
        return np.eye(3, dtype=f32)
    s = 2.0 / f64(Nq)
    w, x, y, z = f64(w), f64(x), f64(y), f64(z)
    X, Y, Z = x * s, y * s, z * s
    wX, wY, wZ = w * X, w * Y, w * Z
    xX, xY, xZ = x * X, x * Y, x * Z
    yY, yZ, zZ = y * Y, y * Z, z * Z
    return np.array([
        [1.0 - (yY + zZ), xY - wZ, xZ + wY],
        [xY + wZ, 1.0 - (xX + zZ), yZ - wX],
        [xZ - wY, yZ + wX, 1.0 - (xX + yY)],
    ]).astype(f32)


def _t_tgt(T_src, T_delta, T_means, T_stds, Rm_delta, rot_coord):
    rc = rot_coord.lower()
    if rc == "naive":
        return pose_np.T_transform_naive(Rm_delta, T_src, T_delta).astype(f32)
    return pose_np.T_transform(T_src, T_delta, T_means, T_stds, rot_coord).astype(f32)


def forward(points, rotation, translation, pose_src, T_means, T_stds, rot_coord="CAMERA"):
    """transform3dOperator.forward.  points [B,3,N] (or any [B,3,...]); returns the same shape."""
    points = np.asarray(points, f32)
    B = points.shape[0]
    bottom = points.reshape(B, 3, -1)
    rotation = np.asarray(rotation, f32)
    translation = np.asarray(translation, f32)
    pose_src = np.asarray(pose_src, f32)
    T_means = np.asarray(T_means, f32)
    T_stds = np.asarray(T_stds, f32)
    out = np.empty_like(bottom)
    for b in range(B):
        Rm_delta = quat2mat_forward(rotation[b])
        Rm_tgt = pose_np.R_transform(pose_src[b, :, :3], Rm_delta, rot_coord).astype(f32)
        T_tgt = _t_tgt(pose_src[b, :, 3], translation[b], T_means, T_stds, Rm_delta, rot_coord)
        out[b] = (np.matmul(Rm_tgt.astype(f64), bottom[b].astype(f64)) +
                  T_tgt.astype(f64)[:, None]).astype(f32)
    return out.reshape(points.shape)


def T_transform_backward(D, T_src, T_delta, T_means, T_stds, rot_coord):
    """transform3d.py:193-225 (float32 NDArray arithmetic)."""
    D, T_src, T_delta = np.asarray(D, f32), np.asarray(T_src, f32), np.asarray(T_delta, f32)
    T_means, T_stds = np.asarray(T_means, f32), np.asarray(T_stds, f32)
    assert T_src[2] != 0
    T_delta_1 = T_delta * T_stds + T_means
    z2 = T_src[2] / np.exp(T_delta_1[2])
    out = np.zeros((3,), f32)
    rc = rot_coord.lower()
    if rc in ("camera", "model"):
        out[0] = D[0] * (T_stds[0] * z2)
        out[1] = D[1] * (T_stds[1] * z2)
        share = -T_stds[2] * z2
        out[2] = (D[0] * (share * (T_delta_1[0] + T_src[0] / T_src[2])) +
                  D[1] * (share * (T_delta_1[1] + T_src[1] / T_src[2])) +
                  D[2] * (-T_stds[2] * z2))
    elif rc == "camera_new":
        out[0] = D[0] * (T_stds[0] * T_src[2])
        out[1] = D[1] * (T_stds[1] * T_src[2])
        out[2] = D[2] * (-T_stds[2] * z2)
    return out


def quat2mat_backward(mat_diff, q):
    """transform3d.py:256-327."""
    w, x, y, z = [f32(v) for v in q]
    Nq = f32(f32(f32(f32(w * w) + f32(x * x)) + f32(y * y)) + f32(z * z))
    Nq_sqrt = np.sqrt(Nq)
    w_, x_, y_, z_ = [f64(v) for v in (np.asarray(q, f32) / Nq_sqrt)]
    s = 2.0
    if not (-1e-4 < f64(Nq) - 1 < 1e-4):
        return np.zeros(4, f32)
    D = np.asarray(mat_diff, f32).astype(f64)
    w_d = (-z_ * D[0, 1] + y_ * D[0, 2] + z_ * D[1, 0] - x_ * D[1, 2] - y_ * D[2, 0] + x_ * D[2, 1]) * s
    x_d = (y_ * D[0, 1] + z_ * D[0, 2] + y_ * D[1, 0] - 2 * x_ * D[1, 1] - w_ * D[1, 2] +
           z_ * D[2, 0] + w_ * D[2, 1] - 2 * x_ * D[2, 2]) * s
    y_d = (-2 * y_ * D[0, 0] + x_ * D[0, 1] + w_ * D[0, 2] + x_ * D[1, 0] + z_ * D[1, 2] -
           w_ * D[2, 0] + z_ * D[2, 1] - 2 * y_ * D[2, 2]) * s
    z_d = (-2 * z_ * D[0, 0] - w_ * D[0, 1] + x_ * D[0, 2] + w_ * D[1, 0] - 2 * z_ * D[1, 1] +
           y_ * D[1, 2] + x_ * D[2, 0] + y_ * D[2, 1]) * s
    ns = f64(Nq_sqrt)
    w, x, y, z = f64(w), f64(x), f64(y), f64(z)
    share = ns ** 3 * (w * w_d + x * x_d + y * y_d + z * z_d)
    return np.array([ns * w_d - w * share, ns * x_d - x * share, ns * y_d - y * share,
                     ns * z_d - z * share]).astype(f32)


def backward(out_grad, points, rotation, translation, pose_src, T_means, T_stds,
             rot_coord="CAMERA"):
    """transform3dOperator.backward -> (grad_rotation [B,4], grad_translation [B,3])."""
    points = np.asarray(points, f32)
    B = points.shape[0]
    X = points.reshape(B, 3, -1).astype(f64)
    G = np.asarray(out_grad, f32).reshape(B, 3, -1).astype(f64)
    rotation = np.asarray(rotation, f32)
    translation = np.asarray(translation, f32)
    pose_src = np.asarray(pose_src, f32)
    rc = rot_coord.lower()
    g_rot = np.zeros((B, 4), f32)
    g_tr = np.zeros((B, 3), f32)
    for b in range(B):
        Rm_src = pose_src[b, :, :3].astype(f64)
        T_src = pose_src[b, :, 3]
        T_tgt_diff = G[b].sum(axis=1).astype(f32)
        if rc == "naive":
            g_tr[b] = T_tgt_diff
        else:
            g_tr[b] = T_transform_backward(T_tgt_diff, T_src, translation[b], T_means, T_stds, rot_coord)
        if rc == "naive":
            src_pts = np.matmul(Rm_src, X[b]) + T_src.astype(f64)[:, None]
            Rm_delta_diff = np.matmul(G[b], src_pts.astype(f32).astype(f64).T)
        else:
            Rm_tgt_diff = np.matmul(G[b], X[b].T).astype(f32).astype(f64)
            if rc == "model":
                Rm_delta_diff = np.matmul(Rm_src.T, Rm_tgt_diff)
            elif rc in ("camera", "camera_new"):
                Rm_delta_diff = np.matmul(Rm_tgt_diff, Rm_src.T)
            else:
                raise Exception("Unknown rot_coord in transform3d operator: {}".format(rot_coord))
        g_rot[b] = quat2mat_backward(Rm_delta_diff.astype(f32), rotation[b])
    return g_rot, g_tr


# ------------------------------------------------------------- loss graph
def _loss_f(x, loss_type, sigma):
    lt = loss_type.lower()
    if lt == "l1":
        return np.abs(x)
    if lt == "l2":
        return np.square(x)
    if lt == "smooth_l1":
        s2 = f32(sigma) * f32(sigma)
        return np.where(np.abs(x) < 1.0 / s2, 0.5 * s2 * x * x, np.abs(x) - 0.5 / s2).astype(f32)
    raise Exception("Unknown Point Matching Loss Type: {}".format(loss_type))


def _loss_df(x, loss_type, sigma):
    lt = loss_type.lower()
    if lt == "l1":
        return np.sign(x)
    if lt == "l2":
        return 2 * x
    s2 = f32(sigma) * f32(sigma)
    return np.where(np.abs(x) < 1.0 / s2, s2 * x, np.sign(x)).astype(f32)


def pm_loss_forward(points, points_obs, weights, rotation, translation, pose_src, T_means, T_stds,
                    rot_coord="CAMERA", loss_type="L1", norm_term=0.1, sl1_scalar=1.0):
    """deepIM_flownet.py:446-486.  Returns (loss_map [B,3,N], est [B,3,N], per-sample sums [B])."""
    est = forward(points, rotation, translation, pose_src, T_means, T_stds, rot_coord)
    x = ((est - np.asarray(points_obs, f32)).astype(f32) / f32(norm_term)).astype(f32)
    loss = (np.asarray(weights, f32) * _loss_f(x, loss_type, sl1_scalar).astype(f32)).astype(f32)
    sums = loss.reshape(loss.shape[0], -1).astype(f64).sum(axis=1)
    return loss, est, sums


def pm_loss_backward(points, points_obs, weights, rotation, translation, pose_src, T_means,
                     T_stds, grad_scale, rot_coord="CAMERA", loss_type="L1", norm_term=0.1,
                     sl1_scalar=1.0):
    """MakeLoss(grad_scale) head gradient (:493-497) pushed through the loss graph and
    transform3dOperator.backward."""
    est = forward(points, rotation, translation, pose_src, T_means, T_stds, rot_coord)
    x = ((est - np.asarray(points_obs, f32)).astype(f32) / f32(norm_term)).astype(f32)
    g = (f32(grad_scale) * np.asarray(weights, f32)).astype(f32)
    g = (g * _loss_df(x, loss_type, sl1_scalar).astype(f32)).astype(f32)
    g = (g / f32(norm_term)).astype(f32)
    return backward(g, points, rotation, translation, pose_src, T_means, T_stds, rot_coord)


def pm_metric(loss_maps, num_3d_sample):
    """PointMatchingLossMetric (metric.py:108-120): sum / NUM_3D_SAMPLE per update."""
    return float(np.sum(loss_maps)) / float(num_3d_sample)
