"""Host-side pose algebra of the render-and-compare path (numpy, float64).

The reference keeps this part on the host too
(/root/reference/lib/pair_matching/RT_transform.py, lib/utils/projection.py);
it is B x (3 x 4) scalars per iteration.  Only what the path's callers need is
provided: composing KT for the flow kernel and applying / inverting the
network's pose update.  The batched on-device versions live in pose.cu.
"""
from __future__ import annotations

import numpy as np


def se3_inverse(RT):
    """RT^-1 for RT = [R|T] (3x4); float32 result like lib/utils/projection.py:12-23."""
    R = np.asarray(RT)[:3, :3]
    T = np.asarray(RT)[:3, 3]
    out = np.zeros((3, 4), dtype=np.float32)
    out[:, :3] = R.T
    out[:, 3] = -R.T.dot(T)
    return out


def se3_mul(RT1, RT2):
    """RT1 o RT2; float32 result like lib/utils/projection.py:26-43."""
    RT1, RT2 = np.asarray(RT1), np.asarray(RT2)
    out = np.zeros((3, 4), dtype=np.float32)
    out[:, :3] = RT1[:, :3].dot(RT2[:, :3])
    out[:, 3] = RT1[:, :3].dot(RT2[:, 3]) + RT1[:, 3]
    return out


def calc_se3(pose_src, pose_tgt):
    """Relative transform src -> tgt (RT_transform.py:186-197)."""
    rel = se3_mul(pose_tgt, se3_inverse(pose_src))
    return rel[:, :3], rel[:, 3].reshape(3)


def make_KT(K, pose_src, pose_tgt):
    """K . [R_rel | t_rel], float32 (batch_updater_py_multi.py:306-312, :343)."""
    R, t = calc_se3(pose_src, pose_tgt)
    m = np.zeros((3, 4))
    m[:, :3] = R
    m[:, 3] = t
    return np.dot(np.asarray(K, dtype=np.float64), m).astype(np.float32)


def make_Kinv(K):
    """batch_updater_py_multi.py:70, :344."""
    return np.linalg.inv(np.asarray(K, dtype=np.float64)).astype(np.float32)


def quat2mat(q):
    """w,x,y,z -> 3x3 (RT_transform.py:393-443)."""
    w, x, y, z = [float(v) for v in q]
    Nq = w * w + x * x + y * y + z * z
    if Nq < np.finfo(np.float64).eps:
        return np.eye(3)
    s = 2.0 / Nq
    X, Y, Z = x * s, y * s, z * s
    return np.array([
        [1.0 - (y * Y + z * Z), x * Y - w * Z, x * Z + w * Y],
        [x * Y + w * Z, 1.0 - (x * X + z * Z), y * Z - w * X],
        [x * Z - w * Y, y * Z + w * X, 1.0 - (x * X + y * Y)],
    ])


def mat2quat(M):
    """3x3 -> w,x,y,z with w >= 0 (RT_transform.py:446-523)."""
    Qxx, Qyx, Qzx, Qxy, Qyy, Qzy, Qxz, Qyz, Qzz = np.asarray(M, dtype=np.float64).flat
    Km = np.array([
        [Qxx - Qyy - Qzz, 0, 0, 0],
        [Qyx + Qxy, Qyy - Qxx - Qzz, 0, 0],
        [Qzx + Qxz, Qzy + Qyz, Qzz - Qxx - Qyy, 0],
        [Qyz - Qzy, Qzx - Qxz, Qxy - Qyx, Qxx + Qyy + Qzz],
    ]) / 3.0
    vals, vecs = np.linalg.eigh(Km)
    q = vecs[[3, 0, 1, 2], np.argmax(vals)]
    return -q if q[0] < 0 else q


def RT_transform(pose_src, r, t, T_means, T_stds, rot_coord="CAMERA"):
    """Apply a predicted (quaternion, translation-delta) update (RT_transform.py:135-161)."""
    pose_src = np.asarray(pose_src, dtype=np.float64)
    r = np.squeeze(np.asarray(r, dtype=np.float64))
    Rd = quat2mat(r / np.linalg.norm(r))
    rc = rot_coord.lower()
    out = np.zeros((3, 4))
    out[:, :3] = pose_src[:, :3].dot(Rd) if rc == "model" else Rd.dot(pose_src[:, :3])
    td = np.squeeze(np.asarray(t, dtype=np.float64)) * np.asarray(T_stds) + np.asarray(T_means)
    Ts = pose_src[:, 3]
    z2 = Ts[2] / np.exp(td[2])
    if rc == "camera_new":
        out[:2, 3] = Ts[2] * td[:2] + Ts[:2]
    else:
        out[:2, 3] = z2 * (td[:2] + Ts[:2] / Ts[2])
    out[2, 3] = z2
    return out


def calc_RT_delta(pose_src, pose_tgt, T_means, T_stds, rot_coord="CAMERA"):
    """Regression labels (quaternion, translation delta) src -> tgt (RT_transform.py:16-48)."""
    pose_src = np.asarray(pose_src, dtype=np.float64)
    pose_tgt = np.asarray(pose_tgt, dtype=np.float64)
    rc = rot_coord.lower()
    if rc == "model":
        Rd = pose_src[:, :3].T.dot(pose_tgt[:, :3])
    else:
        Rd = pose_tgt[:, :3].dot(pose_src[:, :3].T)
    Ts, Tt = pose_src[:, 3], pose_tgt[:, 3]
    td = np.zeros(3)
    if rc == "camera_new":
        td[:2] = (Tt[:2] - Ts[:2]) / Ts[2]
    else:
        td[:2] = Tt[:2] / Tt[2] - Ts[:2] / Ts[2]
    td[2] = np.log(Ts[2] / Tt[2])
    return mat2quat(Rd), (td - np.asarray(T_means)) / np.asarray(T_stds)
