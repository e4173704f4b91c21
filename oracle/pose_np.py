"""oracle/pose_np.py -- TEST INFRASTRUCTURE ONLY (never on the product path).

numpy restatement of the pose algebra on the render-and-compare path:
  /root/reference/lib/pair_matching/RT_transform.py:16-48    calc_RT_delta
  /root/reference/lib/pair_matching/RT_transform.py:51-79    R_transform, R_inv_transform
  /root/reference/lib/pair_matching/RT_transform.py:82-132   T_transform(_naive), T_inv_transform
  /root/reference/lib/pair_matching/RT_transform.py:135-161  RT_transform
  /root/reference/lib/pair_matching/RT_transform.py:186-197  calc_se3
  /root/reference/lib/pair_matching/RT_transform.py:393-523  quat2mat, mat2quat
  /root/reference/lib/utils/projection.py:12-43              se3_inverse, se3_mul (float32 outputs)

The reference module itself does not import under numpy >= 2
(RT_transform.py:246-247 use np.maximum_sctype / np.float).  With a two-line
shim it does; tests/golden/make_golden.py imports it that way in the build
container and stores its outputs in tests/golden/pose_*.npz, which pin this
restatement (tests/test_oracle_cpu.py, tests/test_golden_ops_cpu.py).  Docstring known answers of the
reference (RT_transform.py:409-414, :478-483) are checked there too.
"""
from __future__ import annotations

import numpy as np

_FLOAT_EPS = np.finfo(np.float64).eps  # RT_transform.py:247 (np.finfo(np.float).eps)


def se3_inverse(RT):
    """projection.py:12-23."""
    R = RT[0:3, 0:3]
    T = RT[0:3, 3].reshape((3, 1))
    RT_new = np.zeros((3, 4), dtype=np.float32)
    RT_new[0:3, 0:3] = R.transpose()
    RT_new[0:3, 3] = -1 * np.dot(R.transpose(), T).reshape((3))
    return RT_new


def se3_mul(RT1, RT2):
    """projection.py:26-43."""
    R1 = RT1[0:3, 0:3]
    T1 = RT1[0:3, 3].reshape((3, 1))
    R2 = RT2[0:3, 0:3]
    T2 = RT2[0:3, 3].reshape((3, 1))
    RT_new = np.zeros((3, 4), dtype=np.float32)
    RT_new[0:3, 0:3] = np.dot(R1, R2)
    T_new = np.dot(R1, T2) + T1
    RT_new[0:3, 3] = T_new.reshape((3))
    return RT_new


def quat2mat(q):
    """RT_transform.py:393-443 (w, x, y, z)."""
    w, x, y, z = [np.float64(v) for v in q]
    Nq = w * w + x * x + y * y + z * z
    if Nq < _FLOAT_EPS:
        return np.eye(3)
    s = 2.0 / Nq
    X, Y, Z = x * s, y * s, z * s
    wX, wY, wZ = w * X, w * Y, w * Z
    xX, xY, xZ = x * X, x * Y, x * Z
    yY, yZ, zZ = y * Y, y * Z, z * Z
    return np.array([
        [1.0 - (yY + zZ), xY - wZ, xZ + wY],
        [xY + wZ, 1.0 - (xX + zZ), yZ - wX],
        [xZ - wY, yZ + wX, 1.0 - (xX + yY)],
    ])


def mat2quat(M):
    """RT_transform.py:446-523 (Bar-Itzhack; largest eigenvector of K, w >= 0)."""
    Qxx, Qyx, Qzx, Qxy, Qyy, Qzy, Qxz, Qyz, Qzz = np.asarray(M, dtype=np.float64).flat
    K = np.array([
        [Qxx - Qyy - Qzz, 0, 0, 0],
        [Qyx + Qxy, Qyy - Qxx - Qzz, 0, 0],
        [Qzx + Qxz, Qzy + Qyz, Qzz - Qxx - Qyy, 0],
        [Qyz - Qzy, Qzx - Qxz, Qxy - Qyx, Qxx + Qyy + Qzz],
    ]) / 3.0
    vals, vecs = np.linalg.eigh(K)
    q = vecs[[3, 0, 1, 2], np.argmax(vals)]
    if q[0] < 0:
        q *= -1
    return q


def R_transform(R_src, R_delta, rot_coord="MODEL"):
    """RT_transform.py:51-69."""
    rc = rot_coord.lower()
    if rc == "model":
        return np.dot(R_src, R_delta)
    if rc in ("camera", "naive", "camera_new"):
        return np.dot(R_delta, R_src)
    raise Exception("Unknown rot_coord in R_transform: {}".format(rot_coord))


def R_inv_transform(R_src, R_tgt, rot_coord):
    """RT_transform.py:72-79."""
    rc = rot_coord.lower()
    if rc == "model":
        return np.dot(R_src.transpose(), R_tgt)
    if rc in ("camera", "camera_new"):
        return np.dot(R_tgt, R_src.transpose())
    raise Exception("Unknown rot_coord in R_inv_transform: {}".format(rot_coord))


def T_transform(T_src, T_delta, T_means, T_stds, rot_coord):
    """RT_transform.py:82-103."""
    assert T_src[2] != 0, "T_src: {}".format(T_src)
    T_delta_1 = T_delta * T_stds + T_means
    T_tgt = np.zeros((3,))
    z2 = T_src[2] / np.exp(T_delta_1[2])
    T_tgt[2] = z2
    rc = rot_coord.lower()
    if rc in ("camera", "model"):
        T_tgt[0] = z2 * (T_delta_1[0] + T_src[0] / T_src[2])
        T_tgt[1] = z2 * (T_delta_1[1] + T_src[1] / T_src[2])
    elif rc == "camera_new":
        T_tgt[0] = T_src[2] * T_delta_1[0] + T_src[0]
        T_tgt[1] = T_src[2] * T_delta_1[1] + T_src[1]
    else:
        raise Exception("Unknown: {}".format(rot_coord))
    return T_tgt


def T_transform_naive(R_delta, T_src, T_delta):
    """RT_transform.py:106-110."""
    T_new = np.dot(R_delta, np.reshape(T_src, (3, 1))) + np.reshape(T_delta, (3, 1))
    return T_new.reshape((3,))


def T_inv_transform(T_src, T_tgt, T_means, T_stds, rot_coord):
    """RT_transform.py:113-132."""
    T_delta = np.zeros((3,))
    rc = rot_coord.lower()
    if rc == "camera_new":
        T_delta[0] = (T_tgt[0] - T_src[0]) / T_src[2]
        T_delta[1] = (T_tgt[1] - T_src[1]) / T_src[2]
    elif rc in ("camera", "model"):
        T_delta[0] = T_tgt[0] / T_tgt[2] - T_src[0] / T_src[2]
        T_delta[1] = T_tgt[1] / T_tgt[2] - T_src[1] / T_src[2]
    else:
        raise Exception("Unknown: {}".format(rot_coord))
    T_delta[2] = np.log(T_src[2] / T_tgt[2])
    return (T_delta - T_means) / T_stds


def RT_transform(pose_src, r, t, T_means, T_stds, rot_coord="MODEL"):
    """RT_transform.py:135-161 (quaternion branch; the euler branch is off-path)."""
    r = np.squeeze(r)
    if r.shape[0] != 4:
        raise Exception("Unknown r shape: {}".format(r.shape))
    quat = r / np.linalg.norm(r)
    Rm_delta = quat2mat(quat)
    t_delta = np.squeeze(t)
    if rot_coord.lower() == "naive":
        se3_mx = np.zeros((3, 4))
        se3_mx[:, :3] = Rm_delta
        se3_mx[:, 3] = t
        return se3_mul(se3_mx, pose_src)
    pose_est = np.zeros((3, 4))
    pose_est[:3, :3] = R_transform(pose_src[:3, :3], Rm_delta, rot_coord)
    pose_est[:3, 3] = T_transform(pose_src[:, 3], t_delta, T_means, T_stds, rot_coord)
    return pose_est


def calc_RT_delta(pose_src, pose_tgt, T_means, T_stds, rot_coord="MODEL", rot_type="MATRIX"):
    """RT_transform.py:16-48."""
    if rot_coord.lower() == "naive":
        se3_src2tgt = se3_mul(pose_tgt, se3_inverse(pose_src))
        Rm_delta = se3_src2tgt[:, :3]
        T_delta = se3_src2tgt[:, 3].reshape((3))
    else:
        Rm_delta = R_inv_transform(pose_src[:3, :3], pose_tgt[:3, :3], rot_coord)
        T_delta = T_inv_transform(pose_src[:, 3], pose_tgt[:, 3], T_means, T_stds, rot_coord)
    if rot_type.lower() == "quat":
        r = mat2quat(Rm_delta)
    elif rot_type.lower() == "matrix":
        r = Rm_delta
    else:
        raise Exception("Unknown rot_type: {}".format(rot_type))
    return r, np.squeeze(T_delta)


def calc_se3(pose_src, pose_tgt):
    """RT_transform.py:186-197."""
    se3_src2tgt = se3_mul(pose_tgt, se3_inverse(pose_src))
    return se3_src2tgt[:, :3], se3_src2tgt[:, 3].reshape((3))


def make_KT(K, pose_refined, pose_tgt):
    """batch_updater_py_multi.py:306-312: KT = K . [R_rel | t_rel] (float64; cast to float32 at :343)."""
    se3_m = np.zeros([3, 4])
    se3_rotm, se3_t = calc_se3(pose_refined, pose_tgt)
    se3_m[:, :3] = se3_rotm
    se3_m[:, 3] = se3_t
    return np.dot(K, se3_m)
