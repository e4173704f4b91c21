"""Evaluation reductions on the device, behind the reference's names:

    add, adi, arp_2d, re, te            /root/reference/lib/utils/pose_error.py:55-147
    estimate_visib_mask(_gt/_est)       /root/reference/lib/utils/visibility.py:7-33
    calc_EPE_one_pair                   /root/reference/deepim/core/tester.py:719-736

The batched forms (`pose_errors`, `visib_mask`, `flow_epe`) take CUDA tensors and do not
synchronise; the reference-named single-pair functions take numpy arrays, like the originals,
and return Python floats / numpy arrays.  No CPU fallback.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .operator_py import _common as C


def pose_errors(pose_est, pose_gt, points, K, n_valid=None, want_adi=True, stream=None):
    """pose_est, pose_gt: CUDA float32 [B,3,4]; points: CUDA float32 [B,3,N] or [3,N] (shared);
    K: 3x3; n_valid: CUDA int32 [B] or None.  Returns CUDA float64 [B,5] =
    (add, adi, arp_2d, re_deg, te)."""
    C.check_tensor("pose_est", pose_est, ndim=3)
    B = int(pose_est.shape[0])
    C.check_tensor("pose_est", pose_est, shape=(B, 3, 4))
    C.check_tensor("pose_gt", pose_gt, shape=(B, 3, 4))
    dev = pose_est.device
    shared = points.dim() == 2
    N = int(points.shape[-1])
    C.check_tensor("points", points, shape=(3, N) if shared else (B, 3, N))
    if n_valid is not None:
        C.check_tensor("n_valid", n_valid, dtype=torch.int32, shape=(B,))
    lib = _lib.load()
    out = torch.empty((B, 5), dtype=torch.float64, device=dev)
    ws = torch.empty((B * lib.rc_pose_error_chunks(N) * 3,), dtype=torch.float64, device=dev)
    Kd = C.device_const(np.asarray(K, np.float32), dev)
    with torch.cuda.device(dev):
        st = lib.rc_pose_error_dev(pose_est.data_ptr(), pose_gt.data_ptr(), points.data_ptr(),
                                   1 if shared else 0,
                                   n_valid.data_ptr() if n_valid is not None else None,
                                   Kd.data_ptr(), 1 if want_adi else 0, out.data_ptr(),
                                   ws.data_ptr(), B, N, _lib.stream_ptr(stream))
    _lib.check(st, "rc_pose_error_dev")
    return out


def visib_mask(d_test, d_model, delta, visib_gt=None, out=None, stream=None):
    """CUDA float32 tensors of equal shape -> float32 0/1 mask (estimate_visib_mask; with
    visib_gt: estimate_visib_mask_est)."""
    C.check_tensor("d_test", d_test)
    C.check_tensor("d_model", d_model, shape=d_test.shape)
    if visib_gt is not None:
        C.check_tensor("visib_gt", visib_gt, shape=d_test.shape)
    if out is None:
        out = torch.empty_like(d_test)
    C.check_tensor("out", out, shape=d_test.shape)
    with torch.cuda.device(d_test.device):
        st = _lib.load().rc_visib_mask_dev(d_test.data_ptr(), d_model.data_ptr(),
                                           visib_gt.data_ptr() if visib_gt is not None else None,
                                           float(delta), out.data_ptr(), d_test.numel(),
                                           _lib.stream_ptr(stream))
    _lib.check(st, "rc_visib_mask_dev")
    return out


def flow_epe(flow_pred, flow_gt, visible, bg=None, depth_rendered=None, stream=None):
    """flow_pred, flow_gt CUDA float32 [B,2,H,W]; visible [B,1,H,W]; bg [B,1,H,W] or derived
    from depth_rendered (tester.py:712).  Returns CUDA float64 [B,6] = epe_all, num_all,
    epe_viz, num_viz, epe_vizbg, num_vizbg."""
    C.check_tensor("flow_pred", flow_pred, ndim=4)
    B, _, H, W = flow_pred.shape
    C.check_tensor("flow_pred", flow_pred, shape=(B, 2, H, W))
    C.check_tensor("flow_gt", flow_gt, shape=(B, 2, H, W))
    C.check_tensor("visible", visible, shape=(B, 1, H, W))
    if bg is not None:
        C.check_tensor("bg", bg, shape=(B, 1, H, W))
    if depth_rendered is not None:
        C.check_tensor("depth_rendered", depth_rendered, shape=(B, 1, H, W))
    dev = flow_pred.device
    lib = _lib.load()
    out = torch.empty((B, 6), dtype=torch.float64, device=dev)
    ws = torch.empty((B * lib.rc_flow_epe_chunks() * 5,), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        st = lib.rc_flow_epe_dev(flow_pred.data_ptr(), flow_gt.data_ptr(), visible.data_ptr(),
                                 bg.data_ptr() if bg is not None else None,
                                 depth_rendered.data_ptr() if depth_rendered is not None else None,
                                 out.data_ptr(), ws.data_ptr(), B, H, W, _lib.stream_ptr(stream))
    _lib.check(st, "rc_flow_epe_dev")
    return out


# ------------------------------------------------- reference-named single-pair wrappers
def _one(R_est, t_est, R_gt, t_gt, pts, K=None, want_adi=False):
    if not torch.cuda.is_available():
        raise _lib.RenderCompareError("evaluation needs a CUDA device (no CPU fallback)")
    dev = torch.device("cuda", torch.cuda.current_device())
    pe = np.zeros((1, 3, 4), np.float32)
    pg = np.zeros((1, 3, 4), np.float32)
    pe[0, :, :3], pe[0, :, 3] = R_est, np.asarray(t_est).reshape(3)
    pg[0, :, :3], pg[0, :, 3] = R_gt, np.asarray(t_gt).reshape(3)
    pts = np.asarray(pts)
    assert pts.shape[1] == 3
    P = torch.from_numpy(np.ascontiguousarray(pts.T, np.float32)).to(dev)
    out = pose_errors(torch.from_numpy(pe).to(dev), torch.from_numpy(pg).to(dev), P,
                      np.eye(3) if K is None else K, want_adi=want_adi)
    torch.cuda.synchronize(dev)
    return out[0].cpu().numpy()


def add(R_est, t_est, R_gt, t_gt, pts):
    return float(_one(R_est, t_est, R_gt, t_gt, pts)[0])


def adi(R_est, t_est, R_gt, t_gt, pts):
    return float(_one(R_est, t_est, R_gt, t_gt, pts, want_adi=True)[1])


def arp_2d(R_est, t_est, R_gt, t_gt, pts, K):
    return float(_one(R_est, t_est, R_gt, t_gt, pts, K)[2])


def re(R_est, R_gt):
    assert np.shape(R_est) == np.shape(R_gt) == (3, 3)
    return float(_one(R_est, np.zeros(3), R_gt, np.zeros(3), np.zeros((1, 3)))[3])


def te(t_est, t_gt):
    assert np.size(t_est) == np.size(t_gt) == 3
    return float(_one(np.eye(3), t_est, np.eye(3), t_gt, np.zeros((1, 3)))[4])


def estimate_visib_mask(d_test, d_model, delta):
    assert d_test.shape == d_model.shape
    dev = torch.device("cuda", torch.cuda.current_device())
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev)
    out = visib_mask(t(d_test), t(d_model), delta)
    torch.cuda.synchronize(dev)
    return out.cpu().numpy() != 0


def estimate_visib_mask_gt(d_test, d_gt, delta):
    return estimate_visib_mask(d_test, d_gt, delta)


def estimate_visib_mask_est(d_test, d_est, visib_gt, delta):
    dev = torch.device("cuda", torch.cuda.current_device())
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev)
    out = visib_mask(t(d_test), t(d_est), delta, visib_gt=t(visib_gt))
    torch.cuda.synchronize(dev)
    return out.cpu().numpy() != 0


def calc_EPE_one_pair(flow_pred_list, flow_gt, flow_type):
    """Same arguments and result dict as tester.py:719-736 (flows are [H,W,2] numpy)."""
    dev = torch.device("cuda", torch.cuda.current_device())
    pred = np.asarray(flow_pred_list[flow_type])
    gt, visible, bg = flow_gt[flow_type][0], flow_gt[flow_type][1], flow_gt[flow_type][2]
    chw = lambda a: torch.from_numpy(np.ascontiguousarray(np.asarray(a, np.float32).transpose(2, 0, 1)[None])).to(dev)
    pl = lambda a: torch.from_numpy(np.ascontiguousarray(np.asarray(a, np.float32)[None, None])).to(dev)
    out = flow_epe(chw(pred), chw(gt), pl(visible), bg=pl(bg))
    torch.cuda.synchronize(dev)
    o = out[0].cpu().numpy()
    return {"epe_all": o[0], "num_all": int(o[1]), "epe_viz": o[2], "num_viz": o[3],
            "epe_vizbg": o[4], "num_vizbg": int(o[5])}
