"""rc_zoom_factor_noalloc_dev -- the crop-box / zoom-factor kernel with its scratch in a static
device array inside the library (no cudaMalloc on the call path; include/rendercompare.h).  It has
no counterpart in the reference, so the checker is
the existing entry point rc_zoom_factor_dev on the same inputs: crop boxes and zoom factors are
integer / bit-exact quantities everywhere in this suite, and they are here.

Collected last on purpose."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call_pair(lib, _lib, torch, dev, host, K, mode, means=None):
    """Both entry points through the C ABI on the same device inputs -> (zf, bbox) x 2."""
    if mode == _lib.RC_ZF_MASK:
        obs = torch.from_numpy(host["mask_gt_observed"]).to(dev)
        rend = torch.from_numpy(host["depth_rendered"]).to(dev)
    else:
        obs = torch.from_numpy(host["image_observed"]).to(dev)
        rend = torch.from_numpy(host["image_rendered"]).to(dev)
    pose = torch.from_numpy(host["src_pose"]).to(dev)
    Kd = torch.from_numpy(np.asarray(K, np.float32)).to(dev)
    pm = torch.from_numpy(np.asarray(means, np.float32)).to(dev) if means is not None else None
    B, C, H, W = obs.shape
    out = []
    stream = _lib.stream_ptr(None)
    for noalloc in (False, True):
        zf = torch.full((B, 4), -7.0, dtype=torch.float32, device=dev)
        bbox = torch.full((B, 8), -7, dtype=torch.int32, device=dev)
        pmp = pm.data_ptr() if pm is not None else None
        if noalloc:
            st = lib.rc_zoom_factor_noalloc_dev(obs.data_ptr(), rend.data_ptr(), pose.data_ptr(),
                                                Kd.data_ptr(), pmp, zf.data_ptr(), bbox.data_ptr(),
                                                B, C, H, W, mode, stream)
        else:
            ws = torch.zeros((B, _lib.RC_ZF_WS_INTS), dtype=torch.int32, device=dev)
            st = lib.rc_zoom_factor_dev(obs.data_ptr(), rend.data_ptr(), pose.data_ptr(),
                                        Kd.data_ptr(), pmp, zf.data_ptr(), bbox.data_ptr(),
                                        ws.data_ptr(), B, C, H, W, mode, stream)
        _lib.check(st, "rc_zoom_factor%s_dev" % ("_noalloc" if noalloc else ""))
        torch.cuda.synchronize()
        out.append((zf.cpu().numpy(), bbox.cpu().numpy()))
    return out


def test_zoom_factor_noalloc_matches_allocating_entry_point(cuda_device):
    import torch

    from mx_deepim_b200 import _lib, synth

    lib = _lib.load()
    assert lib.rc_zoom_factor_noalloc_max_batch() == 4096
    # two calls in a row on the shared static scratch: a larger batch first, then a smaller one
    # in the other mode (rows the first call used must not leak into the second)
    for B, seed, mode in ((9, 301, _lib.RC_ZF_MASK), (4, 302, _lib.RC_ZF_IMAGE),
                          (9, 303, _lib.RC_ZF_MASK)):
        host = synth.make_compare_batch(B, seed=seed, occlusion=(seed % 2 == 0))
        means = synth.PIXEL_MEANS[::-1].copy() if mode == _lib.RC_ZF_IMAGE else None
        (zf0, bb0), (zf1, bb1) = _call_pair(lib, _lib, torch, cuda_device, host, synth.K_LINEMOD,
                                            mode, means)
        assert np.isfinite(zf0).all() and (zf0 != -7.0).all()
        assert np.array_equal(bb0, bb1), (B, mode)
        assert np.array_equal(zf0.view(np.int32), zf1.view(np.int32)), (B, mode)  # bit-exact


def test_zoom_factor_noalloc_rejects_oversized_batch(cuda_device):
    from mx_deepim_b200 import _lib

    lib = _lib.load()
    too_many = lib.rc_zoom_factor_noalloc_max_batch() + 1
    st = lib.rc_zoom_factor_noalloc_dev(None, None, None, None, None, None, None, too_many, 1, 480,
                                        640, _lib.RC_ZF_MASK, None)
    assert st == _lib.RC_ERR_CAPACITY
    assert b"capacity" in lib.rc_status_string(st).lower()
    with pytest.raises(_lib.RenderCompareError):
        _lib.check(st, "rc_zoom_factor_noalloc_dev")
    assert lib.rc_zoom_factor_noalloc_dev(None, None, None, None, None, None, None, 0, 1, 480, 640,
                                          _lib.RC_ZF_MASK, None) == 0  # empty batch: nothing to do
