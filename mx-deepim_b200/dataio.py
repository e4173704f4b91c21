"""On-disk formats either side of the path -> packed host batches for
`pipeline.HostFeed(packed=True)` and device decode through prep.py.

The reference's loaders decode every frame to float on ONE host thread before the executor
uploads it (lib/utils/image.py:65-310): `cv2.imread` -> `.astype(np.float32)` -> mask multiply
-> `/ DEPTH_FACTOR` -> `image.transform`, then `nd.array` copies 4 bytes per value over PCIe.
Here the files are read into their STORAGE dtypes (uint8 BGR, 16-bit depth, uint8 label masks),
stacked into pinned batches, uploaded as they are (3.06 instead of 12.4 MB per pair) and decoded
by the prep kernels on the device.

    <frame>-color.png   8-bit BGR       image.py:96,101  (cv2.IMREAD_COLOR)
    <frame>-depth.png   16-bit, 1/DEPTH_FACTOR m   image.py:200-205,228-262  (IMREAD_UNCHANGED)
    <frame>-label.png   8-bit instance ids         image.py:235-244,300-304
    <frame>-pose.txt    one header row + 3x4       np.loadtxt(path, skiprows=1), e.g.
                                                   lib/flow_c/flow.py:58, render_py_multi.py:253
    <model>/points.xyz  M x 3 text                 lib/pair_matching/load_object_points.py:11-15
"""
from __future__ import annotations

import os

import numpy as np

PAIR_IMAGE_KEYS = ("image_observed", "image_rendered")
PAIR_DEPTH_KEYS = ("depth_observed", "depth_rendered", "depth_gt_observed")
PAIR_MASK_KEYS = ("mask_gt_observed",)


def _exists(path):
    assert os.path.exists(path), "{} does not exist".format(path)  # image.py:197-199
    return path


def load_object_points(point_path):
    """load_object_points.py:11-15: `points.xyz`, M rows of x y z (metres)."""
    assert os.path.exists(point_path), "Path does not exist: {}".format(point_path)
    return np.loadtxt(point_path)


def load_pose(pose_path):
    """`-pose.txt`: one header row (the class index), then the 3x4 [R|t] rows."""
    return np.loadtxt(_exists(pose_path), skiprows=1)


def read_color(path):
    """uint8 [H,W,3] BGR, as cv2.imread(path, cv2.IMREAD_COLOR) (image.py:96)."""
    import cv2

    im = cv2.imread(_exists(path), cv2.IMREAD_COLOR)
    assert im is not None and im.dtype == np.uint8, path
    return im


def read_depth(path):
    """uint16 [H,W] raw depth, as cv2.imread(path, cv2.IMREAD_UNCHANGED) before the reference's
    .astype(np.float32) (image.py:200-202)."""
    import cv2

    d = cv2.imread(_exists(path), cv2.IMREAD_UNCHANGED)
    assert d is not None and d.ndim == 2, path
    if d.dtype != np.uint16:
        assert d.dtype == np.uint8, "depth PNG must be 8- or 16-bit: {}".format(path)
        d = d.astype(np.uint16)
    return d


def read_label(path):
    """uint8 [H,W] instance ids (image.py:235-237)."""
    import cv2

    m = cv2.imread(_exists(path), cv2.IMREAD_UNCHANGED)
    assert m is not None, path
    if m.ndim == 3:
        m = m[:, :, 0]
    return np.ascontiguousarray(m.astype(np.uint8))


def read_pair(pair_rec):
    """One pairdb record (the keys lib/dataset/*.py put into `pair_rec`) -> storage-dtype arrays.
    Optional keys are skipped when absent.  `mask_idx` defaults to 1."""
    out = {"mask_idx": np.int32(pair_rec.get("mask_idx", 1))}
    for k in PAIR_IMAGE_KEYS:
        if k in pair_rec:
            out[k] = read_color(pair_rec[k])
    for k in PAIR_DEPTH_KEYS:
        if k in pair_rec:
            out[k] = read_depth(pair_rec[k])
    for k in PAIR_MASK_KEYS:
        if k in pair_rec:
            out[k] = read_label(pair_rec[k])
    for k in ("pose_observed", "pose_rendered"):
        if k in pair_rec:
            v = pair_rec[k]
            out[k] = np.asarray(load_pose(v) if isinstance(v, str) else v, np.float32)
    return out


def stack_pairs(records, pin=False):
    """List of read_pair() dicts -> one batch dict of torch CPU tensors in storage dtypes:
    images uint8 [B,H,W,3], depths int16 VIEWS of the uint16 data [B,H,W] (torch has no uint16
    arithmetic; prep.decode_depth reinterprets), labels uint8 [B,H,W], mask_idx int32 [B],
    poses float32 [B,3,4].  pin=True allocates page-locked memory for async uploads."""
    import torch

    batch = {}
    for k in records[0]:
        arr = np.stack([np.asarray(r[k]) for r in records])
        if arr.dtype == np.uint16:
            arr = arr.view(np.int16)
        t = torch.from_numpy(np.ascontiguousarray(arr))
        batch[k] = t.pin_memory() if pin else t
    return batch


def decode_batch(batch, device, depth_factor=1000.0, pixel_means_bgr=(0.0, 0.0, 0.0),
                 mask_inputs=True, stream=None):
    """Upload a stack_pairs() batch and decode it on the device into the float32 blobs of the
    reference's loaders:

        image_*            [B,3,H,W]  image.transform (image.py:709-720)
        depth_gt_observed  [B,1,H,W]  raw / DEPTH_FACTOR (image.py:200-205)
        depth_observed     [B,1,H,W]  raw * (label == mask_idx) / DEPTH_FACTOR when
                                      `mask_inputs` (config.network.MASK_INPUTS, image.py:228-262)
        depth_rendered     [B,1,H,W]  raw / DEPTH_FACTOR
        mask_gt_observed   [B,1,H,W]  label == mask_idx (image.py:300-304)
    """
    import torch

    from . import prep

    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("dataio.decode_batch needs a CUDA device (no CPU fallback)")
    up = {k: v.to(dev, non_blocking=True) for k, v in batch.items()}
    out = {}
    for k in PAIR_IMAGE_KEYS:
        if k in up:
            out[k] = prep.transform_batch(up[k], np.asarray(pixel_means_bgr, np.float64), stream=stream)
    label, idx = up.get("mask_gt_observed"), up.get("mask_idx")
    for k in PAIR_DEPTH_KEYS:
        if k in up:
            masked = k == "depth_observed" and mask_inputs and label is not None
            out[k] = prep.decode_depth(up[k], depth_factor, label if masked else None,
                                       idx if masked else None, stream=stream)
    if label is not None:
        out["mask_gt_observed"] = prep.decode_mask(label, idx, stream=stream)
    for k in ("pose_observed", "pose_rendered"):
        if k in up:
            out[k] = up[k]
    return out
