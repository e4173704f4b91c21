"""oracle/prep_np.py -- TEST INFRASTRUCTURE ONLY (never on the product path).

numpy restatement of the iteration-0 mask preparation of the reference's loader:
  /root/reference/lib/utils/mask_dilate.py:10-58          mask_dilate
  /root/reference/lib/utils/get_min_rect.py:9-24          get_min_rect
  /root/reference/lib/utils/image.py:272-492              get_pair_mask, for the policies the
      shipped configs use (experiments/deepim/cfgs/*.yaml): TRAIN.INIT_MASK = "box_gt"
      (+ MASK_DILATE), TEST.INIT_MASK = "box_rendered", and mask_rendered of both phases.

mask_dilate is restated per pixel with the random draws taken out (`draw_dilate`), which is how
the device kernel gets them.  PINNED: tests/golden/make_golden_masks.py runs the reference's
own get_pair_mask / mask_dilate (imported unmodified, seeded np.random) and
tests/test_golden_ops_cpu.py compares bit for bit.
"""
from __future__ import annotations

import numpy as np


def draw_dilate(max_thickness=10, rng=np.random):
    """The random draws of one mask_dilate call, in the reference's order (mask_dilate.py:19-47):
    returns the four block thicknesses (down, up, right, left), 0 where `direction` skips it."""
    direction = rng.randint(10)
    t = [0, 0, 0, 0]
    for i, skip in enumerate(([0, 1, 4], [1, 2, 5], [2, 3, 6], [0, 3, 7])):
        if direction not in skip:
            t[i] = rng.randint(max_thickness) + 1
    return t


def mask_dilate_from_draws(mask_origin, t):
    """mask_dilate.py:21-57 for given draws, per pixel: a zero pixel becomes 1 when the pixel
    `t` rows above / below / columns left / right of it (per block) is non-zero; a non-zero
    pixel keeps its value; everything is clamped at 1."""
    m = np.asarray(mask_origin)
    H, W = m.shape
    td, tu, tr, tl = [int(v) for v in t]
    grow = np.zeros((H, W), bool)
    if td > 0:
        grow[td:, :] |= m[:-td, :] != 0
    if tu > 0:
        grow[:-tu, :] |= m[tu:, :] != 0
    if tr > 0:
        grow[:, tr:] |= m[:, :-tr] != 0
    if tl > 0:
        grow[:, :-tl] |= m[:, tl:] != 0
    out = np.where(m != 0, np.minimum(m, 1.0), np.where(grow, 1.0, 0.0))
    return out.astype(m.dtype if m.dtype.kind == "f" else np.float64)


def get_min_rect(mask):
    """get_min_rect.py:9-24."""
    nz_x = np.nonzero(np.max(mask, 0))[0]
    nz_y = np.nonzero(np.max(mask, 1))[0]
    return np.min(nz_x), np.min(nz_y), np.max(nz_x), np.max(nz_y)


def mask_rendered_from_depth(depth):
    """image.py:487-492: depth (metres) with everything above 0.2 set to 1."""
    d = np.array(depth, dtype=np.float32, copy=True)
    d[d > 0.2] = 1
    return d


def pair_masks_train_box_gt(label, mask_idx, depth_rendered, dilate_draws=None):
    """image.py:289-349 with INIT_MASK = "box_gt" for one pair.  Returns (mask_observed,
    mask_gt_observed, mask_rendered), each [H,W]."""
    gt = np.zeros(label.shape)
    gt[label.astype(np.float32) == mask_idx] = 1.0
    obs = np.zeros(label.shape)
    x0, y0, x1, y1 = get_min_rect(gt)
    obs[y0:y1, x0:x1] = 1.0
    if dilate_draws is not None:
        obs = mask_dilate_from_draws(obs, dilate_draws)
    return obs, gt, mask_rendered_from_depth(depth_rendered)


def pair_masks_test_box_rendered(depth_rendered, dilate_draws=None):
    """image.py:351-481 with TEST.INIT_MASK = "box_rendered" for one pair."""
    d = np.asarray(depth_rendered, np.float32)
    rend = np.zeros(d.shape)
    rend[d > 0.2] = 1.0
    obs = np.zeros(d.shape)
    if np.sum(d) != 0 and np.count_nonzero(rend) != 0:
        x0, y0, x1, y1 = get_min_rect(rend)
        obs[y0:y1, x0:x1] = 1.0
    if dilate_draws is not None:
        obs = mask_dilate_from_draws(obs, dilate_draws)
    return obs, obs, mask_rendered_from_depth(d)
