"""ZoomMask -- crop box / zoom factor from the masks, then zoom three masks.

Mirror of /root/reference/deepim/operator_py/zoom_mask.py:22-177 (operator
:22-140, prop :143-177).  Same argument/outputs names, kwargs and `req`
handling; the numpy bbox block (:51-105), the per-sample GridGenerator loop
(:107-119) and the three BilinearSampler + round calls (:121-129) are two
kernel launches here (rc_zoom_factor_dev, rc_zoom_sample_dev).
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from .. import operator as mxop
from . import _common as C


class ZoomMaskOperator(mxop.CustomOp):
    def __init__(self, K, height, width):
        super(ZoomMaskOperator, self).__init__()
        self.K = K
        self.height = height
        self.width = width

    def forward(self, is_train, req, in_data, out_data, aux):
        mask_observed, mask_gt_observed, mask_rendered, src_pose = in_data[:4]
        B = mask_observed.shape[0]
        zf = out_data[3] if req[3] in ("write", "inplace") else torch.empty(
            (B, 4), dtype=torch.float32, device=mask_observed.device)
        C.zoom_factor(mask_gt_observed, mask_rendered, src_pose, self.K, None, zf, _lib.RC_ZF_MASK)
        C.zoom_sample(zf, [
            C.Group(mask_observed, out_data[0], _lib.RC_PRE_NONE, _lib.RC_POST_ROUND, req[0]),
            C.Group(mask_gt_observed, out_data[1], _lib.RC_PRE_NONE, _lib.RC_POST_ROUND, req[1]),
            C.Group(mask_rendered, out_data[2], _lib.RC_PRE_THRESH02, _lib.RC_POST_ROUND, req[2]),
        ], self.height, self.width)
        if zf is not out_data[3]:
            self.assign(out_data[3], req[3], zf)

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        for i in range(4):
            C.fill_zero(in_grad[i], req[i])


@mxop.register("ZoomMask")
class ZoomMaskProp(mxop.CustomOpProp):
    def __init__(self, K, width=640, height=480):
        super(ZoomMaskProp, self).__init__(True)
        self.K = mxop.parse_array(K, np.float32).reshape([3, 3])
        self.height = int(height)
        self.width = int(width)

    def list_arguments(self):
        return ["mask_observed", "mask_gt_observed", "mask_rendered", "src_pose"]

    def list_outputs(self):
        return ["zoom_mask_observed", "zoom_mask_gt_observed", "zoom_mask_rendered", "zoom_factor"]

    def infer_shape(self, in_shape):
        batch_size = in_shape[0][0]
        out_shape = in_shape[:-1]
        out_shape.append([batch_size, 4])
        return in_shape, out_shape, []

    def infer_type(self, in_type):
        dtype = in_type[0]
        return [dtype] * 4, [dtype] * 4, []

    def create_operator(self, ctx, shapes, dtypes):
        return ZoomMaskOperator(self.K, self.height, self.width)
