"""ZoomImage -- crop box / zoom factor from image non-zero pixels, zoom two images.

Mirror of /root/reference/deepim/operator_py/zoom_image.py:18-161.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from .. import operator as mxop
from . import _common as C


class ZoomImageOperator(mxop.CustomOp):
    def __init__(self, K, height, width, pixel_means):
        super(ZoomImageOperator, self).__init__()
        self.K = K
        self.height = height
        self.width = width
        self.pixel_means = pixel_means  # [1,3,1,1], already reversed by the Prop

    def forward(self, is_train, req, in_data, out_data, aux):
        image_observed, image_rendered, src_pose = in_data[:3]
        B = image_observed.shape[0]
        means = np.asarray(self.pixel_means, dtype=np.float32).reshape(-1)
        zf = out_data[2] if req[2] in ("write", "inplace") else torch.empty(
            (B, 4), dtype=torch.float32, device=image_observed.device)
        C.zoom_factor(image_observed, image_rendered, src_pose, self.K, means, zf,
                      _lib.RC_ZF_IMAGE)
        C.zoom_sample(zf, [
            C.Group(image_observed, out_data[0], _lib.RC_PRE_ADD_MEAN, _lib.RC_POST_SUB_MEAN,
                    req[0], means),
            C.Group(image_rendered, out_data[1], _lib.RC_PRE_ADD_MEAN, _lib.RC_POST_SUB_MEAN,
                    req[1], means),
        ], self.height, self.width)
        if zf is not out_data[2]:
            self.assign(out_data[2], req[2], zf)

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        for i in range(3):
            C.fill_zero(in_grad[i], req[i])


@mxop.register("ZoomImage")
class ZoomImageProp(mxop.CustomOpProp):
    def __init__(self, K, width=640, height=480, pixel_means="[0 0 0]"):
        super(ZoomImageProp, self).__init__(True)
        self.K = mxop.parse_array(K, np.float32).reshape([3, 3])
        self.height = int(height)
        self.width = int(width)
        self.pixel_means = mxop.parse_array(pixel_means, np.float32).reshape(3)
        self.pixel_means = self.pixel_means[::-1]
        self.pixel_means = self.pixel_means[np.newaxis, :, np.newaxis, np.newaxis]

    def list_arguments(self):
        return ["image_observed", "image_rendered", "src_pose"]

    def list_outputs(self):
        return ["zoom_image_observed", "zoom_image_rendered", "zoom_factor"]

    def infer_shape(self, in_shape):
        batch_size = in_shape[0][0]
        out_shape = in_shape[:]
        out_shape[2] = [batch_size, 4]
        return in_shape, out_shape, []

    def infer_type(self, in_type):
        dtype = in_type[0]
        return [dtype] * 3, [dtype] * 3, []

    def create_operator(self, ctx, shapes, dtypes):
        return ZoomImageOperator(self.K, self.height, self.width, self.pixel_means)
