"""ZoomImageWithFactor -- zoom two images with a given zoom factor.

Mirror of /root/reference/deepim/operator_py/zoom_image_with_factor.py:21-120.
Note the reference's output order: out[0] = observed, out[1] = rendered, and
the optional red centre spot is applied to the RENDERED image only (:67-70).
"""
from __future__ import annotations

import numpy as np

from .. import _lib
from .. import operator as mxop
from . import _common as C


class ZoomImageWithFactorOperator(mxop.CustomOp):
    def __init__(self, height, width, pixel_means, high_light_center):
        super(ZoomImageWithFactorOperator, self).__init__()
        self.height = height
        self.width = width
        self.pixel_means = pixel_means
        self.high_light_center = high_light_center
        self.spot_radius = 5

    def forward(self, is_train, req, in_data, out_data, aux):
        zoom_factor, image_observed, image_rendered = in_data[:3]
        means = np.asarray(self.pixel_means, dtype=np.float32).reshape(-1)
        post_rendered = (_lib.RC_POST_SPOT_SUB_MEAN if self.high_light_center
                         else _lib.RC_POST_SUB_MEAN)
        C.zoom_sample(zoom_factor, [
            C.Group(image_observed, out_data[0], _lib.RC_PRE_ADD_MEAN, _lib.RC_POST_SUB_MEAN,
                    req[0], means),
            C.Group(image_rendered, out_data[1], _lib.RC_PRE_ADD_MEAN, post_rendered, req[1],
                    means),
        ], self.height, self.width)

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        for i in range(3):
            C.fill_zero(in_grad[i], req[i])


@mxop.register("ZoomImageWithFactor")
class ZoomImageWithFactorProp(mxop.CustomOpProp):
    def __init__(self, width=640, height=480, pixel_means="[0 0 0]", high_light_center="False"):
        super(ZoomImageWithFactorProp, self).__init__(True)
        self.height = int(height)
        self.width = int(width)
        self.pixel_means = mxop.parse_array(pixel_means, np.float32).reshape(3)
        self.pixel_means = self.pixel_means[::-1]
        self.pixel_means = self.pixel_means[np.newaxis, :, np.newaxis, np.newaxis]
        self.hight_light_center = high_light_center.lower() == "true"

    def list_arguments(self):
        return ["zoom_factor", "image_observed", "image_rendered"]

    def list_outputs(self):
        return ["zoom_image_observed", "zoom_image_rendered"]

    def infer_shape(self, in_shape):
        out_shape = [in_shape[1], in_shape[2]]
        return in_shape, out_shape, []

    def infer_type(self, in_type):
        dtype = in_type[0]
        return [dtype] * 3, [dtype] * 2, []

    def create_operator(self, ctx, shapes, dtypes):
        return ZoomImageWithFactorOperator(self.height, self.width, self.pixel_means,
                                           self.hight_light_center)
