"""ZoomMaskWithFactor -- threshold (> 0.2), zoom or inverse-zoom, round one mask.

Mirror of /root/reference/deepim/operator_py/zoom_mask_with_factor.py:22-102.
"""
from __future__ import annotations

from .. import _lib
from .. import operator as mxop
from . import _common as C


class ZoomMaskWithFactorOperator(mxop.CustomOp):
    def __init__(self, height, width, b_inv_zoom):
        super(ZoomMaskWithFactorOperator, self).__init__()
        self.height = height
        self.width = width
        self.b_inv_zoom = b_inv_zoom

    def forward(self, is_train, req, in_data, out_data, aux):
        zoom_factor, mask = in_data[0], in_data[1]
        C.zoom_sample(zoom_factor, [
            C.Group(mask, out_data[0], _lib.RC_PRE_THRESH02, _lib.RC_POST_ROUND, req[0]),
        ], self.height, self.width, inverse=self.b_inv_zoom)

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        for i in range(2):
            C.fill_zero(in_grad[i], req[i])


@mxop.register("ZoomMaskWithFactor")
class ZoomMaskWithFactorProp(mxop.CustomOpProp):
    def __init__(self, width=640, height=480, b_inv_zoom="False"):
        super(ZoomMaskWithFactorProp, self).__init__(True)
        self.height = int(height)
        self.width = int(width)
        self.b_inv_zoom = b_inv_zoom.lower() == "true"

    def list_arguments(self):
        return ["zoom_factor", "mask"]

    def list_outputs(self):
        return ["zoom_mask"]

    def infer_shape(self, in_shape):
        out_shape = [in_shape[1]]
        return in_shape, out_shape, []

    def infer_type(self, in_type):
        dtype = in_type[0]
        return [dtype, dtype], [dtype], []

    def create_operator(self, ctx, shapes, dtypes):
        return ZoomMaskWithFactorOperator(self.height, self.width, self.b_inv_zoom)
