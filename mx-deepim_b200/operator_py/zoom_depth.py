"""ZoomDepth -- zoom two depth maps with a given zoom factor.

Mirror of /root/reference/deepim/operator_py/zoom_depth.py:18-84 (the
per-sample debug print at :32 is not reproduced).
"""
from __future__ import annotations

from .. import _lib
from .. import operator as mxop
from . import _common as C


class ZoomDepthOperator(mxop.CustomOp):
    def __init__(self, height, width):
        super(ZoomDepthOperator, self).__init__()
        self.height = height
        self.width = width

    def forward(self, is_train, req, in_data, out_data, aux):
        zoom_factor, depth_observed, depth_rendered = in_data[:3]
        C.zoom_sample(zoom_factor, [
            C.Group(depth_observed, out_data[0], req=req[0]),
            C.Group(depth_rendered, out_data[1], req=req[1]),
        ], self.height, self.width)

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        for i in range(3):
            C.fill_zero(in_grad[i], req[i])


@mxop.register("ZoomDepth")
class ZoomDepthProp(mxop.CustomOpProp):
    def __init__(self, width=640, height=480):
        super(ZoomDepthProp, self).__init__(True)
        self.height = int(height)
        self.width = int(width)

    def list_arguments(self):
        return ["zoom_factor", "depth_observed", "depth_rendered"]

    def list_outputs(self):
        return ["zoom_depth_observed", "zoom_depth_rendered"]

    def infer_shape(self, in_shape):
        out_shape = in_shape[1:]
        return in_shape, out_shape, []

    def infer_type(self, in_type):
        dtype = in_type[0]
        return [dtype] * 3, [dtype] * 2, []

    def create_operator(self, ctx, shapes, dtypes):
        return ZoomDepthOperator(self.height, self.width)
