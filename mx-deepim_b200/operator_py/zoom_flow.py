"""ZoomFlow -- zoom (or inverse-zoom) flow and binarised flow weights.

Mirror of /root/reference/deepim/operator_py/zoom_flow.py:21-118: flow is
bilinearly resampled and rescaled by 1/wx (zoom) or wx (inverse zoom, :59-66);
flow weights are resampled and binarised with round(v - 0.45) (:74-76, forward
zoom only).
"""
from __future__ import annotations

from .. import _lib
from .. import operator as mxop
from . import _common as C


class ZoomFlowOperator(mxop.CustomOp):
    def __init__(self, height, width, b_inv_zoom):
        super(ZoomFlowOperator, self).__init__()
        self.height = height
        self.width = width
        self.b_inv_zoom = b_inv_zoom

    def forward(self, is_train, req, in_data, out_data, aux):
        zoom_factor, flow = in_data[0], in_data[1]
        groups = [C.Group(flow, out_data[0], _lib.RC_PRE_NONE,
                          _lib.RC_POST_MUL_WX if self.b_inv_zoom else _lib.RC_POST_DIV_WX,
                          req[0])]
        if not self.b_inv_zoom:
            groups.append(C.Group(in_data[2], out_data[1], _lib.RC_PRE_NONE,
                                  _lib.RC_POST_ROUND_M045, req[1]))
        C.zoom_sample(zoom_factor, groups, self.height, self.width, inverse=self.b_inv_zoom)

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        for i in range(len(in_grad)):
            C.fill_zero(in_grad[i], req[i])


@mxop.register("ZoomFlow")
class ZoomFlowProp(mxop.CustomOpProp):
    def __init__(self, width=640, height=480, b_inv_zoom="False"):
        super(ZoomFlowProp, self).__init__(True)
        self.height = int(height)
        self.width = int(width)
        self.b_inv_zoom = b_inv_zoom.lower() == "true"

    def list_arguments(self):
        if self.b_inv_zoom:
            return ["zoom_factor", "flow"]
        return ["zoom_factor", "flow", "flow_weights"]

    def list_outputs(self):
        if self.b_inv_zoom:
            return ["zoom_flow"]
        return ["zoom_flow", "zoom_flow_weights"]

    def infer_shape(self, in_shape):
        out_shape = in_shape[1:]
        return in_shape, out_shape, []

    def infer_type(self, in_type):
        dtype = in_type[0]
        n_in = 2 if self.b_inv_zoom else 3
        return [dtype] * n_in, [dtype] * (n_in - 1), []

    def create_operator(self, ctx, shapes, dtypes):
        return ZoomFlowOperator(self.height, self.width, self.b_inv_zoom)
