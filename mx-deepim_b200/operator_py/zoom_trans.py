"""ZoomTrans -- scale (tx, ty) of the translation delta by the zoom factor.

Mirror of /root/reference/deepim/operator_py/zoom_trans.py:16-105.  The only
zoom plugin with a real gradient (:50-76).
"""
from __future__ import annotations

import torch

from .. import _lib
from .. import operator as mxop
from . import _common as C


def _launch(zoom_factor, src, dst, inverse, scale_xy, req):
    skip, acc = C._req_flag(req)
    if skip:
        return
    C.check_tensor("zoom_factor", zoom_factor, 2)
    C.check_tensor("trans", src, 2)
    C.check_tensor("trans_out", dst, 2)
    lib = _lib.load()
    with torch.cuda.device(src.device):
        st = lib.rc_zoom_trans_dev(zoom_factor.data_ptr(), src.data_ptr(), dst.data_ptr(),
                                   src.shape[0], 1 if inverse else 0, 1 if scale_xy else 0, acc,
                                   _lib.stream_ptr())
    _lib.check(st, "rc_zoom_trans_dev")


class ZoomTransOperator(mxop.CustomOp):
    def __init__(self, b_inv_zoom, b_zoom_grad):
        super(ZoomTransOperator, self).__init__()
        self.b_inv_zoom = b_inv_zoom
        self.b_zoom_grad = b_zoom_grad

    def forward(self, is_train, req, in_data, out_data, aux):
        _launch(in_data[0], in_data[1], out_data[0], self.b_inv_zoom, True, req[0])

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        C.fill_zero(in_grad[0], req[0])
        _launch(in_data[0], out_grad[0], in_grad[1], self.b_inv_zoom, self.b_zoom_grad, req[1])


@mxop.register("ZoomTrans")
class ZoomTransProp(mxop.CustomOpProp):
    def __init__(self, b_inv_zoom="False", b_zoom_grad="False"):
        super(ZoomTransProp, self).__init__(True)
        self.b_inv_zoom = b_inv_zoom.lower() == "true"
        self.b_zoom_grad = b_zoom_grad.lower() == "true"

    def list_arguments(self):
        return ["zoom_factor", "trans_delta"]

    def list_outputs(self):
        return ["zoom_trans_delta"]

    def infer_shape(self, in_shape):
        out_shape = [in_shape[1]]
        return in_shape, out_shape, []

    def infer_type(self, in_type):
        dtype = in_type[0]
        return [dtype, dtype], [dtype], []

    def create_operator(self, ctx, shapes, dtypes):
        return ZoomTransOperator(self.b_inv_zoom, self.b_zoom_grad)
