"""``mx.operator``-shaped plugin shim over torch CUDA tensors.

The reference registers its zoom / Transform3D plugins as MXNet CustomOps:

    @mx.operator.register("ZoomMask")                     (zoom_mask.py:143)
    class ZoomMaskProp(mx.operator.CustomOpProp): ...
    class ZoomMaskOperator(mx.operator.CustomOp): ...

MXNet is not importable in this environment, so this module supplies the same
three names with the same contract -- ``CustomOp.assign(dst, req, src)`` with
``req`` in {"null", "write", "inplace", "add"}; ``CustomOpProp(need_top_grad)``
with ``list_arguments / list_outputs / infer_shape / infer_type /
create_operator``; ``register(name)``; kwargs crossing the boundary as strings
-- so that the operator classes in ``operator_py`` read like the reference's
and would register unchanged under the real ``mx.operator`` (swap the import).
``in_data`` / ``out_data`` are float32 CUDA ``torch.Tensor`` objects instead of
``mx.nd.NDArray``.

``Custom(...)`` is the eager counterpart of ``mx.nd.Custom``: it stringifies
kwargs the way MXNet does, builds the Prop, infers shapes, allocates outputs
and runs ``forward``; the returned handle can run ``backward``.
"""
from __future__ import annotations

from typing import Dict, List

import numpy as np

_REGISTRY: Dict[str, type] = {}


class CustomOp(object):
    """Base class for operators implemented in Python (mx.operator.CustomOp)."""

    def __init__(self):
        pass

    def forward(self, is_train, req, in_data, out_data, aux):
        raise NotImplementedError()

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        raise NotImplementedError()

    def assign(self, dst, req, src):
        """Helper for assigning into dst depending on requirements."""
        if req == "null":
            return
        import torch

        if req in ("write", "inplace"):
            if isinstance(src, torch.Tensor):
                if src.data_ptr() != dst.data_ptr():
                    dst.copy_(src.reshape(dst.shape))
            else:
                dst.fill_(float(src))
        elif req == "add":
            if isinstance(src, torch.Tensor):
                dst.add_(src.reshape(dst.shape))
            else:
                dst.add_(float(src))
        else:
            raise ValueError("unknown req %r" % (req,))


class CustomOpProp(object):
    """Base class for operator property classes (mx.operator.CustomOpProp)."""

    def __init__(self, need_top_grad=True):
        self.need_top_grad_ = need_top_grad

    def infer_shape(self, in_shape):
        return in_shape, (in_shape[0],) * len(self.list_outputs()), ()

    def infer_type(self, in_type):
        return in_type, [in_type[0]] * len(self.list_outputs()), [in_type[0]] * len(
            self.list_auxiliary_states())

    def list_outputs(self):
        return ["output"]

    def list_arguments(self):
        return ["data"]

    def list_auxiliary_states(self):
        return []

    def declare_backward_dependency(self, out_grad, in_data, out_data):
        deps = []
        if self.need_top_grad_:
            deps.extend(out_grad)
        deps.extend(in_data)
        deps.extend(out_data)
        return deps

    def create_operator(self, ctx, in_shapes, in_dtypes):
        return CustomOp()


def register(reg_name):
    """Register a subclass of CustomOpProp under ``reg_name`` (mx.operator.register)."""

    def do_register(prop_cls):
        _REGISTRY[reg_name] = prop_cls
        return prop_cls

    return do_register


def get_all_registered_operators() -> List[str]:
    return sorted(_REGISTRY)


def get_prop_class(op_type: str):
    if op_type not in _REGISTRY:
        raise KeyError("Cannot find custom operator type %s" % op_type)
    return _REGISTRY[op_type]


def parse_array(text, dtype=np.float32):
    """Parse the ``str(ndarray)`` form MXNet hands to a Prop (``"[a b c]"``).

    Stands in for ``np.fromstring(K[1:-1], dtype=np.float32, sep=" ")``
    (zoom_mask.py:147), whose text mode is deprecated in numpy 2."""
    if isinstance(text, str):
        body = text.replace("[", " ").replace("]", " ").replace(",", " ")
        return np.array(body.split(), dtype=np.float64).astype(dtype)
    return np.asarray(text, dtype=dtype).reshape(-1)


def _stringify(value) -> str:
    # MXNet serialises every Custom kwarg with str() before the Prop sees it.
    if isinstance(value, np.ndarray):
        with np.printoptions(precision=17, floatmode="maxprec", linewidth=1 << 20,
                             threshold=1 << 20):
            # numpy's default repr prints float32 shortest-round-trip digits; keep that
            return str(value)
    return str(value)


class CustomHandle(object):
    """Result of an eager ``Custom`` call; keeps what ``backward`` needs."""

    def __init__(self, prop, op, in_data, out_data, arg_names, out_names):
        self.prop = prop
        self.op = op
        self.in_data = in_data
        self.out_data = out_data
        self.arg_names = arg_names
        self.out_names = out_names

    @property
    def outputs(self):
        return self.out_data

    def backward(self, out_grad, req="write", in_grad=None):
        import torch

        if in_grad is None:
            in_grad = [torch.empty_like(t) for t in self.in_data]
        reqs = [req] * len(self.in_data) if isinstance(req, str) else list(req)
        self.op.backward(reqs, list(out_grad), self.in_data, self.out_data, in_grad, [])
        return in_grad


def Custom(*args, **kwargs):
    """Eager stand-in for ``mx.nd.Custom(..., op_type=...)``.

    Tensor kwargs / positional args are the operator inputs (matched by
    ``list_arguments``); every other kwarg is stringified and passed to the
    Prop constructor.  Returns a :class:`CustomHandle`.
    """
    import torch

    op_type = kwargs.pop("op_type")
    kwargs.pop("name", None)
    out = kwargs.pop("out", None)
    req = kwargs.pop("req", "write")
    is_train = bool(kwargs.pop("is_train", True))
    tensors = {k: v for k, v in kwargs.items() if isinstance(v, torch.Tensor)}
    params = {k: _stringify(v) for k, v in kwargs.items() if not isinstance(v, torch.Tensor)}
    prop = get_prop_class(op_type)(**params)
    arg_names = prop.list_arguments()
    in_data = list(args)
    for name in arg_names[len(in_data):]:
        if name not in tensors:
            raise ValueError("Custom(%s): missing input %r" % (op_type, name))
        in_data.append(tensors[name])
    if len(in_data) != len(arg_names):
        raise ValueError("Custom(%s): expected %d inputs, got %d"
                         % (op_type, len(arg_names), len(in_data)))
    in_shapes = [list(t.shape) for t in in_data]
    _, out_shapes, _ = prop.infer_shape(in_shapes)
    out_names = prop.list_outputs()
    device = in_data[0].device
    if out is None:
        out_data = [torch.empty(tuple(s), dtype=torch.float32, device=device) for s in out_shapes]
    else:
        out_data = list(out)
    op = prop.create_operator(device, in_shapes, [np.float32] * len(in_data))
    reqs = [req] * len(out_data) if isinstance(req, str) else list(req)
    op.forward(is_train, reqs, in_data, out_data, [])
    return CustomHandle(prop, op, in_data, out_data, arg_names, out_names)
