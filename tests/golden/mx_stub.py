"""A numpy-backed stand-in for the handful of `mxnet` names the reference's operator plugins
touch, so that /root/reference/deepim/operator_py/*.py can be IMPORTED AND RUN UNMODIFIED in
the build container (MXNet itself is an un-vendored pip dependency that cannot be installed
here).  Used only by tests/golden/make_golden_ops.py to produce golden vectors; never imported
by the product or by the tests.

What this pins and what it does not:
  * everything the plugins do in Python/numpy -- the crop-box / zoom-factor arithmetic, the
    mask thresholds, the mean add / subtract, the centre spot, the inverse-zoom affine, the
    flow rescale, the 0.45 weight threshold, ZoomTrans, the Transform3D forward / backward
    scalar tails, `req` handling and the argument / output lists -- is the REFERENCE'S OWN
    CODE running;
  * `GridGenerator` and `BilinearSampler` are MXNet built-ins (C++/CUDA, not in the reference
    tree); here they are served by oracle/zoom_np.py's restatement, so the sampler itself
    stays "parity unpinned" (DESIGN.md section 4);
  * elementwise NDArray arithmetic, batch_dot, sum, exp, round are float32 numpy with MXNet's
    semantics (scalar operands rounded to float32, round half away from zero).
"""
from __future__ import annotations

import sys
import types

import numpy as np

f32 = np.float32


class Context(object):
    def __init__(self, kind="cpu", idx=0):
        self.device_type, self.device_id = kind, idx

    def __repr__(self):
        return "%s(%d)" % (self.device_type, self.device_id)


def _raw(x):
    if isinstance(x, NDArray):
        return x._a
    if isinstance(x, (int, float, np.integer, np.floating)):
        return f32(x)  # MXNet scalar ops: the scalar is cast to the array's dtype
    return np.asarray(x, dtype=f32)


class NDArray(object):
    """float32 array with the slice of mx.nd.NDArray the plugins use."""

    def __init__(self, a, ctx=None):
        self._a = a if (isinstance(a, np.ndarray) and a.dtype == f32) else np.array(a, dtype=f32)
        self.context = ctx or Context()

    # -- introspection
    @property
    def shape(self):
        return tuple(self._a.shape)

    @property
    def dtype(self):
        return f32

    def asnumpy(self):
        return self._a.copy()

    def asscalar(self):
        return self._a.reshape(-1)[0]

    def reshape(self, shape, *more):
        if more:
            shape = (shape,) + tuple(more)
        return NDArray(self._a.reshape(shape), self.context)

    def copy(self):
        return NDArray(self._a.copy(), self.context)

    # -- indexing: integer index on a 1-D array yields shape (1,), as in MXNet
    def __getitem__(self, key):
        if self._a.ndim == 1 and isinstance(key, (int, np.integer)):
            return NDArray(self._a[key:key + 1], self.context)
        return NDArray(self._a[key], self.context)

    def __setitem__(self, key, value):
        v = _raw(value)
        if self._a.ndim == 1 and isinstance(key, (int, np.integer)):
            self._a[key] = np.asarray(v, dtype=f32).reshape(-1)[0]
        else:
            self._a[key] = v

    # -- arithmetic (float32)
    def _bin(self, other, fn, swap=False):
        a, b = self._a, _raw(other)
        if swap:
            a, b = b, a
        return NDArray(np.asarray(fn(a, b), dtype=f32), self.context)

    def __add__(self, o):
        return self._bin(o, np.add)

    __radd__ = __add__

    def __sub__(self, o):
        return self._bin(o, np.subtract)

    def __rsub__(self, o):
        return self._bin(o, np.subtract, swap=True)

    def __mul__(self, o):
        return self._bin(o, np.multiply)

    __rmul__ = __mul__

    def __truediv__(self, o):
        return self._bin(o, np.divide)

    def __rtruediv__(self, o):
        return self._bin(o, np.divide, swap=True)

    __div__ = __truediv__

    def __neg__(self):
        return NDArray(-self._a, self.context)

    def _ibin(self, other, fn):
        self._a[...] = fn(self._a, _raw(other))
        return self

    def __iadd__(self, o):
        return self._ibin(o, np.add)

    def __isub__(self, o):
        return self._ibin(o, np.subtract)

    def __imul__(self, o):
        return self._ibin(o, np.multiply)

    def __itruediv__(self, o):
        return self._ibin(o, np.divide)

    __idiv__ = __itruediv__


# ------------------------------------------------------------------ mx.nd
def _array(source_array, ctx=None, dtype=None):
    return NDArray(np.array(_raw(source_array), dtype=f32, copy=True), ctx)


def _zeros(shape, ctx=None, dtype=None, **kw):
    if isinstance(shape, (int, np.integer)):
        shape = (shape,)
    return NDArray(np.zeros(tuple(shape), dtype=f32), ctx)


def _zeros_like(data, **kw):
    return NDArray(np.zeros_like(_raw(data)), kw.get("ctx"))


def _round(x):
    a = _raw(x)
    return NDArray((np.sign(a) * np.floor(np.abs(a).astype(np.float64) + 0.5)).astype(f32),
                   getattr(x, "context", None))


def _grid_generator(data, transform_type, target_shape):
    from oracle import zoom_np

    assert transform_type == "affine"
    m = _raw(data).reshape(-1, 6)
    h, w = target_shape
    out = np.zeros((m.shape[0], 2, h, w), dtype=f32)
    for b in range(m.shape[0]):
        assert m[b, 1] == 0 and m[b, 3] == 0  # the plugins only build axis-aligned affines
        gx, gy = zoom_np.grid_generator_affine(m[b, 0], m[b, 4], m[b, 2], m[b, 5], h, w)
        out[b, 0], out[b, 1] = gx, gy
    return NDArray(out, data.context)


def _bilinear_sampler(data, grid):
    from oracle import zoom_np

    d, g = _raw(data), _raw(grid)
    out = np.zeros(d.shape[:2] + g.shape[2:], dtype=f32)
    for b in range(d.shape[0]):
        out[b] = zoom_np.bilinear_sampler(d[b], g[b, 0], g[b, 1])
    return NDArray(out, data.context)


def _batch_dot(lhs, rhs, transpose_a=False, transpose_b=False):
    a, b = _raw(lhs), _raw(rhs)
    if transpose_a:
        a = a.transpose(0, 2, 1)
    if transpose_b:
        b = b.transpose(0, 2, 1)
    return NDArray(np.matmul(a, b).astype(f32), lhs.context)


def _slice_axis(data, axis, begin, end):
    idx = [slice(None)] * _raw(data).ndim
    idx[axis] = slice(begin, end)
    return NDArray(np.ascontiguousarray(_raw(data)[tuple(idx)]), data.context)


def _make_nd():
    nd = types.ModuleType("mxnet.ndarray")
    nd.NDArray = NDArray
    nd.array = _array
    nd.zeros = _zeros
    nd.zeros_like = _zeros_like
    nd.round = _round
    nd.GridGenerator = _grid_generator
    nd.BilinearSampler = _bilinear_sampler
    nd.batch_dot = _batch_dot
    nd.slice_axis = _slice_axis
    nd.add = lambda a, b: NDArray((_raw(a) + _raw(b)).astype(f32), a.context)
    nd.maximum = lambda a, b: NDArray(np.maximum(_raw(a), _raw(b)).astype(f32), a.context)
    nd.exp = lambda a: NDArray(np.exp(_raw(a)).astype(f32), a.context)
    nd.sum = lambda a, axis=None: NDArray(np.sum(_raw(a), axis=axis, dtype=f32), a.context)
    nd.expand_dims = lambda a, axis: NDArray(np.expand_dims(_raw(a), axis), a.context)
    nd.transpose = lambda a, axes=None: NDArray(np.ascontiguousarray(np.transpose(_raw(a), axes)),
                                                a.context)
    return nd


# ------------------------------------------------------------ mx.operator
REGISTRY = {}


class CustomOp(object):
    def __init__(self):
        pass

    def assign(self, dst, req, src):
        if req == "null":
            return
        if req in ("write", "inplace"):
            dst[:] = src
        elif req == "add":
            dst[:] += src


class CustomOpProp(object):
    def __init__(self, need_top_grad=False):
        self.need_top_grad_ = need_top_grad


def _register(name):
    def deco(cls):
        REGISTRY[name] = cls
        return cls

    return deco


def install():
    """Put the stand-in under sys.modules['mxnet'] (idempotent) and return it."""
    if "mxnet" in sys.modules and getattr(sys.modules["mxnet"], "_rc_stub", False):
        return sys.modules["mxnet"]
    mx = types.ModuleType("mxnet")
    mx._rc_stub = True
    mx.nd = mx.ndarray = _make_nd()
    op = types.ModuleType("mxnet.operator")
    op.CustomOp, op.CustomOpProp, op.register = CustomOp, CustomOpProp, _register
    mx.operator = op
    mx.cpu = lambda i=0: Context("cpu", i)
    mx.gpu = lambda i=0: Context("gpu", i)
    sys.modules["mxnet"] = mx
    sys.modules["mxnet.ndarray"] = mx.nd
    sys.modules["mxnet.operator"] = op
    return mx
