"""render_py_depth.Render_Py: single-model renderer whose render() has no K override and
returns a uint8 image (/root/reference/lib/render_glumpy/render_py_depth.py:48-128)."""
from ..render import Render_Py_Depth as Render_Py  # noqa: F401
