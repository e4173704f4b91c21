"""render_py_depth.Render_Py_depth: single-model renderer whose render() has no K override and
returns a uint8 image (/root/reference/lib/render_glumpy/render_py_depth.py:47-128)."""
from ..render import Render_Py_Depth as Render_Py_depth  # noqa: F401

Render_Py = Render_Py_depth  # earlier spelling of this module's export, kept for callers
