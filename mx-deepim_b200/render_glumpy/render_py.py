"""render_py.Render_Py: single-model textured renderer
(/root/reference/lib/render_glumpy/render_py.py:48-128)."""
from ..render import Render_Py_Single as Render_Py  # noqa: F401
