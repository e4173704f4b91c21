"""render_py_multi.Render_Py (multi-class textured renderer)."""
from ..render import Render_Py  # noqa: F401
