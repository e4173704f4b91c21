"""render_py_light.Render_Py_Light (/root/reference/lib/render_glumpy/render_py_light.py:80-213)."""
from ..render import Render_Py_Light  # noqa: F401
