"""render_py_light_modelnet.Render_Py_Light_ModelNet
(/root/reference/lib/render_glumpy/render_py_light_modelnet.py:82-211)."""
from ..render import Render_Py_Light_ModelNet  # noqa: F401
