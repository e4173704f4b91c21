"""render_py_light_modelnet_multi.Render_Py_Light_ModelNet_Multi."""
from ..render import Render_Py_Light_ModelNet_Multi  # noqa: F401
