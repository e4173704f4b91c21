"""render_py_light_multi_program.Render_Py_Light_MultiProgram
(/root/reference/lib/render_glumpy/render_py_light_multi_program.py:84-240)."""
from ..render import Render_Py_Light_MultiProgram  # noqa: F401
