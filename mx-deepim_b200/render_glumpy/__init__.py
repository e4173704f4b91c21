"""Drop-in module layout of /root/reference/lib/render_glumpy: the same module names, class
names, constructor and `render` signatures and return values, backed by the CUDA rasteriser
(csrc/raster.cu) instead of glumpy / OpenGL.

    render_py.Render_Py                                   render_py.py:48-128
    render_py_depth.Render_Py_depth                       render_py_depth.py:47-128
    render_py_multi.Render_Py                             render_py_multi.py:49-147
    render_py_light.Render_Py_Light                       render_py_light.py:80-213
    render_py_light_multi_program.Render_Py_Light_MultiProgram   render_py_light_multi_program.py:84-240
    render_py_light_modelnet.Render_Py_Light_ModelNet     render_py_light_modelnet.py:82-211
    render_py_light_modelnet_multi.Render_Py_Light_ModelNet_Multi  render_py_light_modelnet_multi.py:82-231
"""
