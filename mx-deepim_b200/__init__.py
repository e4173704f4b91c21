"""rendercompare-b200: the per-iteration render-and-compare data path of DeepIM
(reference: wangg12/mx-DeepIM) as hand-written sm_100a kernels behind the
reference's own plugin surfaces.

Sub-modules
-----------
flow          gpu_flow / gpu_flow_wrapper          (reference lib/flow_c/flow.py)
operator      mx.operator-shaped CustomOp shim     (reference uses mxnet.operator)
operator_py   ZoomMask, ZoomImage, ... Transform3D (reference deepim/operator_py)
render        Render_Py, Render_Py_Light_ModelNet_Multi (reference lib/render_glumpy)
pipeline      fused on-device render -> flow -> zoom -> loss iteration
"""
__version__ = "0.1.0"
