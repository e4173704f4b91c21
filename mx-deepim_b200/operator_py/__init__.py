"""CustomOp plugins of the render-and-compare path (reference: deepim/operator_py).

Importing this package registers every operator with ``mx_deepim_b200.operator``."""
from . import (transform3d, zoom_depth, zoom_flow, zoom_image, zoom_image_with_factor,  # noqa: F401
               zoom_mask, zoom_mask_with_factor, zoom_trans)
