"""paddle.incubate names of the fused ops on this path (python/paddle/incubate/operators/__init__.py:20-23,
python/paddle/incubate/nn/functional/__init__.py): thin aliases of the functions in nn/functional.py."""
from .nn.functional import (fused_bias_dropout_residual_layer_norm, fused_layer_norm, fused_rms_norm, softmax_mask_fuse,  # noqa: F401
                            softmax_mask_fuse_upper_triangle)
