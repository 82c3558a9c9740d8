"""paddlepaddle_b200 — B200-native (sm_100a) implementation of PaddlePaddle's PHI
broadcast/elementwise + reduce hot path, behind Paddle's own operator names.

    import paddlepaddle_b200 as paddle
    paddle.add(x, y); paddle.sum(x, axis=-1); paddle.nn.functional.layer_norm(x, [8192], w, b)

The compute lives in libpdb200.so (hand-written CUDA, C-ABI in include/pdb200.h).  Importing
this package does not need a GPU; calling an op without the built library or without a CUDA
device raises — there is no fallback.
"""
from . import distributed, nn  # noqa: F401
from .ops import (add, add_, argmax, divide, max, maximum, mean, min, minimum, multiply, multiply_, prod,  # noqa: F401,A004
                  subtract, sum)

__all__ = ["add", "add_", "multiply", "multiply_", "subtract", "divide", "maximum", "minimum", "sum", "mean", "max",
           "min", "prod", "argmax", "nn"]
