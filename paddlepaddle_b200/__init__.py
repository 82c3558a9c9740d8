"""paddlepaddle_b200 — B200-native (sm_100a) implementation of PaddlePaddle's PHI
broadcast/elementwise + reduce hot path, behind Paddle's own operator names.

    import paddlepaddle_b200 as paddle
    paddle.add(x, y); paddle.sum(x, axis=-1); paddle.nn.functional.layer_norm(x, [8192], w, b)

The compute lives in libpdb200.so (hand-written CUDA, C-ABI in include/pdb200.h).  Importing
this package does not need a GPU; calling an op without the built library or without a CUDA
device raises — there is no fallback.
"""
from . import distributed, incubate, nn  # noqa: F401
from .ops import (add, add_, add_grad, argmax, argmin, bitwise_and, bitwise_left_shift, bitwise_not, bitwise_or, bitwise_right_shift, bitwise_xor, cast, divide, equal,  # noqa: F401,A004
                  greater_equal, greater_than, less_equal, less_than, logical_and, logical_not, logical_or,
                  logical_xor, max, max_grad, maximum, mean, mean_grad, min, min_grad, minimum, multiply, multiply_, multiply_grad, not_equal, prod, prod_grad, scale,
                  subtract, sum, sum_grad, where)

__all__ = ["add", "add_", "multiply", "multiply_", "subtract", "divide", "maximum", "minimum", "sum", "mean", "max",
           "min", "prod", "argmax", "argmin", "less_than", "less_equal", "greater_than", "greater_equal", "equal", "not_equal",
           "logical_and", "logical_or", "logical_xor", "logical_not", "bitwise_and", "bitwise_or", "bitwise_xor",
           "bitwise_not", "bitwise_left_shift", "bitwise_right_shift", "scale", "cast", "where", "add_grad", "multiply_grad", "sum_grad", "mean_grad", "max_grad", "min_grad", "prod_grad", "nn"]
