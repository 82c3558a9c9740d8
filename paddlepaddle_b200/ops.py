"""Host-side mirror of Paddle's Python operator API for the hot path.

Same names, argument meaning and error behaviour as the reference front-end:
  paddle.add / paddle.multiply        python/paddle/tensor/math.py:697, :1203 (multiply)
  paddle.sum                          python/paddle/tensor/math.py:1494
  paddle.max / min / prod             python/paddle/tensor/math.py:2906 / 3065 / 4830
  paddle.mean                         python/paddle/tensor/stat.py:45
  paddle.argmax                       python/paddle/tensor/search.py:175
What Paddle does in C++ between `_C_ops.<op>` and the kernel (InferMeta: output shape and
dtype; `dev_ctx.Alloc`; scratch through phi::Empty) is done here in Python, then the call goes
through the C-ABI (include/pdb200.h) with raw device pointers.  torch is used ONLY as the
owner of device memory and of the CUDA stream; no torch operator computes anything here.
"""
import builtins
import ctypes

import torch

from . import _lib as L

_TORCH2PDB = {
    torch.bool: L.BOOL, torch.uint8: L.UINT8, torch.int8: L.INT8, torch.int16: L.INT16,
    torch.int32: L.INT32, torch.int64: L.INT64, torch.float16: L.FLOAT16,
    torch.bfloat16: L.BFLOAT16, torch.float32: L.FLOAT32, torch.float64: L.FLOAT64,
}
_STR2TORCH = {
    "bool": torch.bool, "uint8": torch.uint8, "int8": torch.int8, "int16": torch.int16,
    "int32": torch.int32, "int64": torch.int64, "float16": torch.float16,
    "bfloat16": torch.bfloat16, "float32": torch.float32, "float64": torch.float64,
}


def _dtype_code(dt):
    try:
        return _TORCH2PDB[dt]
    except KeyError:
        raise TypeError(f"unsupported dtype {dt}") from None


def _as_torch_dtype(dt):
    if dt is None or isinstance(dt, torch.dtype):
        return dt
    s = str(dt)
    s = s.split(".")[-1]
    if s in _STR2TORCH:
        return _STR2TORCH[s]
    raise TypeError(f"unsupported dtype {dt!r}")


def _require_cuda(*tensors):
    for t in tensors:
        if t is None:
            continue
        if not isinstance(t, torch.Tensor):
            raise TypeError(f"expected a Tensor, got {type(t).__name__}")
        if not t.is_cuda:
            # The reference would dispatch to its CPU kernel; this build has none, on purpose.
            raise RuntimeError("paddlepaddle_b200 has no CPU kernels: tensors must live on a CUDA device")


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream():
    """The current CUDA stream of the current device as a void* (dev_ctx.stream() in the reference)."""
    if _raw_stream is not None:  # one C call instead of building a torch.cuda.Stream object per launch
        return ctypes.c_void_p(_raw_stream(torch.cuda.current_device()))
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


_workspaces = {}


def _workspace(nbytes, device):
    """Scratch from the caching allocator (stands in for phi::Empty on dev_ctx)."""
    if nbytes == 0:
        return None
    key = (device.index, torch.cuda.current_stream().cuda_stream)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(builtins.max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = buf
    return buf


# ---------------------------------------------------------------------------------------
# elementwise
# ---------------------------------------------------------------------------------------
_bcast_cache = {}


def _bcast_plan(xs, ys, lib):
    """(x dims array, rank, y dims array, rank, output shape) for a pair of shapes — InferMeta's broadcast rule, asked of
    the library once per shape pair and remembered (a call then costs one dict lookup instead of three ctypes trips)."""
    key = (xs, ys)
    plan = _bcast_cache.get(key)
    if plan is None:
        xd, xr = L.dims_array(xs)
        yd, yr = L.dims_array(ys)
        od = (ctypes.c_int64 * L.MAX_RANK)()
        orank = ctypes.c_int(0)
        L.check(lib.pdb200_broadcast_shape(xd, xr, yd, yr, od, ctypes.byref(orank)), lib)
        plan = (xd, xr, yd, yr, tuple(od[i] for i in range(orank.value)))
        if len(_bcast_cache) < 4096:
            _bcast_cache[key] = plan
    return plan


def _elementwise(fn_name, x, y, out=None, op=None, out_dtype=None):
    """One binary broadcast call.  `op`: leading op-code argument of the coded entry points (compare /
    logical / bitwise); `out_dtype`: dtype of the result when it is not x's (bool for compare / logical)."""
    _require_cuda(x, y)
    if x.dtype != y.dtype:
        raise TypeError(f"(InvalidType) the dtypes of X ({x.dtype}) and Y ({y.dtype}) must match "
                        "(type promotion is done by the caller in Paddle's eager layer)")
    lib = L.load()
    if not x.is_contiguous():
        x = x.contiguous()
    if not y.is_contiguous():
        y = y.contiguous()
    xd, xr, yd, yr, oshape = _bcast_plan(tuple(x.shape), tuple(y.shape), lib)
    if out is None:
        out = torch.empty(oshape, dtype=out_dtype or x.dtype, device=x.device)
    elif tuple(out.shape) != oshape:
        # test_elementwise_add_op.py:695-720 (inplace broadcast errors)
        raise ValueError(f"(InvalidArgument) The shape of tensor assigned value must be the same: {tuple(out.shape)} vs {oshape}")
    lead = () if op is None else (tuple(op) if isinstance(op, tuple) else (op,))
    L.check(getattr(lib, fn_name)(*lead, _ptr(x), xd, xr, _ptr(y), yd, yr, _ptr(out), _dtype_code(x.dtype), _stream()),
            lib)
    return out


def add(x, y, name=None):
    """paddle.add(x, y): elementwise x + y with NumPy broadcasting."""
    return _elementwise("pdb200_add", x, y)


def add_(x, y, name=None):
    """In-place add: x must already have the broadcast shape."""
    return _elementwise("pdb200_add", x, y, out=x)


def multiply(x, y, name=None):
    """paddle.multiply(x, y): elementwise x * y with NumPy broadcasting (bool: logical AND)."""
    return _elementwise("pdb200_multiply", x, y)


def multiply_(x, y, name=None):
    return _elementwise("pdb200_multiply", x, y, out=x)


def subtract(x, y, name=None):
    """paddle.subtract(x, y) (python/paddle/tensor/math.py): x - y with broadcasting."""
    return _elementwise("pdb200_subtract", x, y)


def divide(x, y, name=None):
    """paddle.divide(x, y): true division for floats, truncating division for integers."""
    return _elementwise("pdb200_divide", x, y)


def maximum(x, y, name=None):
    """paddle.maximum(x, y): elementwise `x > y ? x : y`."""
    return _elementwise("pdb200_maximum", x, y)


def minimum(x, y, name=None):
    """paddle.minimum(x, y): elementwise `x < y ? x : y`."""
    return _elementwise("pdb200_minimum", x, y)


# compare / logical / bitwise: python/paddle/tensor/logic.py (equal :476, less_than :884, logical_and :132,
# bitwise_and :1263, ...).  Compare and logical results are bool tensors.
def less_than(x, y, name=None):
    return _elementwise("pdb200_compare", x, y, op=L.LESS_THAN, out_dtype=torch.bool)


def less_equal(x, y, name=None):
    return _elementwise("pdb200_compare", x, y, op=L.LESS_EQUAL, out_dtype=torch.bool)


def greater_than(x, y, name=None):
    return _elementwise("pdb200_compare", x, y, op=L.GREATER_THAN, out_dtype=torch.bool)


def greater_equal(x, y, name=None):
    return _elementwise("pdb200_compare", x, y, op=L.GREATER_EQUAL, out_dtype=torch.bool)


def equal(x, y, name=None):
    """paddle.equal: for float32/float64 "equal" means |x - y| < 1e-8 (the reference's EqualFunctor)."""
    return _elementwise("pdb200_compare", x, y, op=L.EQUAL, out_dtype=torch.bool)


def not_equal(x, y, name=None):
    return _elementwise("pdb200_compare", x, y, op=L.NOT_EQUAL, out_dtype=torch.bool)


def logical_and(x, y, out=None, name=None):
    return _elementwise("pdb200_logical", x, y, out=out, op=L.LOGICAL_AND, out_dtype=torch.bool)


def logical_or(x, y, out=None, name=None):
    return _elementwise("pdb200_logical", x, y, out=out, op=L.LOGICAL_OR, out_dtype=torch.bool)


def logical_xor(x, y, out=None, name=None):
    return _elementwise("pdb200_logical", x, y, out=out, op=L.LOGICAL_XOR, out_dtype=torch.bool)


def bitwise_and(x, y, out=None, name=None):
    return _elementwise("pdb200_bitwise", x, y, out=out, op=L.BITWISE_AND)


def bitwise_or(x, y, out=None, name=None):
    return _elementwise("pdb200_bitwise", x, y, out=out, op=L.BITWISE_OR)


def bitwise_xor(x, y, out=None, name=None):
    return _elementwise("pdb200_bitwise", x, y, out=out, op=L.BITWISE_XOR)


def bitwise_left_shift(x, y, is_arithmetic=True, out=None, name=None):
    """paddle.bitwise_left_shift (python/paddle/tensor/math.py): x << y; counts outside [0, bits) give 0."""
    return _elementwise("pdb200_bitwise_shift", x, y, out=out, op=(0, int(bool(is_arithmetic))))


def bitwise_right_shift(x, y, is_arithmetic=True, out=None, name=None):
    """paddle.bitwise_right_shift: sign-filling when is_arithmetic, else a shift of the unsigned representation."""
    return _elementwise("pdb200_bitwise_shift", x, y, out=out, op=(1, int(bool(is_arithmetic))))


# ---------------------------------------------------------------------------------------
# backward of add / multiply (what `_C_ops.add_grad` / `_C_ops.multiply_grad` run under autograd;
# paddle/phi/ops/yaml/backward.yaml: add_grad(x, y, out_grad, axis=-1) -> (x_grad, y_grad))
# ---------------------------------------------------------------------------------------
def _binary_grad(op, x, y, out_grad, axis, want_x, want_y):
    _require_cuda(x, y, out_grad)
    if axis != -1:
        raise NotImplementedError("(Unimplemented) only axis=-1 (NumPy right-aligned broadcasting) is on this path")
    if not (x.dtype == y.dtype == out_grad.dtype):
        raise TypeError("(InvalidType) x, y and out_grad must share one dtype")
    lib = L.load()
    x, y, out_grad = x.contiguous(), y.contiguous(), out_grad.contiguous()
    xd, xr = L.dims_array(x.shape)
    yd, yr = L.dims_array(y.shape)
    od = (ctypes.c_int64 * L.MAX_RANK)()
    orank = ctypes.c_int(0)
    L.check(lib.pdb200_broadcast_shape(xd, xr, yd, yr, od, ctypes.byref(orank)), lib)
    oshape = tuple(od[i] for i in range(orank.value))
    if tuple(out_grad.shape) != oshape:
        raise ValueError(f"(InvalidArgument) out_grad has shape {tuple(out_grad.shape)}, the forward output has {oshape}")
    dx = torch.empty_like(x) if want_x else None
    dy = torch.empty_like(y) if want_y else None
    code = _dtype_code(x.dtype)
    nbytes = lib.pdb200_elementwise_grad_workspace_bytes(op, xd, xr, yd, yr, code, int(want_x), int(want_y))
    ws = _workspace(nbytes, x.device)
    if op == L.GRAD_ADD:
        rc = lib.pdb200_add_grad(xd, xr, yd, yr, _ptr(out_grad), code, _ptr(dx), _ptr(dy), _ptr(ws), nbytes, _stream())
    else:
        rc = lib.pdb200_multiply_grad(_ptr(x), xd, xr, _ptr(y), yd, yr, _ptr(out_grad), code, _ptr(dx), _ptr(dy), _ptr(ws),
                                      nbytes, _stream())
    L.check(rc, lib)
    return dx, dy


def add_grad(x, y, out_grad, axis=-1, want_x_grad=True, want_y_grad=True):
    """(x_grad, y_grad) of out = x + y: out_grad summed over the dimensions each operand was broadcast along."""
    return _binary_grad(L.GRAD_ADD, x, y, out_grad, axis, want_x_grad, want_y_grad)


def multiply_grad(x, y, out_grad, axis=-1, want_x_grad=True, want_y_grad=True):
    """(x_grad, y_grad) of out = x * y: sum(out_grad * y) and sum(out_grad * x) over the broadcast dimensions."""
    return _binary_grad(L.GRAD_MULTIPLY, x, y, out_grad, axis, want_x_grad, want_y_grad)


# ---------------------------------------------------------------------------------------
# one-input maps and where
# ---------------------------------------------------------------------------------------
def _unary(fn_name, x, out_dtype, *mid):
    _require_cuda(x)
    lib = L.load()
    x = x.contiguous()
    out = torch.empty(x.shape, dtype=out_dtype, device=x.device)
    L.check(getattr(lib, fn_name)(_ptr(x), x.numel(), _dtype_code(x.dtype), *mid, _ptr(out), _stream()), lib)
    return out


def logical_not(x, out=None, name=None):
    """paddle.logical_not (python/paddle/tensor/logic.py:331): bool tensor, true where x == 0."""
    return _unary("pdb200_logical_not", x, torch.bool)


def bitwise_not(x, out=None, name=None):
    """paddle.bitwise_not (logic.py:1500): ~x for integers, !x for bool."""
    return _unary("pdb200_bitwise_not", x, x.dtype)


def scale(x, scale=1.0, bias=0.0, bias_after_scale=True, act=None, name=None):
    """paddle.scale (python/paddle/tensor/math.py:243): scale*x + bias, or scale*(x + bias)."""
    if act is not None:
        raise NotImplementedError("(Unimplemented) scale(act=...) is a static-graph legacy option outside this path")
    if isinstance(scale, torch.Tensor):
        scale = scale.item()  # ScaleKernel takes a Scalar; a tensor scale is read on the host in the reference too
    return _unary("pdb200_scale", x, x.dtype, float(scale), float(bias), int(bool(bias_after_scale)))


def cast(x, dtype):
    """paddle.cast (python/paddle/tensor/manipulation.py:200): static_cast per element; same dtype returns a copy."""
    dt = _as_torch_dtype(dtype)
    return _unary("pdb200_cast", x, dt, _dtype_code(dt))


def where(condition, x, y, name=None):
    """paddle.where(condition, x, y) (python/paddle/tensor/search.py:704).  WhereKernel takes three tensors of ONE
    shape; the reference's Python layer first materialises differently-shaped operands with its `expand` op
    (search.py:776-797), which is outside this path — so here the three shapes must already agree."""
    _require_cuda(condition, x, y)
    if condition.dtype != torch.bool:
        raise TypeError("(InvalidType) the dtype of condition must be bool")
    if x.dtype != y.dtype:
        raise TypeError(f"(InvalidType) the dtypes of X ({x.dtype}) and Y ({y.dtype}) must match")
    if not (tuple(condition.shape) == tuple(x.shape) == tuple(y.shape)):
        raise ValueError("(InvalidArgument) where: condition, x and y must have the same shape "
                         f"(got {tuple(condition.shape)}, {tuple(x.shape)}, {tuple(y.shape)}); expand them first")
    condition, x, y = condition.contiguous(), x.contiguous(), y.contiguous()
    lib = L.load()
    out = torch.empty(x.shape, dtype=x.dtype, device=x.device)
    L.check(lib.pdb200_where(_ptr(condition), _ptr(x), _ptr(y), out.numel(), _dtype_code(x.dtype), _ptr(out), _stream()),
            lib)
    return out


# ---------------------------------------------------------------------------------------
# reductions
# ---------------------------------------------------------------------------------------
def _norm_axes(axis, rank):
    if axis is None:
        return []
    if isinstance(axis, (list, tuple)):
        axes = [int(a) for a in axis]
    else:
        axes = [int(axis)]
    for a in axes:
        if rank == 0:
            if a not in (0, -1):
                raise ValueError(f"(InvalidArgument) axis {a} out of range for a 0-D tensor")
        elif a < -rank or a >= rank:
            raise ValueError(f"(InvalidArgument) The reduce dim index should be in the range [-{rank}, {rank}), but received {a}")
    return axes


def _reduced_shape(shape, axes, keepdim):
    """ReduceIntArrayAxisInferMeta / SumRawInferMeta shape rule (infermeta/unary.cc:3771, 4725-4753)."""
    rank = len(shape)
    if rank == 0:
        return ()
    red = set(a % rank for a in axes) if axes else set(range(rank))
    if keepdim:
        return tuple(1 if i in red else d for i, d in enumerate(shape))
    return tuple(d for i, d in enumerate(shape) if i not in red)


_REDUCE_FN = {
    L.REDUCE_SUM: "pdb200_sum_raw", L.REDUCE_MEAN: "pdb200_mean_raw", L.REDUCE_MAX: "pdb200_max_raw",
    L.REDUCE_MIN: "pdb200_min_raw", L.REDUCE_PROD: "pdb200_prod",
}


_reduce_cache = {}


def _reduce_plan(op, shape, axis, keepdim, x_dtype, out_dtype, lib):
    """Everything about a reduction that depends only on (op, shape, axes, dtypes): normalised axes, output shape, the
    dims arrays handed to the C-ABI and the workspace size.  Remembered per key, so a repeated call does no ctypes query."""
    akey = tuple(axis) if isinstance(axis, (list, tuple)) else axis
    key = (op, shape, akey, keepdim, x_dtype, out_dtype)
    plan = _reduce_cache.get(key)
    if plan is None:
        axes = _norm_axes(axis, len(shape))
        oshape = _reduced_shape(shape, axes, keepdim)
        n_out = 1
        for d_ in oshape:
            n_out *= d_
        xd, xr = L.dims_array(shape)
        ad, an = L.dims_array(axes)
        reduce_all = 1 if not axes else 0
        xc, oc = _dtype_code(x_dtype), _dtype_code(out_dtype)
        nbytes = lib.pdb200_reduce_workspace_bytes(op, xd, xr, ad, an, reduce_all, xc, oc)
        plan = (oshape, n_out, xd, xr, ad, an, reduce_all, xc, oc, nbytes, getattr(lib, _REDUCE_FN[op]))
        if len(_reduce_cache) < 4096:
            _reduce_cache[key] = plan
    return plan


def _reduce(op, x, axis, keepdim, out_dtype, out=None):
    """`out`: write the result into this buffer (right element count and dtype) instead of a new tensor — used by the
    distributed layer to let the kernel write straight into a peer-visible slot."""
    _require_cuda(x)
    lib = L.load()
    if not x.is_contiguous():
        x = x.contiguous()
    if x.numel() == 0:
        _norm_axes(axis, x.dim())
        raise ValueError("(InvalidArgument) Tensor need be reduced must not empty.")
    oshape, n_out, xd, xr, ad, an, reduce_all, xc, oc, nbytes, fn = _reduce_plan(op, tuple(x.shape), axis, bool(keepdim),
                                                                                 x.dtype, out_dtype, lib)
    if out is None:
        out = torch.empty(oshape, dtype=out_dtype, device=x.device)
    elif out.numel() != n_out or out.dtype != out_dtype or not out.is_contiguous():
        raise ValueError("(InvalidArgument) out must be a contiguous tensor of the result's element count and dtype")
    ws = _workspace(nbytes, x.device)
    L.check(fn(_ptr(x), xd, xr, ad, an, reduce_all, xc, oc, _ptr(out), _ptr(ws), nbytes, _stream()), lib)
    return out


def sum(x, axis=None, dtype=None, keepdim=False, name=None):
    """paddle.sum: bool/int32 inputs produce int64 unless `dtype` is given
    (SumRawInferMeta, infermeta/unary.cc:4743-4747)."""
    dt = _as_torch_dtype(dtype)
    if dt is None:
        dt = torch.int64 if x.dtype in (torch.bool, torch.int32) else x.dtype
    return _reduce(L.REDUCE_SUM, x, axis, keepdim, dt)


def _reduce_grad(fn_name, x, out_grad, axis, keepdim, reduce_all):
    """sum_grad / mean_grad(x, out_grad, axis, keepdim, reduce_all) -> x_grad (paddle/phi/ops/yaml/backward.yaml)."""
    _require_cuda(x, out_grad)
    if x.dtype != out_grad.dtype:
        raise TypeError("(InvalidType) x and out_grad must share one dtype")
    axes = _norm_axes(axis, x.dim())
    want = _reduced_shape(tuple(x.shape), [] if reduce_all else axes, keepdim)
    if tuple(out_grad.shape) != tuple(want):
        raise ValueError(f"(InvalidArgument) out_grad has shape {tuple(out_grad.shape)}, the forward output has {tuple(want)}")
    lib = L.load()
    out_grad = out_grad.contiguous()
    x_grad = torch.empty(x.shape, dtype=x.dtype, device=x.device)
    xd, xr = L.dims_array(x.shape)
    ad, an = L.dims_array(axes)
    L.check(getattr(lib, fn_name)(_ptr(out_grad), xd, xr, ad, an, int(bool(reduce_all) or not axes), _dtype_code(x.dtype),
                                  _ptr(x_grad), _stream()), lib)
    return x_grad


def sum_grad(x, out_grad, axis=None, keepdim=False, reduce_all=False):
    """Backward of paddle.sum: out_grad repeated along the reduced axes (only x's shape and dtype are used)."""
    return _reduce_grad("pdb200_sum_grad", x, out_grad, axis, keepdim, reduce_all)


def mean_grad(x, out_grad, axis=None, keepdim=False, reduce_all=False):
    """Backward of paddle.mean: out_grad / n repeated along the reduced axes."""
    return _reduce_grad("pdb200_mean_grad", x, out_grad, axis, keepdim, reduce_all)


def _select_grad(fn_name, x, out, out_grad, axis, keepdim, reduce_all):
    """max_grad / min_grad / prod_grad(x, out, out_grad, axis, keepdim, reduce_all) -> x_grad (backward.yaml)."""
    _require_cuda(x, out, out_grad)
    if not (x.dtype == out.dtype == out_grad.dtype):
        raise TypeError("(InvalidType) x, out and out_grad must share one dtype")
    axes = _norm_axes(axis, x.dim())
    want = _reduced_shape(tuple(x.shape), [] if reduce_all else axes, keepdim)
    for name, t in (("out", out), ("out_grad", out_grad)):
        if tuple(t.shape) != tuple(want):
            raise ValueError(f"(InvalidArgument) {name} has shape {tuple(t.shape)}, the forward output has {tuple(want)}")
    lib = L.load()
    x, out, out_grad = x.contiguous(), out.contiguous(), out_grad.contiguous()
    x_grad = torch.empty(x.shape, dtype=x.dtype, device=x.device)
    xd, xr = L.dims_array(x.shape)
    ad, an = L.dims_array(axes)
    L.check(getattr(lib, fn_name)(_ptr(x), xd, xr, _ptr(out), _ptr(out_grad), ad, an, int(bool(reduce_all) or not axes),
                                  _dtype_code(x.dtype), _ptr(x_grad), _stream()), lib)
    return x_grad


def max_grad(x, out, out_grad, axis=None, keepdim=False, reduce_all=False):
    """Backward of paddle.max: out_grad at every position that holds the maximum (ties all receive it), 0 elsewhere."""
    return _select_grad("pdb200_max_grad", x, out, out_grad, axis, keepdim, reduce_all)


def min_grad(x, out, out_grad, axis=None, keepdim=False, reduce_all=False):
    """Backward of paddle.min."""
    return _select_grad("pdb200_min_grad", x, out, out_grad, axis, keepdim, reduce_all)


def prod_grad(x, out, out_grad, axis=None, keepdim=False, reduce_all=False):
    """Backward of paddle.prod: out_grad * out / x."""
    return _select_grad("pdb200_prod_grad", x, out, out_grad, axis, keepdim, reduce_all)


def mean(x, axis=None, keepdim=False, name=None):
    return _reduce(L.REDUCE_MEAN, x, axis, keepdim, x.dtype)


def max(x, axis=None, keepdim=False, name=None):
    return _reduce(L.REDUCE_MAX, x, axis, keepdim, x.dtype)


def min(x, axis=None, keepdim=False, name=None):
    return _reduce(L.REDUCE_MIN, x, axis, keepdim, x.dtype)


def prod(x, axis=None, keepdim=False, dtype=None, name=None):
    dt = _as_torch_dtype(dtype)
    if dt is not None and dt != x.dtype:
        raise NotImplementedError("prod(dtype=...) casts first in Paddle (math.py:4830); cast the input yourself")
    return _reduce(L.REDUCE_PROD, x, axis, keepdim, x.dtype)


def _arg(fn_name, x, axis, keepdim, dtype):
    _require_cuda(x)
    lib = L.load()
    x = x.contiguous()
    odt = _as_torch_dtype(dtype)
    if odt not in (torch.int64, torch.int32):
        raise ValueError("(InvalidArgument) The attribute of dtype in argmin/argmax must be [int32] or [int64]")
    if x.numel() == 0:
        raise ValueError("(InvalidArgument) argmin/argmax input numel must > 0")
    rank = x.dim()
    flatten = axis is None
    ax = 0 if flatten else int(axis)
    if not flatten and rank > 0 and (ax < -rank or ax >= rank):
        raise ValueError(f"(InvalidArgument) 'axis'({ax}) must be in [-{rank}, {rank})")
    if flatten or rank == 0:
        oshape = tuple([1] * rank) if (keepdim and flatten) else ()
    else:
        a = ax % rank
        oshape = tuple(1 if i == a else d for i, d in enumerate(x.shape)) if keepdim else \
            tuple(d for i, d in enumerate(x.shape) if i != a)
    out = torch.empty(oshape, dtype=odt, device=x.device)
    xd, xr = L.dims_array(x.shape)
    xc = _dtype_code(x.dtype)
    nbytes = lib.pdb200_argmax_workspace_bytes(xd, xr, ax, int(flatten), xc)
    ws = _workspace(nbytes, x.device)
    L.check(getattr(lib, fn_name)(_ptr(x), xd, xr, ax, int(flatten), xc, _dtype_code(odt), _ptr(out), _ptr(ws),
                                  nbytes, _stream()), lib)
    return out


def argmax(x, axis=None, keepdim=False, dtype="int64", name=None):
    """paddle.argmax: axis=None flattens; lowest index wins ties; dtype int64 or int32
    (ArgMinMaxInferMeta, infermeta/unary.cc:257)."""
    return _arg("pdb200_argmax", x, axis, keepdim, dtype)


def argmin(x, axis=None, keepdim=False, dtype="int64", name=None):
    """paddle.argmin (python/paddle/tensor/search.py:283): the same rules with the minimum."""
    return _arg("pdb200_argmin", x, axis, keepdim, dtype)
