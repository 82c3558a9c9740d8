"""paddle.nn.functional.softmax / layer_norm on the B200 kernels.

Mirrors python/paddle/nn/functional/activation.py:1111 (softmax) and
python/paddle/nn/functional/norm.py:320 (layer_norm); the C++ in between (SoftmaxInferMeta,
LayerNormInferMeta, Alloc) is done here, then the call goes through the C-ABI.
"""
import ctypes
import numbers

import torch

from .. import _lib as L
from ..ops import _as_torch_dtype, _dtype_code, _ptr, _require_cuda, _stream, _workspace


def softmax(x, axis=-1, dtype=None, name=None):
    """softmax along `axis`.  `dtype` casts the input first (activation.py:1245-1247); the cast itself
    is not on the hot path and only the no-cast call is supported here."""
    _require_cuda(x)
    dt = _as_torch_dtype(dtype)
    if dt is not None and dt != x.dtype:
        raise NotImplementedError("softmax(dtype=...) needs phi::CastKernel, which is outside the hot path")
    lib = L.load()
    x = x.contiguous()
    out = torch.empty_like(x)
    d, r = L.dims_array(x.shape)
    L.check(lib.pdb200_softmax(_ptr(x), d, r, int(axis), _dtype_code(x.dtype), _ptr(out), _stream()), lib)
    return out


def log_softmax(x, axis=-1, dtype=None, name=None):
    """log(softmax(x)) along `axis` (python/paddle/nn/functional/activation.py: log_softmax), computed in one pass as
    (x - max) - log(sum(exp(x - max)))."""
    _require_cuda(x)
    dt = _as_torch_dtype(dtype)
    if dt is not None and dt != x.dtype:
        raise NotImplementedError("log_softmax(dtype=...): cast first with paddle.cast")
    lib = L.load()
    x = x.contiguous()
    out = torch.empty_like(x)
    d, r = L.dims_array(x.shape)
    L.check(lib.pdb200_log_softmax(_ptr(x), d, r, int(axis), _dtype_code(x.dtype), _ptr(out), _stream()), lib)
    return out


def softmax_mask_fuse(x, mask, name=None):
    """paddle.incubate.softmax_mask_fuse (python/paddle/incubate/operators/softmax_mask_fuse.py:26): softmax(x + mask)
    along the last axis for attention scores x [batch, heads, q_len, k_len] and a mask [batch, 1, q_len, k_len] shared by
    the heads, in one pass (the sum is never written)."""
    _require_cuda(x, mask)
    if x.dim() != 4 or mask.dim() != 4:
        raise ValueError("(InvalidArgument) softmax_mask_fuse expects 4-D x [batch, heads, q_len, k_len] and 4-D mask")
    if mask.dtype != x.dtype and mask.dtype != torch.float32:
        raise TypeError("(InvalidType) the mask must have the dtype of x or float32")
    lib = L.load()
    x, mask = x.contiguous(), mask.contiguous()
    out = torch.empty_like(x)
    xd, _ = L.dims_array(x.shape)
    md, _ = L.dims_array(mask.shape)
    L.check(lib.pdb200_fused_softmax_mask(_ptr(x), xd, _ptr(mask), md, _dtype_code(x.dtype), _dtype_code(mask.dtype), _ptr(out),
                                          _stream()), lib)
    return out


def softmax_mask_fuse_upper_triangle(x):
    """paddle.incubate.softmax_mask_fuse_upper_triangle (python/paddle/incubate/operators/softmax_mask_fuse_upper_triangle.py:26):
    causal softmax of square attention scores x [batch, heads, seq, seq] — row q is normalised over columns <= q, the rest
    of the row is zero."""
    _require_cuda(x)
    if x.dim() != 4:
        raise ValueError("(InvalidArgument) softmax_mask_fuse_upper_triangle expects a 4-D x [batch, heads, seq, seq]")
    lib = L.load()
    x = x.contiguous()
    out = torch.empty_like(x)
    xd, _ = L.dims_array(x.shape)
    L.check(lib.pdb200_fused_softmax_mask_upper_triangle(_ptr(x), xd, _dtype_code(x.dtype), _ptr(out), _stream()), lib)
    return out


def softmax_with_cross_entropy(logits, label, soft_label=False, ignore_index=-100, numeric_stable_mode=True,
                               return_softmax=False, axis=-1):
    """paddle.nn.functional.softmax_with_cross_entropy (python/paddle/nn/functional/loss.py): loss, or (loss, softmax).
    Hard labels: integer class indices with the logits' shape except a 1 along `axis`; soft labels: distributions with
    the logits' shape and dtype."""
    return _cross_entropy(logits, label, soft_label, True, ignore_index, axis, return_softmax)


def _cross_entropy(logits, label, soft_label, use_softmax, ignore_index, axis, return_softmax):
    _require_cuda(logits, label)
    rank = logits.dim()
    if rank == 0:
        raise ValueError("(InvalidArgument) cross_entropy needs at least one dimension")
    if not -rank <= axis < rank:
        raise ValueError(f"(InvalidArgument) The axis is expected to be in range of [{-rank}, {rank}), but got {axis}")
    ax = axis % rank
    lshape = tuple(1 if i == ax else d for i, d in enumerate(logits.shape))
    if soft_label:
        if tuple(label.shape) != tuple(logits.shape) or label.dtype != logits.dtype:
            raise ValueError("(InvalidArgument) soft labels must have the shape and dtype of the logits")
    else:
        if label.dtype not in (torch.int64, torch.int32):
            raise TypeError("(InvalidType) hard labels must be int64 or int32")
        if tuple(label.shape) != lshape:
            raise ValueError(f"(InvalidArgument) label has shape {tuple(label.shape)}, expected {lshape}")
    lib = L.load()
    logits, label = logits.contiguous(), label.contiguous()
    sm = torch.empty_like(logits)
    loss = torch.empty(lshape, dtype=logits.dtype, device=logits.device)
    d, r = L.dims_array(logits.shape)
    nbytes = lib.pdb200_cross_entropy_workspace_bytes(d, r, int(axis), _dtype_code(logits.dtype)) if use_softmax else 0
    ws = _workspace(nbytes, logits.device)
    L.check(lib.pdb200_cross_entropy_with_softmax(_ptr(logits), d, r, _ptr(label), _dtype_code(label.dtype), int(bool(soft_label)),
                                                  int(bool(use_softmax)), int(ignore_index), int(axis),
                                                  _dtype_code(logits.dtype), _ptr(sm), _ptr(loss), _ptr(ws), nbytes, _stream()),
            lib)
    return (loss, sm) if return_softmax else loss


def softmax_with_cross_entropy_grad(label, softmax, loss_grad, soft_label=False, ignore_index=-100, axis=-1):
    """logits_grad of softmax_with_cross_entropy (what _C_ops.cross_entropy_with_softmax_grad computes) from the saved
    softmax, the labels and the gradient of the loss (the logits' shape with a 1 along `axis`)."""
    _require_cuda(label, softmax, loss_grad)
    rank = softmax.dim()
    if rank == 0:
        raise ValueError("(InvalidArgument) cross_entropy needs at least one dimension")
    if not -rank <= axis < rank:
        raise ValueError(f"(InvalidArgument) The axis is expected to be in range of [{-rank}, {rank}), but got {axis}")
    ax = axis % rank
    lshape = tuple(1 if i == ax else d for i, d in enumerate(softmax.shape))
    if tuple(loss_grad.shape) != lshape or loss_grad.dtype != softmax.dtype:
        raise ValueError(f"(InvalidArgument) loss_grad must have shape {lshape} and the softmax's dtype")
    if soft_label:
        if tuple(label.shape) != tuple(softmax.shape) or label.dtype != softmax.dtype:
            raise ValueError("(InvalidArgument) soft labels must have the shape and dtype of the logits")
    else:
        if label.dtype not in (torch.int64, torch.int32):
            raise TypeError("(InvalidType) hard labels must be int64 or int32")
        if tuple(label.shape) != lshape:
            raise ValueError(f"(InvalidArgument) label has shape {tuple(label.shape)}, expected {lshape}")
    lib = L.load()
    label, softmax, loss_grad = label.contiguous(), softmax.contiguous(), loss_grad.contiguous()
    dx = torch.empty_like(softmax)
    d, r = L.dims_array(softmax.shape)
    L.check(lib.pdb200_cross_entropy_with_softmax_grad(_ptr(label), _dtype_code(label.dtype), _ptr(softmax), _ptr(loss_grad), d, r,
                                                       int(bool(soft_label)), 1, int(ignore_index), int(axis),
                                                       _dtype_code(softmax.dtype), _ptr(dx), _stream()), lib)
    return dx


def fused_rms_norm(x, norm_weight, norm_bias=None, epsilon=1e-6, begin_norm_axis=1, bias=None, residual=None,
                   quant_scale=-1, quant_round_type=0, quant_max_bound=0, quant_min_bound=0):
    """paddle.incubate.nn.functional.fused_rms_norm (python/paddle/incubate/nn/functional/fused_rms_norm.py):
    returns (out, inv_var) or, with a residual, (out, residual_out, inv_var); v = x + residual + bias is normalised by
    its root mean square over the trailing dims from begin_norm_axis.  Quantised outputs are not on this path."""
    _require_cuda(x, norm_weight, norm_bias, bias, residual)
    if quant_scale > 0:
        raise NotImplementedError("(Unimplemented) quantised rms_norm outputs stay with the reference kernel")
    for t in (norm_weight, norm_bias, bias, residual):
        if t is not None and t.dtype != x.dtype:
            raise TypeError("(InvalidType) every operand of rms_norm must have the dtype of x")
    if bias is not None and residual is None:
        raise ValueError("(InvalidArgument) bias is only added together with a residual")
    if residual is not None and residual.shape != x.shape:
        raise ValueError("(InvalidArgument) residual must have the shape of x")
    rank = x.dim()
    if not 0 <= begin_norm_axis <= rank:
        raise ValueError(f"(InvalidArgument) begin_norm_axis must be in [0, {rank}]")
    D = 1
    for d_ in x.shape[begin_norm_axis:]:
        D *= d_
    rows = x.numel() // D if D else 0
    for name_, t in (("norm_weight", norm_weight), ("norm_bias", norm_bias), ("bias", bias)):
        if t is not None and t.numel() != D:
            raise ValueError(f"(InvalidArgument) {name_} must have {D} elements, got {t.numel()}")
    lib = L.load()
    x = x.contiguous()
    cont = lambda t: t.contiguous() if t is not None else None
    norm_weight, norm_bias, bias, residual = cont(norm_weight), cont(norm_bias), cont(bias), cont(residual)
    out = torch.empty_like(x)
    res_out = torch.empty_like(x) if residual is not None else None
    inv_var = torch.empty(tuple(x.shape[:begin_norm_axis]), dtype=torch.float32, device=x.device)
    d, r = L.dims_array(x.shape)
    L.check(lib.pdb200_rms_norm(_ptr(x), d, r, _ptr(bias), _ptr(residual), _ptr(norm_weight), _ptr(norm_bias),
                                ctypes.c_float(epsilon), int(begin_norm_axis), _dtype_code(x.dtype), _ptr(out), _ptr(res_out),
                                _ptr(inv_var), _stream()), lib)
    return (out, res_out, inv_var) if residual is not None else (out, inv_var)


def fused_bias_dropout_residual_layer_norm(x, residual, bias=None, ln_scale=None, ln_bias=None, dropout_rate=0.5,
                                           ln_epsilon=1e-5, training=True, mode="upscale_in_train", name=None,
                                           return_all=False):
    """paddle.incubate.nn.functional.fused_bias_dropout_residual_layer_norm
    (python/paddle/incubate/nn/functional/fused_transformer.py:334): y = layer_norm(residual + dropout(bias + x)) over the
    last axis, in one pass.  return_all=True also returns the op's other outputs (bias_dropout_residual_out, ln_mean,
    ln_variance) that Paddle keeps for the backward pass.  Deterministic cases only (dropout_rate == 0 or training=False):
    a random dropout mask comes from the framework's generator, which is outside this path."""
    _require_cuda(x, residual, bias, ln_scale, ln_bias)
    if mode not in ("upscale_in_train", "downscale_in_infer"):
        raise ValueError("mode argument should be 'downscale_in_infer' or 'upscale_in_train'")  # fused_transformer.py:425-428
    if tuple(residual.shape) != tuple(x.shape) or residual.dtype != x.dtype:
        raise ValueError("(InvalidArgument) residual must have the shape and dtype of x")
    if x.dim() < 1:
        raise ValueError("(InvalidArgument) x needs at least one dimension")
    D = x.shape[-1]
    if bias is not None and (bias.numel() != D or bias.dtype != x.dtype):
        raise ValueError(f"(InvalidArgument) bias must have {D} elements of x's dtype")
    pdt = None
    for p in (ln_scale, ln_bias):
        if p is None:
            continue
        if p.numel() != D:
            raise ValueError(f"(InvalidArgument) ln_scale / ln_bias must have {D} elements, got {p.numel()}")
        if pdt is not None and p.dtype != pdt:
            raise ValueError("(InvalidArgument) ln_scale and ln_bias must share one dtype")
        pdt = p.dtype
    if pdt is not None and pdt != x.dtype and pdt != torch.float32:
        raise ValueError("(InvalidArgument) ln_scale / ln_bias must have x's dtype or float32")
    if training and dropout_rate != 0.0:
        raise NotImplementedError("(Unimplemented) training with dropout_rate > 0 draws its mask from the framework's generator; "
                                  "this path runs dropout_rate == 0 and training=False")
    lib = L.load()
    cont = lambda t: t.contiguous() if t is not None else None
    x, residual, bias, ln_scale, ln_bias = cont(x), cont(residual), cont(bias), cont(ln_scale), cont(ln_bias)
    rows = x.numel() // D if D else 0
    stat_dt = torch.float64 if x.dtype == torch.float64 else torch.float32
    y = torch.empty_like(x)
    dro = torch.empty_like(x)
    stats = torch.empty((2, rows), dtype=stat_dt, device=x.device)
    d, r = L.dims_array(x.shape)
    pcode = _dtype_code(pdt) if pdt is not None else _dtype_code(x.dtype)
    L.check(lib.pdb200_fused_bias_dropout_residual_layer_norm(
        _ptr(x), _ptr(residual), d, r, _ptr(bias), _ptr(ln_scale), _ptr(ln_bias), pcode, ctypes.c_float(float(dropout_rate)),
        int(not training), int(mode == "upscale_in_train"), ctypes.c_float(float(ln_epsilon)), _dtype_code(x.dtype), _ptr(y),
        _ptr(dro), _ptr(stats[0]), _ptr(stats[1]), _stream()), lib)
    if return_all:
        return y, dro, stats[0], stats[1]
    return y


def fused_layer_norm(x, norm_weight, norm_bias, epsilon, residual_alpha=1.0, begin_norm_axis=1, bias=None, residual=None,
                     quant_scale=-1, quant_round_type=0, quant_max_bound=0, quant_min_bound=0, return_all=False):
    """paddle.incubate.nn.functional.fused_layer_norm (python/paddle/incubate/nn/functional/fused_layer_norm.py:65) =
    the op fused_bias_residual_layernorm: layer_norm(bias + residual_alpha * residual + x) in one pass; with
    norm_weight = norm_bias = None it returns the fused sum alone.  Returns what the eager op returns: (out, residual_out,
    mean, variance) — residual_out / mean / variance are None where the kernel does not write them.  float32 parameters
    (LayerNormParamType); the quantised outputs (quant_scale > 0) are outside this path."""
    _require_cuda(x, norm_weight, norm_bias, bias, residual)
    if quant_scale > 0:
        raise NotImplementedError("(Unimplemented) quantised fused_layer_norm outputs (quant_scale > 0) are not on this path")
    nd = x.dim()
    if not (0 <= begin_norm_axis <= nd) or nd < 1:
        raise ValueError(f"(InvalidArgument) begin_norm_axis must be in [0, {nd}]")
    D = 1
    for s in x.shape[begin_norm_axis:]:
        D *= int(s)
    rows = x.numel() // D if D else 0
    for name_, p in (("norm_weight", norm_weight), ("norm_bias", norm_bias)):
        if p is not None and (p.numel() != D or p.dtype != torch.float32):
            raise ValueError(f"(InvalidArgument) {name_} must be float32 with {D} elements")
    if bias is not None and (bias.numel() != D or bias.dtype != x.dtype):
        raise ValueError(f"(InvalidArgument) bias must have {D} elements of x's dtype")
    if residual is not None and (tuple(residual.shape) != tuple(x.shape) or residual.dtype != x.dtype):
        raise ValueError("(InvalidArgument) residual must have the shape and dtype of x")
    lib = L.load()
    cont = lambda t: t.contiguous() if t is not None else None
    x, residual, bias, norm_weight, norm_bias = cont(x), cont(residual), cont(bias), cont(norm_weight), cont(norm_bias)
    has_param = norm_weight is not None or norm_bias is not None
    out = torch.empty_like(x)
    sum_only = residual is not None and not has_param
    res_out = torch.empty_like(x) if (residual is not None and has_param) else None
    stats = None if sum_only else torch.empty((2, rows), dtype=torch.float32, device=x.device)
    d, r = L.dims_array(x.shape)
    L.check(lib.pdb200_fused_bias_residual_layernorm(
        _ptr(x), d, r, _ptr(bias), _ptr(residual), _ptr(norm_weight), _ptr(norm_bias), ctypes.c_float(float(epsilon)),
        ctypes.c_float(float(residual_alpha)), int(begin_norm_axis), _dtype_code(x.dtype), _ptr(out), _ptr(res_out),
        _ptr(stats[0]) if stats is not None else None, _ptr(stats[1]) if stats is not None else None, _stream()), lib)
    if return_all:
        return out, res_out, (stats[0] if stats is not None else None), (stats[1] if stats is not None else None)
    return (out, res_out) if residual is not None and has_param else out


def softmax_(x, axis=-1, dtype=None, name=None):
    _require_cuda(x)
    lib = L.load()
    if not x.is_contiguous():
        raise ValueError("softmax_ needs a contiguous tensor")
    d, r = L.dims_array(x.shape)
    L.check(lib.pdb200_softmax(_ptr(x), d, r, int(axis), _dtype_code(x.dtype), _ptr(x), _stream()), lib)
    return x


_ln_cache = {}


def _ln_plan(ishape, normalized_shape, x_dtype, w, b):
    """Shape / dtype checks of LayerNormInferMeta and of gpu/layer_norm_kernel.cu:516-538, done once per
    (input shape, normalized_shape, dtypes) and remembered: -> (begin_norm_axis, rows, D, dims array, rank, param code)."""
    if isinstance(normalized_shape, numbers.Integral):
        normalized_shape = (normalized_shape,)
    key = (ishape, tuple(normalized_shape), x_dtype, None if w is None else (w.dtype, w.numel()),
           None if b is None else (b.dtype, b.numel()))
    plan = _ln_cache.get(key)
    if plan is not None:
        return plan
    normalized_shape = list(normalized_shape)
    ish = list(ishape)
    n = len(normalized_shape)
    if n == 0 or n > len(ish) or ish[len(ish) - n:] != normalized_shape:
        # norm.py:380-390
        raise ValueError(
            "Given normalized_shape is " + str(normalized_shape) + ", expected input with shape [*, " +
            str(normalized_shape)[1:] + ", but got input shape " + str(ish))
    begin_norm_axis = len(ish) - n
    D = 1
    for d in normalized_shape:
        D *= d
    numel = 1
    for d in ish:
        numel *= d
    rows = numel // D if D else 0
    pdt = None
    for p, nm in ((w, "Scale"), (b, "Bias")):
        if p is None:
            continue
        if p.numel() != D:
            raise ValueError(f"(InvalidArgument) {nm} must have {D} elements, got {p.numel()}")
        if pdt is None:
            pdt = p.dtype
        elif p.dtype != pdt:
            raise ValueError("(InvalidArgument) Scale and Bias must share one dtype")
    if pdt is not None and pdt != x_dtype and pdt != torch.float32:
        raise ValueError("(InvalidArgument) Unsupported data type of Scale and Bias")
    d, r = L.dims_array(ishape)
    pcode = _dtype_code(pdt) if pdt is not None else _dtype_code(x_dtype)
    plan = (begin_norm_axis, rows, D, d, r, pcode, _dtype_code(x_dtype))
    if len(_ln_cache) < 4096:
        _ln_cache[key] = plan
    return plan


def layer_norm(x, normalized_shape, weight=None, bias=None, epsilon=1e-05, name=None, return_stats=False):
    """layer_norm over the trailing `normalized_shape` dims (norm.py:320-396).
    return_stats=True also returns the intermediate outputs `mean` and `variance`
    (ops.yaml:2839-2851) that Paddle keeps for the backward pass."""
    _require_cuda(x, weight, bias)
    begin_norm_axis, rows, D, d, r, pcode, xcode = _ln_plan(tuple(x.shape), normalized_shape, x.dtype, weight, bias)
    lib = L.load()
    if not x.is_contiguous():
        x = x.contiguous()
    if weight is not None and not weight.is_contiguous():
        weight = weight.contiguous()
    if bias is not None and not bias.is_contiguous():
        bias = bias.contiguous()
    stat_dt = torch.float64 if x.dtype == torch.float64 else torch.float32
    y = torch.empty_like(x)
    stats = torch.empty((2, rows), dtype=stat_dt, device=x.device)  # mean and variance: one allocation
    p_mean = stats.data_ptr()
    p_var = p_mean + rows * stats.element_size()  # views are only built when the caller wants them (2 us each)
    L.check(lib.pdb200_layer_norm(_ptr(x), d, r, _ptr(weight), _ptr(bias), pcode, ctypes.c_float(float(epsilon)),
                                  begin_norm_axis, xcode, _ptr(y), ctypes.c_void_p(p_mean), ctypes.c_void_p(p_var), _stream()), lib)
    if return_stats:
        return y, stats[0], stats[1]
    return y


# ---------------------------------------------------------------------------------------------------
# backward kernels (what Paddle's autograd nodes call: _C_ops.softmax_grad / layer_norm_grad,
# paddle/phi/ops/yaml/backward.yaml)
# ---------------------------------------------------------------------------------------------------
def softmax_grad(out, out_grad, axis=-1):
    """x_grad of softmax given its output and the output's gradient."""
    _require_cuda(out, out_grad)
    if out.shape != out_grad.shape or out.dtype != out_grad.dtype:
        raise ValueError("(InvalidArgument) out and out_grad must have the same shape and dtype")
    lib = L.load()
    out, out_grad = out.contiguous(), out_grad.contiguous()
    x_grad = torch.empty_like(out)
    d, r = L.dims_array(out.shape)
    L.check(lib.pdb200_softmax_grad(_ptr(out), _ptr(out_grad), d, r, int(axis), _dtype_code(out.dtype), _ptr(x_grad),
                                    _stream()), lib)
    return x_grad


def log_softmax_grad(out, out_grad, axis=-1):
    """x_grad of log_softmax given its output and the output's gradient: out_grad - exp(out) * sum(out_grad, axis)."""
    _require_cuda(out, out_grad)
    if out.shape != out_grad.shape or out.dtype != out_grad.dtype:
        raise ValueError("(InvalidArgument) out and out_grad must have the same shape and dtype")
    lib = L.load()
    out, out_grad = out.contiguous(), out_grad.contiguous()
    x_grad = torch.empty_like(out)
    d, r = L.dims_array(out.shape)
    L.check(lib.pdb200_log_softmax_grad(_ptr(out), _ptr(out_grad), d, r, int(axis), _dtype_code(out.dtype), _ptr(x_grad),
                                        _stream()), lib)
    return x_grad


def layer_norm_grad(x, normalized_shape, weight, mean, variance, out_grad, epsilon=1e-05, need_weight_grad=None,
                    need_bias_grad=True):
    """(x_grad, weight_grad or None, bias_grad or None) of layer_norm.  mean / variance are the forward
    outputs (layer_norm(..., return_stats=True)).  weight_grad is produced when weight is given."""
    _require_cuda(x, weight, mean, variance, out_grad)
    if isinstance(normalized_shape, numbers.Integral):
        normalized_shape = [normalized_shape]
    n = len(list(normalized_shape))
    begin_norm_axis = x.dim() - n
    lib = L.load()
    x, out_grad = x.contiguous(), out_grad.contiguous()
    D = 1
    for dd in normalized_shape:
        D *= dd
    pdt = weight.dtype if weight is not None else x.dtype
    if need_weight_grad is None:
        need_weight_grad = weight is not None
    x_grad = torch.empty_like(x)
    wg = torch.empty((D,), dtype=pdt, device=x.device) if need_weight_grad else None
    bg = torch.empty((D,), dtype=pdt, device=x.device) if need_bias_grad else None
    d, r = L.dims_array(x.shape)
    nbytes = lib.pdb200_layer_norm_grad_workspace_bytes(d, r, begin_norm_axis, _dtype_code(x.dtype)) if (wg is not None or bg is not None) else 0
    ws = _workspace(nbytes, x.device)
    L.check(lib.pdb200_layer_norm_grad(_ptr(x), d, r, _ptr(weight.contiguous() if weight is not None else None),
                                       _dtype_code(pdt), _ptr(mean.contiguous()), _ptr(variance.contiguous()),
                                       _ptr(out_grad), ctypes.c_float(float(epsilon)), begin_norm_axis,
                                       _dtype_code(x.dtype), _ptr(x_grad), _ptr(wg), _ptr(bg), _ptr(ws), nbytes,
                                       _stream()), lib)
    return x_grad, wg, bg
