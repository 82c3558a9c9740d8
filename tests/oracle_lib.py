"""Test-side access to the CPU oracles (TEST INFRASTRUCTURE — never imported by the product).

  port : oracle/liboracle.so          plain-C restatement (oracle/phi_cpu_oracle.c), always available
  ref  : oracle/_ref/libphi_cpu_ref.so the reference's own CPU PHI kernels compiled from
         /root/reference (oracle/build_ref.sh); present in the build container and shipped
         to the GPU box as a prebuilt .so

Both expose the same numpy-level API through `Oracle`.  bf16/fp16 tensors travel as uint16 bit
patterns (numpy has no bfloat16).
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PORT_SO = os.path.join(ROOT, "oracle", "liboracle.so")
PORT_SRC = os.path.join(ROOT, "oracle", "phi_cpu_oracle.c")
PORT_SRCS = [PORT_SRC, os.path.join(ROOT, "oracle", "phi_cpu_oracle_family.c")]
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libphi_cpu_ref.so")

BOOL, UINT8, INT8, INT16, INT32, INT64, FLOAT16, BFLOAT16, FLOAT32, FLOAT64 = range(10)
NP2CODE = {np.dtype(np.bool_): BOOL, np.dtype(np.uint8): UINT8, np.dtype(np.int8): INT8,
           np.dtype(np.int16): INT16, np.dtype(np.int32): INT32, np.dtype(np.int64): INT64,
           np.dtype(np.float32): FLOAT32, np.dtype(np.float64): FLOAT64}
CODE2NP = {BOOL: np.bool_, UINT8: np.uint8, INT8: np.int8, INT16: np.int16, INT32: np.int32, INT64: np.int64,
           FLOAT16: np.uint16, BFLOAT16: np.uint16, FLOAT32: np.float32, FLOAT64: np.float64}
OPS = {"sum": 0, "mean": 1, "max": 2, "min": 3, "prod": 4}

_i64p = ctypes.POINTER(ctypes.c_int64)


def build_port(force=False):
    if force or not os.path.exists(PORT_SO) or os.path.getmtime(PORT_SO) < max(os.path.getmtime(p) for p in PORT_SRCS):
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-std=gnu11", "-fno-fast-math",
                               "-ffp-contract=off", "-o", PORT_SO, *PORT_SRCS, "-lm"])
    return PORT_SO


def _dims(shape):
    n = len(shape)
    return (ctypes.c_int64 * max(n, 1))(*[int(s) for s in shape]), n


def _vp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def reduced_shape(shape, axes, keepdim):
    rank = len(shape)
    if rank == 0:
        return ()
    red = set(a % rank for a in axes) if axes else set(range(rank))
    if keepdim:
        return tuple(1 if i in red else d for i, d in enumerate(shape))
    return tuple(d for i, d in enumerate(shape) if i not in red)


class Oracle:
    """kind = 'port' (C restatement) or 'ref' (compiled reference)."""

    def __init__(self, kind="port"):
        self.kind = kind
        if kind == "port":
            self.lib = ctypes.CDLL(build_port())
            self.p = "oracle_"
        else:
            if not os.path.exists(REF_SO):
                raise FileNotFoundError(REF_SO)
            self.lib = ctypes.CDLL(REF_SO)
            self.p = "phi_ref_"

    # ---- bf16 / fp16 -------------------------------------------------------------------
    def _conv(self, name, src, dst_dtype):
        src = np.ascontiguousarray(src)
        dst = np.empty(src.shape, dst_dtype)
        getattr(self.lib, self.p + name)(_vp(src), _vp(dst), ctypes.c_int64(src.size))
        return dst

    def f32_to_bf16(self, a):
        return self._conv("float_to_bf16", np.asarray(a, np.float32), np.uint16)

    def bf16_to_f32(self, a):
        return self._conv("bf16_to_float", np.asarray(a, np.uint16), np.float32)

    def f32_to_fp16(self, a):
        return self._conv("float_to_fp16", np.asarray(a, np.float32), np.uint16)

    def fp16_to_f32(self, a):
        return self._conv("fp16_to_float", np.asarray(a, np.uint16), np.float32)

    # ---- ops -----------------------------------------------------------------------------
    def elementwise(self, op, x, y, code=None):
        """op 'add'|'multiply'|'subtract'|'divide'|'maximum'|'minimum'.  code overrides the dtype
        (BFLOAT16/FLOAT16 on uint16 arrays)."""
        x = np.ascontiguousarray(x)
        y = np.ascontiguousarray(y)
        code = NP2CODE[x.dtype] if code is None else code
        oshape = np.broadcast_shapes(x.shape, y.shape)
        out = np.empty(oshape, x.dtype)
        xd, xr = _dims(x.shape)
        yd, yr = _dims(y.shape)
        opc = {"add": 0, "multiply": 1, "subtract": 2, "divide": 3, "maximum": 4, "minimum": 5}[op]
        if self.kind == "port":
            rc = self.lib.oracle_elementwise(opc, code, _vp(x), xd, xr, _vp(y), yd, yr, _vp(out))
        else:
            od, orank = _dims(oshape)
            rc = self.lib.phi_ref_elementwise(opc, code, _vp(x), xd, xr, _vp(y), yd, yr, od, orank, _vp(out))
        self._check(rc)
        return out

    def reduce(self, op, x, axis=None, keepdim=False):
        x = np.ascontiguousarray(x)
        axes = [] if axis is None else ([int(a) for a in axis] if isinstance(axis, (list, tuple)) else [int(axis)])
        oshape = reduced_shape(x.shape, axes, keepdim)
        out = np.empty(oshape, x.dtype)
        xd, xr = _dims(x.shape)
        ad, an = _dims(axes)
        code = NP2CODE[x.dtype]
        if self.kind == "port":
            rc = self.lib.oracle_reduce(OPS[op], code, _vp(x), xd, xr, ad, an, int(not axes), _vp(out))
        else:
            od, orank = _dims(oshape)
            rc = self.lib.phi_ref_reduce(OPS[op], code, _vp(x), xd, xr, ad, an, int(keepdim), int(not axes),
                                         od, orank, _vp(out))
        self._check(rc)
        return out

    def argmin(self, x, axis=None, keepdim=False, dtype=np.int64, code=None):
        return self.argmax(x, axis, keepdim, dtype, code, _fn="argmin")

    def argmax(self, x, axis=None, keepdim=False, dtype=np.int64, code=None, _fn="argmax"):
        x = np.ascontiguousarray(x)
        code = NP2CODE[x.dtype] if code is None else code
        flatten = axis is None
        rank = x.ndim
        if flatten or rank == 0:
            oshape = tuple([1] * rank) if (keepdim and flatten) else ()
        else:
            a = axis % rank
            oshape = tuple(1 if i == a else d for i, d in enumerate(x.shape)) if keepdim else \
                tuple(d for i, d in enumerate(x.shape) if i != a)
        out = np.empty(oshape, dtype)
        xd, xr = _dims(x.shape)
        ocode = NP2CODE[np.dtype(dtype)]
        ax = 0 if flatten else int(axis)
        if self.kind == "port":
            rc = getattr(self.lib, "oracle_" + _fn)(code, _vp(x), xd, xr, ctypes.c_int64(ax), int(flatten), ocode, _vp(out))
        else:
            od, orank = _dims(oshape)
            rc = getattr(self.lib, "phi_ref_" + _fn)(code, _vp(x), xd, xr, ctypes.c_int64(ax), int(keepdim), int(flatten),
                                                     ocode, od, orank, _vp(out))
        self._check(rc)
        return out

    def softmax(self, x, axis=-1):
        x = np.ascontiguousarray(x)
        out = np.empty_like(x)
        xd, xr = _dims(x.shape)
        rc = getattr(self.lib, self.p + "softmax")(NP2CODE[x.dtype], _vp(x), xd, xr, int(axis), _vp(out))
        self._check(rc)
        return out

    def layer_norm(self, x, begin_norm_axis, scale=None, bias=None, epsilon=1e-5):
        x = np.ascontiguousarray(x)
        y = np.empty_like(x)
        rows = int(np.prod(x.shape[:begin_norm_axis], dtype=np.int64))
        mean = np.empty((rows,), x.dtype)
        var = np.empty((rows,), x.dtype)
        xd, xr = _dims(x.shape)
        s = None if scale is None else np.ascontiguousarray(scale, x.dtype)
        b = None if bias is None else np.ascontiguousarray(bias, x.dtype)
        rc = getattr(self.lib, self.p + "layer_norm")(
            NP2CODE[x.dtype], _vp(x), xd, xr, _vp(s) if s is not None else None,
            _vp(b) if b is not None else None, ctypes.c_float(epsilon), int(begin_norm_axis), _vp(y), _vp(mean),
            _vp(var))
        self._check(rc)
        return y, mean, var

    def softmax_grad(self, out, out_grad, axis=-1):
        out = np.ascontiguousarray(out)
        out_grad = np.ascontiguousarray(out_grad, out.dtype)
        dx = np.empty_like(out)
        d, r = _dims(out.shape)
        rc = getattr(self.lib, self.p + "softmax_grad")(NP2CODE[out.dtype], _vp(out), _vp(out_grad), d, r, int(axis), _vp(dx))
        self._check(rc)
        return dx

    def log_softmax_grad(self, out, out_grad, axis=-1):
        out = np.ascontiguousarray(out)
        out_grad = np.ascontiguousarray(out_grad, out.dtype)
        dx = np.empty_like(out)
        d, r = _dims(out.shape)
        rc = getattr(self.lib, self.p + "log_softmax_grad")(NP2CODE[out.dtype], _vp(out), _vp(out_grad), d, r, int(axis), _vp(dx))
        self._check(rc)
        return dx

    def layer_norm_grad(self, x, begin_norm_axis, scale, mean, var, out_grad, epsilon=1e-5, want_bias_grad=True):
        """Returns (dx, dscale or None, dbias or None).  dscale is produced exactly when scale is given."""
        x = np.ascontiguousarray(x)
        g = np.ascontiguousarray(out_grad, x.dtype)
        D = int(np.prod(x.shape[begin_norm_axis:], dtype=np.int64))
        dx = np.empty_like(x)
        s = None if scale is None else np.ascontiguousarray(scale, x.dtype)
        ds = None if scale is None else np.empty((D,), x.dtype)
        db = np.empty((D,), x.dtype) if want_bias_grad else None
        m = np.ascontiguousarray(mean, x.dtype)
        v = np.ascontiguousarray(var, x.dtype)
        d, r = _dims(x.shape)
        rc = getattr(self.lib, self.p + "layer_norm_grad")(
            NP2CODE[x.dtype], _vp(x), d, r, _vp(s) if s is not None else None, _vp(m), _vp(v), _vp(g),
            ctypes.c_float(epsilon), int(begin_norm_axis), _vp(dx), _vp(ds) if ds is not None else None,
            _vp(db) if db is not None else None)
        self._check(rc)
        return dx, ds, db

    # ---- the rest of the elementwise family (oracle/phi_cpu_oracle_family.c, oracle/ref_shim_family.cc) ----
    def _binary_coded(self, fam, opc, x, y, code, out_dtype):
        x = np.ascontiguousarray(x)
        y = np.ascontiguousarray(y)
        code = NP2CODE[x.dtype] if code is None else code
        oshape = np.broadcast_shapes(x.shape, y.shape)
        out = np.empty(oshape, out_dtype or x.dtype)
        xd, xr = _dims(x.shape)
        yd, yr = _dims(y.shape)
        if self.kind == "port":
            rc = getattr(self.lib, "oracle_" + fam)(opc, code, _vp(x), xd, xr, _vp(y), yd, yr, _vp(out))
        else:
            od, orank = _dims(oshape)
            rc = getattr(self.lib, "phi_ref_" + fam)(opc, code, _vp(x), xd, xr, _vp(y), yd, yr, od, orank, _vp(out))
        self._check(rc)
        return out

    def compare(self, op, x, y, code=None):
        """op 'less_than'|'less_equal'|'greater_than'|'greater_equal'|'equal'|'not_equal' -> bool array."""
        opc = ["less_than", "less_equal", "greater_than", "greater_equal", "equal", "not_equal"].index(op)
        return self._binary_coded("compare", opc, x, y, code, np.bool_)

    def logical(self, op, x, y, code=None):
        return self._binary_coded("logical", ["and", "or", "xor"].index(op), x, y, code, np.bool_)

    def bitwise(self, op, x, y):
        return self._binary_coded("bitwise", ["and", "or", "xor"].index(op), x, y, None, None)

    def bitwise_shift(self, right, x, y, is_arithmetic=True):
        x = np.ascontiguousarray(x)
        y = np.ascontiguousarray(y)
        oshape = np.broadcast_shapes(x.shape, y.shape)
        out = np.empty(oshape, x.dtype)
        xd, xr = _dims(x.shape)
        yd, yr = _dims(y.shape)
        if self.kind == "port":
            rc = self.lib.oracle_bitwise_shift(int(right), int(is_arithmetic), NP2CODE[x.dtype], _vp(x), xd, xr, _vp(y), yd, yr, _vp(out))
        else:
            od, orank = _dims(oshape)
            rc = self.lib.phi_ref_bitwise_shift(int(right), int(is_arithmetic), NP2CODE[x.dtype], _vp(x), xd, xr, _vp(y), yd, yr,
                                                od, orank, _vp(out))
        self._check(rc)
        return out

    def logical_not(self, x, code=None):
        x = np.ascontiguousarray(x)
        code = NP2CODE[x.dtype] if code is None else code
        out = np.empty(x.shape, np.bool_)
        self._check(getattr(self.lib, self.p + "not")(0, code, _vp(x), ctypes.c_int64(x.size), _vp(out)))
        return out

    def bitwise_not(self, x):
        x = np.ascontiguousarray(x)
        out = np.empty(x.shape, x.dtype)
        self._check(getattr(self.lib, self.p + "not")(1, NP2CODE[x.dtype], _vp(x), ctypes.c_int64(x.size), _vp(out)))
        return out

    def scale(self, x, scale=1.0, bias=0.0, bias_after_scale=True, code=None):
        x = np.ascontiguousarray(x)
        code = NP2CODE[x.dtype] if code is None else code
        out = np.empty(x.shape, x.dtype)
        self._check(getattr(self.lib, self.p + "scale")(code, _vp(x), ctypes.c_int64(x.size), ctypes.c_double(scale),
                                                        ctypes.c_double(bias), int(bool(bias_after_scale)), _vp(out)))
        return out

    def cast(self, x, out_code, code=None):
        """out_code: dtype code of the result; bf16/fp16 results come back as uint16 bit patterns."""
        x = np.ascontiguousarray(x)
        code = NP2CODE[x.dtype] if code is None else code
        out = np.empty(x.shape, CODE2NP[out_code])
        self._check(getattr(self.lib, self.p + "cast")(code, out_code, _vp(x), ctypes.c_int64(x.size), _vp(out)))
        return out

    def where(self, cond, x, y, code=None):
        cond = np.ascontiguousarray(cond, np.bool_)
        x = np.ascontiguousarray(x)
        y = np.ascontiguousarray(y)
        code = NP2CODE[x.dtype] if code is None else code
        out = np.empty(x.shape, x.dtype)
        self._check(getattr(self.lib, self.p + "where")(code, _vp(cond), _vp(x), _vp(y), ctypes.c_int64(x.size), _vp(out)))
        return out

    def elementwise_grad(self, op, x, y, dout, want_dx=True, want_dy=True, code=None, gpu_arithmetic=False):
        """op 'add'|'multiply' -> (dx, dy) (None where not wanted).  gpu_arithmetic (port only): sums of 16-bit types
        in float, rounded once, as the reference's GPU kernels do (the CPU kernels round after every addition)."""
        x, y, dout = (np.ascontiguousarray(a) for a in (x, y, dout))
        code = NP2CODE[x.dtype] if code is None else code
        dx = np.empty(x.shape, x.dtype) if want_dx else None
        dy = np.empty(y.shape, y.dtype) if want_dy else None
        xd, xr = _dims(x.shape)
        yd, yr = _dims(y.shape)
        opc = {"add": 0, "multiply": 1}[op]
        pdx = _vp(dx) if dx is not None else None
        pdy = _vp(dy) if dy is not None else None
        if self.kind == "port":
            rc = self.lib.oracle_elementwise_grad(opc, code, int(gpu_arithmetic), _vp(x), xd, xr, _vp(y), yd, yr,
                                                  _vp(dout), pdx, pdy)
        else:
            od, orank = _dims(dout.shape)
            rc = self.lib.phi_ref_elementwise_grad(opc, code, _vp(x), xd, xr, _vp(y), yd, yr, _vp(dout), od, orank,
                                                   pdx, pdy)
        self._check(rc)
        return dx, dy

    def reduce_grad(self, op, x_shape, out_grad, axis=None, keepdim=False, code=None):
        """op 'sum'|'mean' -> x_grad of shape x_shape; out_grad has the forward output's shape."""
        out_grad = np.ascontiguousarray(out_grad)
        code = NP2CODE[out_grad.dtype] if code is None else code
        rank = len(x_shape)
        axes = [] if axis is None else ([int(a) for a in axis] if isinstance(axis, (list, tuple)) else [int(axis)])
        x_grad = np.empty(x_shape, out_grad.dtype)
        xd, xr = _dims(x_shape)
        opc = {"sum": 0, "mean": 1}[op]
        if self.kind == "port":
            red = [1] * rank if not axes else [int(i in {a % rank for a in axes}) for i in range(rank)]
            rc = self.lib.oracle_reduce_grad(opc, code, xd, xr, (ctypes.c_int * max(rank, 1))(*red), _vp(out_grad), _vp(x_grad))
        else:
            x = np.zeros(x_shape, out_grad.dtype)
            gd, gr = _dims(out_grad.shape if out_grad.shape == reduced_shape(x_shape, axes, keepdim) else reduced_shape(x_shape, axes, keepdim))
            ad, an = _dims(axes)
            rc = self.lib.phi_ref_reduce_grad(opc, code, _vp(x), xd, xr, _vp(out_grad), gd, gr, ad, an, int(keepdim),
                                              int(not axes), _vp(x_grad))
        self._check(rc)
        return x_grad

    def select_grad(self, op, x, out, out_grad, axis=None, keepdim=False, code=None, tol=True):
        """op 'max'|'min'|'prod' -> x_grad.  port: tol=True is the GPU kernel's EqualFunctor (|a - b| < 1e-8), tol=False the
        CPU kernel's exact ==; ref: the compiled CPU kernel (exact ==; float32 / float64 / int32 / int64)."""
        x, out, out_grad = (np.ascontiguousarray(t) for t in (x, out, out_grad))
        code = NP2CODE[x.dtype] if code is None else code
        rank = x.ndim
        axes = [] if axis is None else ([int(a) for a in axis] if isinstance(axis, (list, tuple)) else [int(axis)])
        x_grad = np.empty(x.shape, x.dtype)
        xd, xr = _dims(x.shape)
        if self.kind == "port":
            red = [1] * rank if not axes else [int(i in {a % rank for a in axes}) for i in range(rank)]
            rc = self.lib.oracle_select_grad(int(op == "prod"), int(tol), code, xd, xr, (ctypes.c_int * max(rank, 1))(*red),
                                             _vp(x), _vp(out), _vp(out_grad), _vp(x_grad))
        else:
            gd, gr = _dims(reduced_shape(x.shape, axes, keepdim))
            ad, an = _dims(axes)
            rc = self.lib.phi_ref_select_grad({"max": 0, "min": 1, "prod": 2}[op], code, _vp(x), xd, xr, _vp(out), _vp(out_grad),
                                              gd, gr, ad, an, int(keepdim), int(not axes), _vp(x_grad))
        self._check(rc)
        return x_grad

    def log_softmax(self, x, axis=-1, code=None):
        x = np.ascontiguousarray(x)
        code = NP2CODE[x.dtype] if code is None else code
        out = np.empty(x.shape, x.dtype)
        d, r = _dims(x.shape)
        self._check(getattr(self.lib, self.p + "log_softmax")(code, _vp(x), d, r, int(axis), _vp(out)))
        return out

    def cross_entropy_with_softmax(self, logits, label, soft_label=False, use_softmax=True, ignore_index=-100, axis=-1):
        """-> (softmax, loss); hard labels are int64 with a 1 along `axis`."""
        logits = np.ascontiguousarray(logits)
        label = np.ascontiguousarray(label if soft_label else np.asarray(label, np.int64))
        code = NP2CODE[logits.dtype]
        ax = axis % logits.ndim
        lshape = tuple(1 if i == ax else d for i, d in enumerate(logits.shape))
        sm = np.empty(logits.shape, logits.dtype)
        loss = np.empty(lshape, logits.dtype)
        d, r = _dims(logits.shape)
        if self.kind == "port":
            rc = self.lib.oracle_cross_entropy_with_softmax(code, _vp(logits), d, r, _vp(label), int(soft_label), int(use_softmax),
                                                            int(ignore_index), int(axis), _vp(sm), _vp(loss))
        else:
            ld, _ = _dims(label.shape)
            rc = self.lib.phi_ref_cross_entropy_with_softmax(code, _vp(logits), d, r, _vp(label), ld, int(soft_label),
                                                             int(use_softmax), int(ignore_index), int(axis), _vp(sm), _vp(loss))
        self._check(rc)
        return sm, loss

    def cross_entropy_with_softmax_grad(self, label, softmax, loss_grad, soft_label=False, ignore_index=-100, axis=-1):
        """-> logits_grad (use_softmax = True); hard labels are int64 with a 1 along `axis`."""
        softmax = np.ascontiguousarray(softmax)
        label = np.ascontiguousarray(label if soft_label else np.asarray(label, np.int64))
        loss_grad = np.ascontiguousarray(loss_grad, softmax.dtype)
        code = NP2CODE[softmax.dtype]
        dx = np.empty_like(softmax)
        d, r = _dims(softmax.shape)
        if self.kind == "port":
            rc = self.lib.oracle_cross_entropy_with_softmax_grad(code, _vp(label), _vp(softmax), _vp(loss_grad), d, r, int(soft_label),
                                                                 int(ignore_index), int(axis), _vp(dx))
        else:
            ld, _ = _dims(label.shape)
            rc = self.lib.phi_ref_cross_entropy_with_softmax_grad(code, _vp(label), ld, _vp(softmax), _vp(loss_grad), d, r,
                                                                  int(soft_label), 1, int(ignore_index), int(axis), _vp(dx))
        self._check(rc)
        return dx

    def rms_norm(self, x, gamma, beta=None, eps=1e-6, begin_norm_axis=1, bias=None, residual=None):
        """float64 restatement of the reference test's formula (port only: the reference has no CPU rms_norm).
        -> (out, residual_out or None, inv_var), all float64."""
        assert self.kind == "port", "rms_norm has no compiled-reference oracle"
        x = np.ascontiguousarray(x, np.float64)
        D = int(np.prod(x.shape[begin_norm_axis:], dtype=np.int64))
        rows = x.size // D
        f = lambda a: None if a is None else np.ascontiguousarray(a, np.float64)
        gamma, beta, bias, residual = f(gamma), f(beta), f(bias), f(residual)
        out = np.empty_like(x)
        res_out = np.empty_like(x) if residual is not None else None
        inv = np.empty(x.shape[:begin_norm_axis], np.float64)
        p = lambda a: _vp(a) if a is not None else None
        rc = self.lib.oracle_rms_norm(_vp(x), ctypes.c_int64(rows), ctypes.c_int64(D), p(bias), p(residual), _vp(gamma), p(beta),
                                      ctypes.c_double(eps), _vp(out), p(res_out), _vp(inv))
        self._check(rc)
        return out, res_out, inv

    # ---- fused neighbours whose reference kernels are GPU-only: pinned THROUGH the reference CPU kernels on
    # pre-combined inputs --------------------------------------------------------------------
    def softmax_mask_fuse(self, x, mask):
        """softmax(float(x) + float(mask)) along the last axis; x [B,H,Sq,Sk], mask [B,1,Sq,Sk], float32 arrays.  The
        sum is formed here (one float addition per element, as the GPU kernel does); the softmax is the CPU kernel,
        whose -64 clip differs from the reference GPU kernel's plain exp only below exp(-64) = 1.6e-28."""
        z = (np.asarray(x, np.float32) + np.asarray(mask, np.float32)).astype(np.float32)
        return self.softmax(z, -1)

    def softmax_mask_fuse_upper_triangle(self, x):
        """Causal softmax of float32 x [B,H,S,S]: row q = CPU softmax of its first q+1 elements, exact zeros after."""
        x = np.asarray(x, np.float32)
        B, H, S, S2 = x.shape
        assert S == S2
        out = np.zeros_like(x)
        for q in range(S):
            out[:, :, q, :q + 1] = self.softmax(np.ascontiguousarray(x[:, :, q, :q + 1]).reshape(B * H, q + 1), -1).reshape(B, H, q + 1)
        return out

    def fused_bias_dropout_residual_layer_norm(self, x, residual, bias, ln_scale, ln_bias, eps=1e-5, factor=1.0, half=None,
                                               conv=None, alpha=1.0, sum_only=False):
        """Deterministic cases of layer_norm(residual + dropout(x + bias)) (keep = 1): d is formed here with the
        reference's T arithmetic — every step rounded to T (half = 'bf16' | 'fp16' | None for float32; x / residual /
        bias then arrive as uint16 bit patterns) — and the layer_norm is the reference CPU kernel on float(d).
        `conv`: the Oracle whose float <-> 16-bit conversions are used (default: this one).  Pass the restatement when this
        is the compiled reference and half == 'fp16': the reference's CPU float -> float16 TRUNCATES on builds without
        F16C (phi/common/float16.h:121-138) while the GPU rounds to nearest even.
        alpha / sum_only: the op fused_bias_residual_layernorm (fused_layernorm_kernel.cu) built on the same helper —
        the residual is multiplied by T(alpha) first; sum_only returns d alone (the form without norm parameters).
        -> (y float32, d as T, mean, variance)."""
        conv = conv or self
        if half is None:
            up = lambda a: np.asarray(a, np.float32)
            rt = lambda a: a.astype(np.float32)
            down = lambda a: a
        else:
            to_h, to_f = (conv.f32_to_bf16, conv.bf16_to_f32) if half == "bf16" else (conv.f32_to_fp16, conv.fp16_to_f32)
            up = to_f
            rt = lambda a: to_f(to_h(a.astype(np.float32)))
            down = lambda a: to_h(a.astype(np.float32))
        t = up(x)
        if bias is not None:
            t = rt(t + up(bias))
        if factor != 1.0:
            t = rt(t * rt(np.float32(factor).reshape(1))[0])
        r = up(residual)
        if alpha != 1.0:  # fused_bias_residual_layernorm: residual * T(residual_alpha), rounded to T before the sum
            r = rt(r * rt(np.float32(alpha).reshape(1))[0])
        d = rt(t + r)
        if sum_only:
            return down(d)
        D = d.shape[-1]
        y, mean, var = self.layer_norm(d.reshape(-1, D), 1, ln_scale, ln_bias, eps)
        return y.reshape(d.shape), down(d), mean, var

    def have_avx512(self):
        return self.kind == "ref" and hasattr(self.lib, "phi_ref_have_avx512") and bool(self.lib.phi_ref_have_avx512())

    def fused_layer_norm_cpu(self, x, bias, residual, norm_weight, norm_bias, eps=1e-5, alpha=1.0):
        """The reference's own CPU kernel of fused_bias_residual_layernorm (fusion/cpu/fused_layer_norm_avx_kernel.cc, float32,
        AVX-512 hosts only; kind == 'ref').  It forms x + alpha * residual + bias — the GPU kernel (x + bias) + alpha * residual —
        and returns 1 / sqrt(var + eps) where the GPU kernel returns the variance: comparisons against the GPU path are
        to rounding, not to the bit.  -> (out, residual_out | None, mean | None, rstd | None)."""
        assert self.kind == "ref"
        x = np.ascontiguousarray(x, np.float32)
        rows, cols = x.shape
        f = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
        bias, residual, norm_weight, norm_bias = f(bias), f(residual), f(norm_weight), f(norm_bias)
        has_param = norm_weight is not None or norm_bias is not None
        out = np.empty_like(x)
        ro = np.empty_like(x) if residual is not None else None
        m = np.empty(rows, np.float32) if has_param else None
        rs = np.empty(rows, np.float32) if has_param else None
        P = lambda a: ctypes.c_void_p(0) if a is None else _vp(a)
        fn = self.lib.phi_ref_fused_layer_norm
        fn.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64] + [ctypes.c_void_p] * 4 + [ctypes.c_float, ctypes.c_float] + \
                      [ctypes.c_void_p] * 4
        self._check(fn(_vp(x), rows, cols, P(bias), P(residual), P(norm_weight), P(norm_bias), float(eps), float(alpha), P(out), P(ro),
                       P(m), P(rs)))
        if residual is not None and not has_param:
            ro = None  # the sum-only form writes `out` alone
        return out, ro, m, rs

    def _check(self, rc):
        if rc != 0:
            msg = ""
            if self.kind == "ref":
                self.lib.phi_ref_last_error.restype = ctypes.c_char_p
                msg = self.lib.phi_ref_last_error().decode("utf-8", "replace")[:400]
            raise ValueError(f"oracle[{self.kind}] rc={rc} {msg}")


def have_ref():
    return os.path.exists(REF_SO)
