"""Child process of tests/test_gpu_phi_dispatch.py (TEST INFRASTRUCTURE): loads Paddle's registry host and the plugin,
calls every registered kernel through SelectKernelOrThrowError -> GetVariadicKernelFn and compares with the ctypes path.
usage: phi_dispatch_child.py <libphi_registry_min.so> <libpdb200_phi.so>; prints one JSON line."""
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch  # noqa: E402

reg = ctypes.CDLL(sys.argv[1], mode=ctypes.RTLD_GLOBAL)
ctypes.CDLL(sys.argv[2], mode=ctypes.RTLD_GLOBAL)  # PD_REGISTER_KERNEL static initialisers run here

import paddlepaddle_b200 as paddle  # noqa: E402

F = paddle.nn.functional

# phi::DataType (paddle/phi/common/data_type.h:44-62)
PHI = {torch.bool: 1, torch.uint8: 2, torch.int8: 3, torch.int16: 5, torch.int32: 7, torch.int64: 9, torch.float32: 10,
       torch.float64: 11, torch.float16: 15, torch.bfloat16: 16}
UNDEFINED = 0


class PmTensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("dims", ctypes.c_int64 * 9), ("rank", ctypes.c_int), ("dtype", ctypes.c_int)]


reg.phi_min_last_error.restype = ctypes.c_char_p
reg.phi_min_alloc_calls.restype = ctypes.c_size_t
reg.phi_min_run.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p, ctypes.POINTER(PmTensor), ctypes.c_int,
                            ctypes.POINTER(PmTensor), ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.c_int,
                            ctypes.POINTER(ctypes.c_int64), ctypes.c_int]
assert reg.phi_min_gpu_init(0) == 0, reg.phi_min_last_error()

ABSENT = object()   # an optional input that is not given / an output that is not wanted
called, failed, n_calls = set(), [], 0


def _pm(t):
    p = PmTensor()
    if t is ABSENT or t is None:
        p.data, p.rank, p.dtype = None, -1, UNDEFINED
        return p
    p.data = t.data_ptr() if t.numel() else None
    p.rank = t.dim()
    for i, d in enumerate(t.shape):
        p.dims[i] = d
    p.dtype = PHI[t.dtype]
    return p


def run(name, family, ins, out_like, f=(), iv=(), backend="GPU", key=None):
    """Call kernel `name` through Paddle's dispatch.  ins: tensors (ABSENT for a missing optional); out_like: tensors
    giving each output's InferMeta result (shape + dtype) — filled with the kernel's results and returned."""
    global n_calls
    key = key if key is not None else PHI[ins[0].dtype]
    ins_c = (PmTensor * max(1, len(ins)))(*[_pm(t.contiguous() if isinstance(t, torch.Tensor) else t) for t in ins])
    outs = [ABSENT if o is ABSENT else torch.full_like(o, 0) for o in out_like]
    outs_c = (PmTensor * max(1, len(outs)))(*[_pm(o) for o in outs])
    for i, o in enumerate(outs):  # zero-size outputs still need their rank (>= 0 means "wanted")
        if o is not ABSENT:
            outs_c[i].rank = o.dim()
    fa = (ctypes.c_double * max(1, len(f)))(*f)
    ia = (ctypes.c_int64 * max(1, len(iv)))(*[int(v) for v in iv])
    torch.cuda.synchronize()
    rc = reg.phi_min_run(name.encode(), backend.encode(), key, family.encode(), ins_c, len(ins), outs_c, len(outs), fa, len(f), ia,
                         len(iv))
    n_calls += 1
    called.add(name)
    if rc != 0:
        raise RuntimeError(f"{name}: {reg.phi_min_last_error().decode()[:300]}")
    return outs


def same(a, b):
    if a.dtype != b.dtype or a.shape != b.shape:
        return False
    if a.dtype in (torch.bfloat16, torch.float16):
        return torch.equal(a.view(torch.int16), b.view(torch.int16))
    if a.is_floating_point():
        return torch.equal(a.view(torch.int32 if a.dtype == torch.float32 else torch.int64),
                           b.view(torch.int32 if b.dtype == torch.float32 else torch.int64))
    return torch.equal(a, b)


def check(label, got, want):
    got = got if isinstance(got, (list, tuple)) else [got]
    want = want if isinstance(want, (list, tuple)) else [want]
    for i, (g, w) in enumerate(zip(got, want)):
        if g is ABSENT or w is None:
            continue
        if not same(g.reshape(w.shape), w):
            failed.append(f"{label} output {i}: differs from the ctypes path")


def attempt(label, fn):
    try:
        fn()
    except Exception as exc:  # noqa: BLE001
        failed.append(f"{label}: {type(exc).__name__}: {str(exc)[:300]}")


g = torch.Generator(device="cuda")
g.manual_seed(7)


def rnd(shape, dt=torch.float32, scale=1.0):
    if dt in (torch.int32, torch.int64, torch.int16, torch.int8, torch.uint8):
        info = torch.iinfo(dt)
        return torch.randint(max(info.min, -1000), min(info.max, 1000), shape, device="cuda", dtype=dt, generator=g)
    if dt == torch.bool:
        return torch.rand(shape, device="cuda", generator=g) < 0.5
    return (torch.randn(shape, device="cuda", generator=g) * scale).to(dt)


# ---- elementwise families -------------------------------------------------------------------------------------------
def case_binary():
    for dt in (torch.bfloat16, torch.int32, torch.float32):
        x, y = rnd((64, 256), dt), rnd((256,), dt)
        if dt != torch.int32:
            y = y + 3  # keep divisors away from zero
        else:
            y = torch.where(y == 0, torch.ones_like(y), y)
        for name, fn in (("add", paddle.add), ("multiply", paddle.multiply), ("subtract", paddle.subtract), ("divide", paddle.divide),
                         ("maximum", paddle.maximum), ("minimum", paddle.minimum)):
            want = fn(x, y)
            check(f"{name}/{dt}", run(name, "binary", [x, y], [want]), want)
        for name, fn in (("less_than", paddle.less_than), ("less_equal", paddle.less_equal), ("greater_than", paddle.greater_than),
                         ("greater_equal", paddle.greater_equal), ("equal", paddle.equal), ("not_equal", paddle.not_equal),
                         ("logical_and", paddle.logical_and), ("logical_or", paddle.logical_or), ("logical_xor", paddle.logical_xor)):
            want = fn(x, y)
            check(f"{name}/{dt}", run(name, "binary", [x, y], [want]), want)
    xi, yi = rnd((64, 256), torch.int32), rnd((256,), torch.int32)
    for name, fn in (("bitwise_and", paddle.bitwise_and), ("bitwise_or", paddle.bitwise_or), ("bitwise_xor", paddle.bitwise_xor)):
        want = fn(xi, yi)
        check(name, run(name, "binary", [xi, yi], [want]), want)
    sh = torch.randint(0, 40, (256,), device="cuda", dtype=torch.int32, generator=g)
    for name, fn in (("bitwise_left_shift", paddle.bitwise_left_shift), ("bitwise_right_shift", paddle.bitwise_right_shift)):
        for arith in (1, 0):
            want = fn(xi, sh, bool(arith))
            check(name, run(name, "binary_bool", [xi, sh], [want], iv=[arith]), want)
    want = paddle.logical_not(xi)
    check("logical_not", run("logical_not", "unary", [xi], [want]), want)
    want = paddle.bitwise_not(xi)
    check("bitwise_not", run("bitwise_not", "unary", [xi], [want]), want)
    xb = rnd((1000, 33), torch.bfloat16)
    want = paddle.scale(xb, 2.5, 1.0)
    check("scale", run("scale", "scale", [xb], [want], f=[2.5, 1.0], iv=[1]), want)
    xf = rnd((1000, 33), torch.float32, 50.0)
    for odt, nm in ((torch.bfloat16, "bfloat16"), (torch.int32, "int32")):
        want = paddle.cast(xf, nm)
        check(f"cast->{nm}", run("cast", "cast", [xf], [want], iv=[PHI[odt]]), want)
    c, a, b = rnd((4096,), torch.bool), rnd((4096,), torch.float32), rnd((4096,), torch.float32)
    want = paddle.where(c, a, b)
    check("where", run("where", "ternary", [c, a, b], [want], key=PHI[torch.float32]), want)


def case_grads():
    for dt in (torch.float32, torch.bfloat16):
        x, y, d = rnd((600, 1024), dt), rnd((1024,), dt), rnd((600, 1024), dt)
        wx, wy = paddle.add_grad(x, y, d)
        check(f"add_grad/{dt}", run("add_grad", "binary_grad", [x, y, d], [wx, wy], iv=[-1]), [wx, wy])
        # autograd does not need dx: the kernel receives a NULL DenseTensor* for it
        got = run("add_grad", "binary_grad", [x, y, d], [ABSENT, wy], iv=[-1])
        check(f"add_grad(dx=NULL)/{dt}", got, [None, wy])
        wx, wy = paddle.multiply_grad(x, y, d)
        check(f"multiply_grad/{dt}", run("multiply_grad", "binary_grad", [x, y, d], [wx, wy], iv=[-1]), [wx, wy])
    x = rnd((64, 96, 32), torch.float32)
    og = rnd((64, 32), torch.float32)
    for name, fn in (("sum_grad", paddle.sum_grad), ("mean_grad", paddle.mean_grad)):
        want = fn(x, og, axis=1)
        check(name, run(name, "reduce_grad", [x, og], [want], iv=[0, 0, 1]), want)
    for dt in (torch.float32, torch.bfloat16):
        xs = torch.randint(-4, 5, (64, 96, 32), device="cuda", generator=g).to(dt)
        ogs = rnd((64, 32), dt)
        for name, fwd, fn in (("max_grad", paddle.max, paddle.max_grad), ("min_grad", paddle.min, paddle.min_grad)):
            out = fwd(xs, axis=1)
            want = fn(xs, out, ogs, axis=1)
            check(f"{name}/{dt}", run(name, "select_grad", [xs, out, ogs], [want], iv=[0, 0, 1]), want)
        xp = rnd((64, 96, 32), dt) + 3
        out = rnd((64, 32), dt)
        want = paddle.prod_grad(xp, out, ogs, axis=1)
        check(f"prod_grad/{dt}", run("prod_grad", "select_grad", [xp, out, ogs], [want], iv=[0, 0, 1]), want)


# ---- reductions -------------------------------------------------------------------------------------------------------
def case_reductions():
    xf = rnd((300, 4096), torch.float32)
    xb = rnd((300, 4096), torch.bfloat16)
    xi = rnd((300, 4096), torch.int32)
    x8 = rnd((300, 4096), torch.int8)
    for x in (xf, xb):
        for axes in ([1], [0], [0, 1]):
            want = paddle.sum(x, axis=axes)
            # sum: (x, IntArray axis, DataType out_dtype = UNDEFINED, keep_dim): the output dtype comes from InferMeta's meta
            check(f"sum/{x.dtype}/{axes}", run("sum", "sum", [x], [want], iv=[UNDEFINED, 0] + axes), want)
            check(f"sum_raw/{x.dtype}/{axes}", run("sum_raw", "sum_raw", [x], [want], iv=[0, 0, UNDEFINED] + axes), want)
            for name, fn in (("mean", paddle.mean), ("max", paddle.max), ("min", paddle.min)):
                want = fn(x, axis=axes)
                check(f"{name}/{x.dtype}/{axes}", run(name, "reduce", [x], [want], iv=[0] + axes), want)
                check(f"{name}_raw/{x.dtype}/{axes}", run(name + "_raw", "reduce_raw", [x], [want], iv=[0, 0] + axes), want)
    want = paddle.sum(xi)  # int32 -> int64 by InferMeta; reduce_all
    check("sum int32->int64", run("sum_raw", "sum_raw", [xi], [want], iv=[0, 1, UNDEFINED]), want)
    want = paddle.sum(x8, axis=0)  # the small integers keep their dtype and wrap
    check("sum int8", run("sum", "sum", [x8], [want], iv=[UNDEFINED, 0, 0]), want)
    # out_dtype outside the built pairs: the binding casts first, like phi::Reduce
    want = paddle.sum(paddle.cast(xf, "float64"), axis=1)
    check("sum fp32->fp64", run("sum", "sum", [xf], [want], iv=[PHI[torch.float64], 0, 1]), want)
    small = torch.randint(-2, 3, (64, 40), device="cuda", dtype=torch.int64, generator=g)
    want = paddle.prod(small, axis=1)
    check("prod", run("prod", "reduce_raw", [small], [want], iv=[0, 0, 1]), want)
    for x, axis, flatten in ((xf, 1, 0), (xb, 0, 0), (xi, 0, 1)):
        want = paddle.argmax(x, axis=None if flatten else axis)
        check(f"argmax/{x.dtype}", run("argmax", "argmax", [x], [want], iv=[axis, 0, flatten, PHI[torch.int64]]), want)
    want = paddle.argmin(xf, axis=1)
    check("argmin", run("argmin", "argmax", [xf], [want], iv=[1, 0, 0, PHI[torch.int64]]), want)
    want = paddle.argmax(xf, axis=1, dtype="int32")
    check("argmax int32 out", run("argmax", "argmax", [xf], [want], iv=[1, 0, 0, PHI[torch.int32]]), want)


# ---- row ops ------------------------------------------------------------------------------------------------------------
def case_rows():
    for dt in (torch.float32, torch.bfloat16):
        x = rnd((700, 2048), dt, 2.0)
        want = F.softmax(x, axis=-1)
        for backend in ("GPUDNN", "GPU"):  # the generated API asks for GPUDNN first (kernel_factory.cc:304-314)
            check(f"softmax/{backend}/{dt}", run("softmax", "axis", [x], [want], iv=[-1], backend=backend), want)
        wl = F.log_softmax(x, axis=-1)
        check(f"log_softmax/{dt}", run("log_softmax", "axis", [x], [wl], iv=[-1]), wl)
        dy = rnd((700, 2048), dt)
        wg = F.softmax_grad(want, dy, axis=-1)
        for backend in ("GPUDNN", "GPU"):
            check(f"softmax_grad/{backend}/{dt}", run("softmax_grad", "axis_grad", [want, dy], [wg], iv=[-1], backend=backend), wg)
        wlg = F.log_softmax_grad(wl, dy, axis=-1)
        check(f"log_softmax_grad/{dt}", run("log_softmax_grad", "axis_grad", [wl, dy], [wlg], iv=[-1]), wlg)
    # layer_norm: bfloat16 key, float32 scale / bias (the mixed-dtype rule), mean / variance outputs are float32
    rows, D = 700, 8192
    x = rnd((rows, D), torch.bfloat16, 3.0)
    w, b = rnd((D,), torch.float32) * 0.1 + 1, rnd((D,), torch.float32) * 0.1
    y, m, v = F.layer_norm(x, [D], w, b, 1e-5, return_stats=True)
    check("layer_norm", run("layer_norm", "layer_norm", [x, w, b], [y, m, v], f=[1e-5], iv=[1]), [y, m, v])
    y0, m0, v0 = F.layer_norm(x, [D], None, None, 1e-5, return_stats=True)
    check("layer_norm no affine", run("layer_norm", "layer_norm", [x, ABSENT, ABSENT], [y0, m0, v0], f=[1e-5], iv=[1]), [y0, m0, v0])
    dy = rnd((rows, D), torch.bfloat16)
    gx, gs, gb = F.layer_norm_grad(x, [D], w, m, v, dy, 1e-5)
    check("layer_norm_grad", run("layer_norm_grad", "layer_norm_grad", [x, w, b, m, v, dy], [gx, gs, gb], f=[1e-5], iv=[1]), [gx, gs, gb])
    # cross_entropy_with_softmax: hard labels
    logits = rnd((512, 4096), torch.float32, 2.0)
    lbl = torch.randint(0, 4096, (512, 1), device="cuda", generator=g)
    loss, sm = F.softmax_with_cross_entropy(logits, lbl, return_softmax=True)
    check("cross_entropy_with_softmax", run("cross_entropy_with_softmax", "cross_entropy", [logits, lbl], [sm, loss],
                                            iv=[0, 1, 1, -100, -1]), [sm, loss])
    lg = rnd((512, 1), torch.float32)
    wdx = F.softmax_with_cross_entropy_grad(lbl, sm, lg)
    check("cross_entropy_with_softmax_grad", run("cross_entropy_with_softmax_grad", "cross_entropy_grad", [lbl, sm, lg], [wdx],
                                                 iv=[0, 1, -100, -1], key=PHI[torch.float32]), wdx)  # key dtype: softmax's
    # rms_norm with the fused residual + bias add
    xr, rr = rnd((700, 4096), torch.bfloat16), rnd((700, 4096), torch.bfloat16)
    wr, br, nb = rnd((4096,), torch.bfloat16) * 0.1 + 1, rnd((4096,), torch.bfloat16), rnd((4096,), torch.bfloat16)
    o, ro, iv_ = F.fused_rms_norm(xr, wr, nb, 1e-6, 1, bias=br, residual=rr)
    check("rms_norm", run("rms_norm", "rms_norm", [xr, br, rr, wr, nb], [o, ro, iv_], f=[1e-6], iv=[1]), [o, ro, iv_])
    o2, iv2 = F.fused_rms_norm(xr, wr, None, 1e-6, 1)
    check("rms_norm plain", run("rms_norm", "rms_norm", [xr, ABSENT, ABSENT, wr, ABSENT], [o2, ABSENT, iv2], f=[1e-6], iv=[1]),
          [o2, None, iv2])
    # masked softmax
    xs = rnd((2, 4, 256, 256), torch.bfloat16, 2.0)
    mk = rnd((2, 1, 256, 256), torch.bfloat16)
    want = F.softmax_mask_fuse(xs, mk)
    check("fused_softmax_mask", run("fused_softmax_mask", "softmax_mask", [xs, mk], [want]), want)
    want = F.softmax_mask_fuse_upper_triangle(xs)
    check("fused_softmax_mask_upper_triangle", run("fused_softmax_mask_upper_triangle", "unary", [xs], [want]), want)
    # layer_norm(residual + dropout(x + bias)), inference form
    xf, rf, bf = rnd((700, 4096), torch.float16), rnd((700, 4096), torch.float16), rnd((4096,), torch.float16)
    ws, wb = rnd((4096,), torch.float32) * 0.1 + 1, rnd((4096,), torch.float32) * 0.1
    y, dro, m, v = F.fused_bias_dropout_residual_layer_norm(xf, rf, bf, ws, wb, dropout_rate=0.3, training=False,
                                                            mode="downscale_in_infer", return_all=True)
    mask_like = torch.empty(xf.shape, dtype=torch.uint8, device="cuda")
    got = run("fused_bias_dropout_residual_layer_norm", "fused_ln", [xf, rf, bf, ws, wb], [y, dro, mask_like, m, v], f=[0.3, 1e-5],
              iv=[1, 0])
    check("fused_bias_dropout_residual_layer_norm", [got[0], got[1], got[3], got[4]], [y, dro, m, v])
    # fused_bias_residual_layernorm (paddle.incubate.nn.functional.fused_layer_norm): the layer_norm form with a residual,
    # the sum-only form (no norm parameters: only `out` is allocated) and the plain form (no residual)
    nw, nb = rnd((4096,), torch.float32) * 0.1 + 1, rnd((4096,), torch.float32) * 0.1
    o, ro, m, v = F.fused_layer_norm(xf, nw, nb, 1e-5, residual_alpha=0.5, begin_norm_axis=1, bias=bf, residual=rf, return_all=True)
    got = run("fused_bias_residual_layernorm", "fused_layernorm", [xf, bf, rf, nw, nb], [o, ro, m, v], f=[1e-5, 0.5], iv=[1])
    check("fused_bias_residual_layernorm/ln+residual", got, [o, ro, m, v])
    o = F.fused_layer_norm(xf, None, None, 1e-5, residual_alpha=0.5, begin_norm_axis=1, bias=bf, residual=rf)
    got = run("fused_bias_residual_layernorm", "fused_layernorm", [xf, bf, rf, ABSENT, ABSENT], [o, ABSENT, ABSENT, ABSENT], f=[1e-5, 0.5], iv=[1])
    check("fused_bias_residual_layernorm/sum only", [got[0]], [o])
    o, _, m, v = F.fused_layer_norm(xf, nw, nb, 1e-5, begin_norm_axis=1, return_all=True)
    got = run("fused_bias_residual_layernorm", "fused_layernorm", [xf, ABSENT, ABSENT, nw, nb], [o, ABSENT, m, v], f=[1e-5, 1.0], iv=[1])
    check("fused_bias_residual_layernorm/plain", [got[0], got[2], got[3]], [o, m, v])


for label, fn in (("elementwise", case_binary), ("grads", case_grads), ("reductions", case_reductions), ("rows", case_rows)):
    attempt(label, fn)

# every kernel name the plugin registered must have been called at least once
reg.phi_min_has_kernel.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]
NAMES = ["add", "multiply", "subtract", "divide", "maximum", "minimum", "less_than", "less_equal", "greater_than", "greater_equal",
         "equal", "not_equal", "logical_and", "logical_or", "logical_xor", "logical_not", "bitwise_and", "bitwise_or", "bitwise_xor",
         "bitwise_not", "bitwise_left_shift", "bitwise_right_shift", "scale", "cast", "where", "add_grad", "multiply_grad", "sum_grad",
         "mean_grad", "max_grad", "min_grad", "prod_grad", "cross_entropy_with_softmax", "cross_entropy_with_softmax_grad", "rms_norm", "log_softmax", "sum_raw", "sum", "mean_raw", "mean", "max_raw", "max",
         "min_raw", "min", "prod", "argmax", "argmin", "softmax", "layer_norm", "softmax_grad", "log_softmax_grad", "layer_norm_grad", "fused_softmax_mask",
         "fused_softmax_mask_upper_triangle", "fused_bias_dropout_residual_layer_norm", "fused_bias_residual_layernorm"]
print(json.dumps({"calls": n_calls, "alloc_calls": reg.phi_min_alloc_calls(), "failed": failed,
                  "names_not_called": sorted(set(NAMES) - called)}))
