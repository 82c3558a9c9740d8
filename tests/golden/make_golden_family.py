"""Generate tests/golden/phi_family_golden.npz from the reference's OWN CPU kernels (the rest of the
elementwise family: compare / logical / bitwise / not / scale / cast / where — SURVEY.md §8f rank 1).

Run in the build container, where /root/reference exists and oracle/build_ref.sh has produced
oracle/_ref/libphi_cpu_ref.so:

    python tests/golden/make_golden_family.py

Every `<case>__out` is an output of that library on the seeded inputs `<case>__in*`.  The shapes follow the
reference's own op tests: test/legacy_test/test_compare_op.py (broadcast cases :560-640), test_logical_op.py
(TEST_META_SHAPE_DATA :60-110), test_bitwise_op.py, test_scale_op.py, test_cast_op.py, test_where_op.py.
CASES / FLOATS are imported by the tests so that inputs never have to be stored twice.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as O  # noqa: E402

BCAST = [("same", (13, 17), (13, 17)), ("vec_last", (2, 10, 12), (12,)), ("four", (10, 3, 1, 4), (10, 1, 12, 1)),
         ("x_small", (10, 12), (2, 2, 10, 12)), ("col", (64, 33), (64, 1)), ("scalar0d", (3, 4), ()), ("both0d", (), ())]
COMPARE_OPS = ["less_than", "less_equal", "greater_than", "greater_equal", "equal", "not_equal"]
INT_TYPES = ["bool", "uint8", "int8", "int16", "int32", "int64"]
SPECIAL = [0.0, -0.0, 1.0, 1.0 + 1e-9, 1e-9, -1e-9, 2.5, -3.0, np.inf, -np.inf, np.nan]


def draw(rng, dt, shape, special=True):
    """Small value sets so that ties, zeros and (for floats) inf / NaN / near-equal pairs all occur."""
    if dt == "bool":
        return rng.random(shape) < 0.5
    if dt in ("uint8", "int8", "int16", "int32", "int64"):
        lo = 0 if dt == "uint8" else -3
        return rng.integers(lo, 4, shape).astype(dt)
    vals = np.array(SPECIAL if special else [0.0, 1.0, 2.5, -3.0, 0.5], dt)
    return rng.choice(vals, shape)


def main():
    ref = O.Oracle("ref")
    rng = np.random.default_rng(20261107)
    g = {}
    for name, xs, ys in BCAST:
        for dt in INT_TYPES + ["float32", "float64"]:
            x, y = draw(rng, dt, xs), draw(rng, dt, ys)
            g[f"fam_{name}_{dt}__in0"], g[f"fam_{name}_{dt}__in1"] = x, y
            for op in COMPARE_OPS:
                g[f"cmp_{op}_{name}_{dt}__out"] = ref.compare(op, x, y)
            if dt != "uint8":
                for op in ("and", "or", "xor"):
                    g[f"log_{op}_{name}_{dt}__out"] = ref.logical(op, x, y)
            if dt in INT_TYPES:
                for op in ("and", "or", "xor"):
                    g[f"bit_{op}_{name}_{dt}__out"] = ref.bitwise(op, x, y)
        for half, conv, code in (("bfloat16", ref.f32_to_bf16, O.BFLOAT16), ("float16", ref.f32_to_fp16, O.FLOAT16)):
            x = conv(draw(rng, "float32", xs).astype(np.float32))
            y = conv(draw(rng, "float32", ys).astype(np.float32))
            g[f"fam_{name}_{half}__in0"], g[f"fam_{name}_{half}__in1"] = x, y
            for op in COMPARE_OPS:
                g[f"cmp_{op}_{name}_{half}__out"] = ref.compare(op, x, y, code=code)
    # one-input maps
    for dt in INT_TYPES + ["float32", "float64"]:
        x = draw(rng, dt, (7, 19))
        g[f"un_{dt}__in0"] = x
        if dt != "uint8":
            g[f"lnot_{dt}__out"] = ref.logical_not(x)
        if dt in INT_TYPES:
            g[f"bnot_{dt}__out"] = ref.bitwise_not(x)
    # scale: test_scale_op.py (scale -2.3, bias 0 / 2.3 / after and before)
    for dt in ["uint8", "int8", "int16", "int32", "int64", "float32", "float64", "bool"]:
        if dt == "bool":
            x = rng.random((10, 10)) < 0.5
        elif dt in ("float32", "float64"):
            x = rng.random((10, 10)).astype(dt)
        else:
            x = rng.integers(0, 120, (10, 10)).astype(dt)
        g[f"scale_{dt}__in0"] = x
        for k, (s, b, after) in enumerate(SCALE_ARGS):
            g[f"scale_{dt}_{k}__out"] = ref.scale(x, s, b, after)
    xb = ref.f32_to_bf16(rng.random((10, 10)).astype(np.float32))
    g["scale_bfloat16__in0"] = xb
    for k, (s, b, after) in enumerate(SCALE_ARGS):
        g[f"scale_bfloat16_{k}__out"] = ref.scale(xb, s, b, after, code=O.BFLOAT16)
    # cast: every ordered pair of the ten dtypes (float sources stay inside every target's range)
    for sc in range(10):
        x = cast_input(rng, ref, sc)
        g[f"cast_{sc}__in0"] = x
        for dc in range(10):
            if dc == O.FLOAT16 and sc != O.FLOAT16:
                continue  # the CPU build's float -> fp16 truncates (no F16C); the GPU rounds: not a golden value
            g[f"cast_{sc}_{dc}__out"] = ref.cast(x, dc, code=sc)
    # where: test_where_op.py (uniform x, y; random condition)
    for dt in ("int32", "int64", "float32", "float64"):
        c = rng.random((6, 5, 20)) < 0.5
        x = (rng.standard_normal((6, 5, 20)) * 9).astype(dt)
        y = (rng.standard_normal((6, 5, 20)) * 9).astype(dt)
        g[f"where_{dt}__in0"], g[f"where_{dt}__in1"], g[f"where_{dt}__in2"] = c, x, y
        g[f"where_{dt}__out"] = ref.where(c, x, y)
    # add_grad / multiply_grad: the broadcast matrix of test_elementwise_add_op.py / test_elementwise_mul_op.py
    for name, xs, ys in GRAD_CASES:
        os_ = np.broadcast_shapes(xs, ys)
        for dt in ("float32", "float64", "int32", "int64"):
            if dt.startswith("int"):
                x, y, d = (rng.integers(-9, 9, s_).astype(dt) for s_ in (xs, ys, os_))
            else:
                x, y, d = (rng.standard_normal(s_).astype(dt) for s_ in (xs, ys, os_))
            g[f"grad_{name}_{dt}__in0"], g[f"grad_{name}_{dt}__in1"], g[f"grad_{name}_{dt}__in2"] = x, y, d
            for op in ("add", "multiply"):
                dx, dy = ref.elementwise_grad(op, x, y, d)
                g[f"grad_{op}_{name}_{dt}__out0"], g[f"grad_{op}_{name}_{dt}__out1"] = dx, dy
        x, y, d = (ref.f32_to_bf16(rng.standard_normal(s_).astype(np.float32)) for s_ in (xs, ys, os_))
        g[f"grad_{name}_bfloat16__in0"], g[f"grad_{name}_bfloat16__in1"], g[f"grad_{name}_bfloat16__in2"] = x, y, d
        dx, dy = ref.elementwise_grad("multiply", x, y, d, code=O.BFLOAT16)  # add_grad has no bf16 kernel on CPU
        g[f"grad_multiply_{name}_bfloat16__out0"], g[f"grad_multiply_{name}_bfloat16__out1"] = dx, dy
    # sum_grad / mean_grad: the axis matrix of test_reduce_op.py
    for name, shape, axis, keep in REDUCE_GRAD_CASES:
        axes = [] if axis is None else ([axis] if isinstance(axis, int) else list(axis))
        gshape = O.reduced_shape(shape, axes, keep)
        for dt in ("float32", "float64", "int32", "int64"):
            d = (rng.standard_normal(gshape) * 100).astype(dt)
            g[f"rgrad_{name}_{dt}__in0"] = d
            for op in ("sum", "mean"):
                g[f"rgrad_{op}_{name}_{dt}__out"] = ref.reduce_grad(op, shape, d, axis, keep)
    # log_softmax: test_log_softmax_op.py shapes + the softmax matrix of make_golden.py
    for name, shape, axis in LOG_SOFTMAX_CASES:
        for dt in ("float32", "float64"):
            x = (rng.standard_normal(shape) * 3).astype(dt)
            if name == "ls_clip":
                x[..., 0] = 200.0  # everything else sits below max - 64: the clip is active
            g[f"lsm_{name}_{dt}__in0"] = x
            g[f"lsm_{name}_{dt}__out"] = ref.log_softmax(x, axis)
    # cross_entropy_with_softmax: test_softmax_with_cross_entropy_op.py (hard / soft labels, axis, ignore_index)
    for name, shape, axis in CE_CASES:
        ax = axis % len(shape)
        lshape = tuple(1 if i == ax else d for i, d in enumerate(shape))
        for dt in ("float32", "float64"):
            x = (rng.standard_normal(shape) * 3).astype(dt)
            hard = rng.integers(0, shape[ax], lshape)
            hard.flat[0] = -100
            soft = rng.random(shape).astype(dt)
            soft /= soft.sum(axis=ax, keepdims=True)
            g[f"ce_{name}_{dt}__in0"], g[f"ce_{name}_{dt}__in1"], g[f"ce_{name}_{dt}__in2"] = x, hard, soft
            g[f"ce_hard_{name}_{dt}__out0"], g[f"ce_hard_{name}_{dt}__out1"] = ref.cross_entropy_with_softmax(x, hard, False, True, -100, axis)
            g[f"ce_soft_{name}_{dt}__out0"], g[f"ce_soft_{name}_{dt}__out1"] = ref.cross_entropy_with_softmax(x, soft, True, True, -100, axis)
    # shifts: test_bitwise_op.py (TestBitwiseLeftShift / RightShift, arithmetic and logic), counts in and out of range
    for name, xs, ys in SHIFT_CASES:
        for dt in ("uint8", "int8", "int16", "int32", "int64"):
            info = np.iinfo(dt)
            x = rng.integers(info.min, info.max, xs, dtype=dt, endpoint=True)
            y = rng.integers(0 if dt == "uint8" else -3, info.bits + 3, ys).astype(dt)
            g[f"sh_{name}_{dt}__in0"], g[f"sh_{name}_{dt}__in1"] = x, y
            for right in (0, 1):
                for ar in (0, 1):
                    g[f"sh_{name}_{dt}_{right}{ar}__out"] = ref.bitwise_shift(right, x, y, bool(ar))
    out = os.path.join(HERE, "phi_family_golden.npz")
    np.savez_compressed(out, **g)
    print(f"wrote {out}: {len(g)} arrays, {os.path.getsize(out)} bytes")


GRAD_CASES = [("same", (13, 17), (13, 17)), ("vec_last", (2, 10, 12), (12,)), ("b_10_12", (2, 10, 12), (10, 12)),
              ("mid", (2, 3, 100), (3, 1)), ("lead", (100, 2, 3), (100, 1, 1)), ("four", (10, 3, 1, 4), (10, 1, 12, 1)),
              ("x_small", (10, 12), (2, 2, 10, 12)), ("scalar0d", (3, 4), ()), ("both0d", (), ()), ("col", (64, 33), (64, 1)),
              ("outer_x", (1, 33), (64, 33)), ("rows", (300, 256), (256,))]
REDUCE_GRAD_CASES = [("r2_last", (5, 6), -1, False), ("r2_first_keep", (5, 6), 0, True), ("r2_all", (5, 6), None, False),
                     ("r3_mid_keep", (5, 6, 10), 1, True), ("r3_02", (5, 6, 10), (0, 2), False), ("r4_12", (2, 5, 3, 2), (1, 2), False),
                     ("r1", (120,), 0, False), ("r3_one", (3, 1, 4), 1, False), ("r0", (), None, False),
                     ("r5", (1, 2, 5, 6, 10), 2, False), ("r2_wide", (3, 901), -1, False)]
LOG_SOFTMAX_CASES = [("ls2", (2, 3), -1), ("ls4_m1", (2, 3, 4, 5), -1), ("ls4_0", (2, 3, 4, 5), 0), ("ls4_1", (2, 3, 4, 5), 1),
                     ("ls4_2", (2, 3, 4, 5), 2), ("ls_2048", (6, 2048), -1), ("ls_odd", (3, 1001), -1), ("ls_clip", (4, 50), -1)]
CE_CASES = [("c2", (5, 7), -1), ("c3_1", (4, 6, 9), 1), ("c4_2", (3, 4, 5, 6), 2), ("c_2048", (8, 2048), -1), ("c3_0", (2, 3, 4), 0),
            ("c4_m1", (2, 3, 4, 41), -1)]
SHIFT_CASES = [("same", (13, 17), (13, 17)), ("vec_last", (2, 10, 12), (12,)), ("x_small", (10, 12), (2, 2, 10, 12)),
               ("four", (10, 3, 1, 4), (10, 1, 12, 1)), ("scalar0d", (3, 4), ())]
SCALE_ARGS = [(-2.3, 0.0, True), (-2.3, 2.3, True), (-2.3, 2.3, False), (3.0, -1.0, True), (0.0, 0.5, False)]


def cast_input(rng, oracle, code):
    """Values every target dtype can hold (non-negative and < 100), so no conversion is undefined."""
    if code == O.BOOL:
        return rng.random((9, 31)) < 0.5
    if code in (O.UINT8, O.INT8, O.INT16, O.INT32, O.INT64):
        lo = 0 if code == O.UINT8 else -100
        return rng.integers(lo, 100, (9, 31)).astype(O.CODE2NP[code])
    f = (rng.random((9, 31)) * 99).astype(np.float32)
    f[0, :4] = [0.0, 0.5, 1.0, 98.75]
    if code == O.FLOAT16:
        return oracle.f32_to_fp16(f)
    if code == O.BFLOAT16:
        return oracle.f32_to_bf16(f)
    return f.astype(O.CODE2NP[code])


if __name__ == "__main__":
    main()
