"""Generate tests/golden/phi_cpu_golden.npz from the reference's OWN CPU kernels.

Run in the build container, where /root/reference exists and oracle/build_ref.sh has produced
oracle/_ref/libphi_cpu_ref.so (Paddle's CPU PHI kernels compiled in place):

    python tests/golden/make_golden.py

Every array `<case>__out*` is an output of that library on the seeded inputs `<case>__in*`;
tests/test_oracle_pin.py replays the inputs through the C restatement (oracle/phi_cpu_oracle.c)
and, on the GPU box, tests/test_gpu_parity.py replays them through the CUDA kernels.
The shape/axis matrices follow the reference's own op tests:
  test/legacy_test/test_reduce_op.py:28-231,268-604,1137+   test_elementwise_add_op.py:318-615
  test_arg_min_max_op.py:30-190   test_softmax_op.py:29-35   test_layer_norm_op.py:35-52
and the docstring known answers of python/paddle/tensor/math.py:1530-1585.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as O  # noqa: E402

REDUCE_CASES = [  # (name, shape, axis, keepdim)
    ("r2_last", (5, 6), -1, False), ("r2_first", (5, 6), 0, False), ("r2_all", (5, 6), None, False),
    ("r3_mid_keep", (5, 6, 10), 1, True), ("r3_02", (5, 6, 10), (0, 2), False),
    ("r4_12", (2, 5, 3, 2), (1, 2), False), ("r5", (1, 2, 5, 6, 10), 2, False),
    ("r6", (1, 1, 2, 5, 6, 10), (3, 5), False), ("r8", (1, 3, 1, 2, 1, 4, 3, 10), (0, 3), False),
    ("r1", (120,), 0, False), ("r1_long", (10007,), None, False), ("r2_wide", (3, 9001), -1, False),
]
BCAST_CASES = [  # from test_elementwise_add_op.py:318-615
    ("same", (13, 17), (13, 17)), ("vec_last", (2, 10, 12), (12,)), ("b_10_12", (2, 10, 12), (10, 12)),
    ("mid", (2, 3, 100), (3, 1)), ("lead", (100, 2, 3), (100, 1, 1)), ("four", (10, 3, 1, 4), (10, 1, 12, 1)),
    ("x_small", (10, 12), (2, 2, 10, 12)), ("scalar0d", (3, 4), ()), ("both0d", (), ()),
    ("one", (2, 5, 3, 12), (2, 5, 1, 12)), ("col", (64, 33), (64, 1)), ("outer_x", (1, 33), (64, 33)),
]


SIBLING_CASES = [("same", (13, 17), (13, 17)), ("vec_last", (2, 10, 12), (12,)), ("four", (10, 3, 1, 4), (10, 1, 12, 1)),
                 ("x_small", (10, 12), (2, 2, 10, 12)), ("col", (64, 33), (64, 1)), ("scalar0d", (3, 4), ())]
SIBLING_OPS = {  # GPU registrations, kps/elementwise_kernel.cu:183-202,338-381
    "float32": ("subtract", "divide", "maximum", "minimum"), "float64": ("subtract", "divide", "maximum", "minimum"),
    "int32": ("subtract", "divide", "maximum", "minimum"), "int64": ("subtract", "divide", "maximum", "minimum"),
    "int16": ("subtract", "divide"), "bfloat16": ("subtract", "divide", "maximum", "minimum"),
}


GRAD_SOFTMAX_CASES = [("g2", (10, 10), -1), ("g4_m1", (2, 3, 4, 5), -1), ("g4_1", (2, 3, 4, 5), 1), ("g4_0", (2, 3, 4, 5), 0),
                      ("g_2048", (6, 2048), -1), ("g_odd", (3, 1001), -1)]
GRAD_LN_CASES = [("g4", (2, 6, 6, 3), 1), ("g4_3", (2, 6, 6, 3), 3), ("g2_8192", (2, 8192), 1), ("g2_odd", (3, 1001), 1),
                 ("g3", (2, 5, 768), 2)]


def main():
    ref = O.Oracle("ref")
    rng = np.random.default_rng(20261106)
    g = {}
    # --- docstring known answers, python/paddle/tensor/math.py:1530-1585 (paddle.sum) ---
    x = np.array([[0.2, 0.3, 0.5, 0.9], [0.1, 0.2, 0.6, 0.7]], np.float32)
    g["doc_sum__in0"] = x
    g["doc_sum_all__out"] = ref.reduce("sum", x)          # 3.5
    g["doc_sum_0__out"] = ref.reduce("sum", x, 0)         # [0.3, 0.5, 1.1, 1.6]
    g["doc_sum_m1__out"] = ref.reduce("sum", x, -1)       # [1.9, 1.6]
    g["doc_sum_m1_keep__out"] = ref.reduce("sum", x, -1, True)
    # --- reductions ---
    for name, shape, axis, keep in REDUCE_CASES:
        for dt in (np.float32, np.float64, np.int32, np.int64):
            if np.issubdtype(dt, np.floating):
                xin = rng.uniform(0.5, 1.5, shape).astype(dt)  # prod stays finite
            else:
                xin = rng.integers(-1000, 1000, shape).astype(dt)
            key = f"red_{name}_{np.dtype(dt).name}"
            g[key + "__in0"] = xin
            for op in ("sum", "mean", "max", "min", "prod"):
                g[f"{key}_{op}__out"] = ref.reduce(op, xin, axis, keep)
    big = rng.integers(-2**40, 2**40, (7, 1000)).astype(np.int64)  # int64 wrap on prod
    g["red_wrap_int64__in0"] = big
    g["red_wrap_int64_prod__out"] = ref.reduce("prod", big, 1)
    g["red_wrap_int64_sum__out"] = ref.reduce("sum", big)
    # --- elementwise ---
    for name, xs, ys in BCAST_CASES:
        for dt in (np.float32, np.int32, np.int64, np.float64, np.int16, np.uint8, np.int8, np.bool_):
            if dt is np.bool_:
                a, b = rng.random(xs) > 0.5, rng.random(ys) > 0.5
            elif np.issubdtype(dt, np.floating):
                a, b = rng.standard_normal(xs).astype(dt), rng.standard_normal(ys).astype(dt)
            else:
                info = np.iinfo(dt)
                a = rng.integers(info.min, info.max, xs, dtype=dt, endpoint=True)
                b = rng.integers(info.min, info.max, ys, dtype=dt, endpoint=True)
            key = f"ew_{name}_{np.dtype(dt).name}"
            g[key + "__in0"], g[key + "__in1"] = a, b
            g[key + "_add__out"] = ref.elementwise("add", a, b)
            g[key + "_multiply__out"] = ref.elementwise("multiply", a, b)
        a = ref.f32_to_bf16(rng.standard_normal(xs).astype(np.float32))
        b = ref.f32_to_bf16(rng.standard_normal(ys).astype(np.float32))
        key = f"ew_{name}_bfloat16"
        g[key + "__in0"], g[key + "__in1"] = a, b
        g[key + "_add__out"] = ref.elementwise("add", a, b, O.BFLOAT16)
        g[key + "_multiply__out"] = ref.elementwise("multiply", a, b, O.BFLOAT16)
    f = np.concatenate([rng.standard_normal(4096).astype(np.float32) * 1e3,
                        np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 1e-40, 3.389e38, 1.00390625, 1.01171875], np.float32)])
    g["bf16_round__in0"] = f
    g["bf16_round__out"] = ref.f32_to_bf16(f)
    # --- argmax ---
    for name, shape, axis in [("a3_0", (3, 4, 5), 0), ("a3_1", (3, 4, 5), 1), ("a3_m1", (3, 4, 5), -1),
                              ("a_flat", (3, 4, 5), None), ("a1", (1000,), 0), ("a2_wide", (4, 5000), 1)]:
        for dt in (np.float32, np.float64, np.int32, np.int64, np.int16, np.uint8):
            xin = rng.integers(0, 50, shape).astype(dt)  # many ties: lowest index must win
            key = f"arg_{name}_{np.dtype(dt).name}"
            g[key + "__in0"] = xin
            g[key + "_i64__out"] = ref.argmax(xin, axis, False, np.int64)
            g[key + "_i32_keep__out"] = ref.argmax(xin, axis, True, np.int32)
    # --- softmax ---
    for name, shape, axis in [("s2", (10, 10), -1), ("s4_m1", (2, 3, 4, 5), -1), ("s4_0", (2, 3, 4, 5), 0),
                              ("s4_1", (2, 3, 4, 5), 1), ("s4_2", (2, 3, 4, 5), 2), ("s5", (2, 3, 4, 5, 6), 3),
                              ("s_2048", (6, 2048), -1), ("s_odd", (3, 1001), -1), ("s_clip", (2, 64), -1)]:
        for dt in (np.float32, np.float64):
            xin = (rng.standard_normal(shape) * 2).astype(dt)
            if name == "s_clip":
                xin[0, 3] = -1e4  # x - max < -64: the CPU kernel clips (softmax_impl.h:36-42)
                xin[1, :] = -200.0
                xin[1, 7] = 0.0
            key = f"sm_{name}_{np.dtype(dt).name}"
            g[key + "__in0"] = xin
            g[key + "__out"] = ref.softmax(xin, axis)
    # --- layer_norm ---
    for name, shape, bna in [("l4", (2, 6, 6, 3), 1), ("l4_3", (2, 6, 6, 3), 3), ("l2_8192", (3, 8192), 1),
                             ("l2_odd", (5, 1001), 1), ("l3", (4, 10, 768), 2)]:
        for dt in (np.float32, np.float64):
            xin = (rng.standard_normal(shape) * 3 + 1).astype(dt)
            D = int(np.prod(shape[bna:]))
            w, b = rng.uniform(0.5, 1.5, D).astype(dt), rng.uniform(-0.5, 0.5, D).astype(dt)
            key = f"ln_{name}_{np.dtype(dt).name}"
            g[key + "__in0"], g[key + "__in1"], g[key + "__in2"] = xin, w, b
            y, m, v = ref.layer_norm(xin, bna, w, b, 1e-5)
            g[key + "_wb__out0"], g[key + "_wb__out1"], g[key + "_wb__out2"] = y, m, v
            y, m, v = ref.layer_norm(xin, bna, None, None, 1e-5)
            g[key + "_none__out0"] = y
    # --- subtract / divide / maximum / minimum, appended last so that every array above
    #     keeps the value it had before these were added ---
    for name, xs, ys in SIBLING_CASES:
        for dt in ("float32", "float64", "int32", "int64", "int16", "bfloat16"):
            if dt == "bfloat16":
                a = ref.f32_to_bf16(rng.standard_normal(xs).astype(np.float32))
                b = ref.f32_to_bf16((rng.standard_normal(ys) + 3).astype(np.float32))
                code = O.BFLOAT16
            elif dt.startswith("float"):
                a, b, code = rng.standard_normal(xs).astype(dt), (rng.standard_normal(ys) + 3).astype(dt), None
            else:
                info = np.iinfo(dt)
                a = rng.integers(info.min, info.max, xs, dtype=dt, endpoint=True)
                b = rng.integers(info.min, info.max, ys, dtype=dt, endpoint=True)
                b[b == 0] = 1
                code = None
            key = f"ew2_{name}_{dt}"
            g[key + "__in0"], g[key + "__in1"] = a, b
            for op in SIBLING_OPS[dt]:
                g[f"{key}_{op}__out"] = ref.elementwise(op, a, b, code)
    # --- backward kernels, appended after everything above ---
    for name, shape, axis in GRAD_SOFTMAX_CASES:
        for dt in (np.float32, np.float64):
            y = ref.softmax((rng.standard_normal(shape) * 2).astype(dt), axis)
            dy = rng.standard_normal(shape).astype(dt)
            key = f"smg_{name}_{np.dtype(dt).name}"
            g[key + "__in0"], g[key + "__in1"] = y, dy
            g[key + "__out"] = ref.softmax_grad(y, dy, axis)
    for name, shape, bna in GRAD_LN_CASES:
        for dt in (np.float32, np.float64):
            xin = (rng.standard_normal(shape) * 3 + 1).astype(dt)
            D = int(np.prod(shape[bna:]))
            w = rng.uniform(0.5, 1.5, D).astype(dt)
            _, m, v = ref.layer_norm(xin, bna, w, None, 1e-5)
            dy = rng.standard_normal(shape).astype(dt)
            key = f"lng_{name}_{np.dtype(dt).name}"
            g[key + "__in0"], g[key + "__in1"], g[key + "__in2"], g[key + "__in3"], g[key + "__in4"] = xin, w, m, v, dy
            dx, ds, db = ref.layer_norm_grad(xin, bna, w, m, v, dy, 1e-5)
            g[key + "_w__out0"], g[key + "_w__out1"], g[key + "_w__out2"] = dx, ds, db
            dx, _, db = ref.layer_norm_grad(xin, bna, None, m, v, dy, 1e-5)
            g[key + "_now__out0"], g[key + "_now__out2"] = dx, db
    path = os.path.join(HERE, "phi_cpu_golden.npz")
    np.savez_compressed(path, **g)
    print(f"wrote {path}: {len(g)} arrays, {os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
