"""The Paddle-side binding really registers: paddlepaddle_b200/plugin/phi_kernels.cc, compiled against the
reference headers, is dlopen'ed next to Paddle's own KernelFactory (built from the reference sources by
oracle/build_registry.sh) and its PD_REGISTER_KERNEL initialisers must leave the B200 kernels under
Paddle's (name, backend, ALL_LAYOUT, dtype) keys — overwriting what was there, which is how an in-tree
kernel gets replaced (paddle/phi/core/kernel_registry.h:147-148).  Runs in a subprocess so that the
registry starts empty.  CPU only; skipped where the two libraries were not built."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REG = os.path.join(ROOT, "oracle", "_ref", "libphi_registry_min.so")
PLUG = os.path.join(ROOT, "oracle", "_ref", "libpdb200_phi.so")

# phi::DataType values (paddle/phi/common/data_type.h:44-62)
DT = {"bool": 1, "uint8": 2, "int8": 3, "int16": 5, "int32": 7, "int64": 9, "float32": 10, "float64": 11,
      "float16": 15, "bfloat16": 16}

CHILD = r'''
import ctypes, json, sys
reg = ctypes.CDLL(sys.argv[1], mode=ctypes.RTLD_GLOBAL)
reg.phi_min_variadic_fn.restype = ctypes.c_void_p
out = {"before": reg.phi_min_kernel_count()}
reg.phi_min_register_dummy(b"sum", b"GPU", 10)          # stand-in for the in-tree GPU kernel
reg.phi_min_register_dummy(b"softmax", b"GPUDNN", 16)
out["dummy_fn"] = reg.phi_min_variadic_fn(b"sum", b"GPU", 10)
ctypes.CDLL(sys.argv[2], mode=ctypes.RTLD_GLOBAL)        # static initialisers run here
out["after"] = reg.phi_min_kernel_count()
out["sum_fn"] = reg.phi_min_variadic_fn(b"sum", b"GPU", 10)
out["softmax_dnn_fn"] = reg.phi_min_variadic_fn(b"softmax", b"GPUDNN", 16)
keys = json.loads(sys.argv[3])
out["has"] = {f"{n}/{b}/{d}": reg.phi_min_has_kernel(n.encode(), b.encode(), d) for n, b, d in keys}
ar = {}
a, b_, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
for n in ["add", "divide", "sum", "sum_raw", "mean", "max_raw", "prod", "argmax", "softmax", "layer_norm", "softmax_grad",
          "layer_norm_grad", "less_than", "logical_not", "logical_xor", "scale", "cast", "where", "add_grad", "multiply_grad", "sum_grad", "mean_grad", "max_grad", "prod_grad", "cross_entropy_with_softmax"]:
    reg.phi_min_arity(n.encode(), b"GPU", 10, ctypes.byref(a), ctypes.byref(b_), ctypes.byref(c))
    ar[n] = [a.value, b_.value, c.value]
out["arity"] = ar
out["ln_out_dtypes"] = [reg.phi_min_output_dtype(b"layer_norm", b"GPU", 16, i) for i in range(3)]
out["sum_out_dtype"] = reg.phi_min_output_dtype(b"sum", b"GPU", 7, 0)
out["argmax_out_dtype"] = reg.phi_min_output_dtype(b"argmax", b"GPU", 10, 0)
out["family_out_dtypes"] = [reg.phi_min_output_dtype(n, b"GPU", 10, 0) for n in (b"equal", b"logical_or", b"cast", b"scale")]
print(json.dumps(out))
'''

EXPECT = {  # the registrations of plugin/phi_kernels.cc
    "add": "float32 float64 int16 int32 bool uint8 int8 int64 float16 bfloat16",
    "multiply": "float32 float64 int16 int32 bool uint8 int8 int64 float16 bfloat16",
    "subtract": "float32 float64 int16 int32 int64 float16 bfloat16",
    "divide": "float32 float64 int8 uint8 int16 int32 int64 float16 bfloat16",
    "maximum": "float32 float64 int32 int64 float16 bfloat16", "minimum": "float32 float64 int32 int64 float16 bfloat16",
    "sum": "bool float32 float64 float16 bfloat16 int8 uint8 int16 int32 int64",
    "sum_raw": "bool float32 float64 float16 bfloat16 int8 uint8 int16 int32 int64",
    "fused_softmax_mask": "float32 float16 bfloat16", "fused_softmax_mask_upper_triangle": "float32 float16 bfloat16",
    "fused_bias_dropout_residual_layer_norm": "float32 float64 float16 bfloat16",
    "mean": "float32 float64 float16 bfloat16 int32 int64", "mean_raw": "float32 float64 float16 bfloat16 int32 int64",
    "max": "float32 float64 int32 int64 float16 bfloat16", "max_raw": "float32 float64 int32 int64 float16 bfloat16",
    "min": "float32 float64 int32 int64 float16 bfloat16", "min_raw": "float32 float64 int32 int64 float16 bfloat16",
    "prod": "float32 float64 int32 int64 float16 bfloat16",
    "argmax": "float16 bfloat16 float32 float64 int32 int64 int16 uint8",
    "argmin": "float16 bfloat16 float32 float64 int32 int64 int16 uint8",
    "rms_norm": "float32 float16 bfloat16",
    "log_softmax": "float32 float64 float16 bfloat16", "cross_entropy_with_softmax": "float32 float64 float16",
    "softmax": "float32 float64 float16 bfloat16", "layer_norm": "float32 float64 float16 bfloat16",
    "softmax_grad": "float32 float64 float16 bfloat16", "layer_norm_grad": "float32 float64 float16 bfloat16",
    "log_softmax_grad": "float32 float64 float16 bfloat16", "cross_entropy_with_softmax_grad": "float32 float64 float16",
    **{n: "bool int32 uint8 int8 int16 int64 float32 float64 float16 bfloat16"
       for n in ("less_than", "less_equal", "greater_than", "greater_equal", "equal", "not_equal")},
    **{n: "float32 float16 bfloat16 float64 bool int64 int32 int8 int16"
       for n in ("logical_and", "logical_or", "logical_xor", "logical_not")},
    **{n: "bool uint8 int8 int16 int32 int64" for n in ("bitwise_and", "bitwise_or", "bitwise_xor", "bitwise_not")},
    "bitwise_left_shift": "uint8 int8 int16 int32 int64", "bitwise_right_shift": "uint8 int8 int16 int32 int64",
    "scale": "bool float32 float64 float16 bfloat16 uint8 int8 int16 int32 int64",
    "cast": "float32 float64 int32 int64 int16 bool int8 uint8 float16 bfloat16",
    "where": "float32 float64 int32 int64 float16 bfloat16",
    "sum_grad": "float32 float64 float16 bfloat16 int32 int64", "mean_grad": "float32 float64 float16 bfloat16 int32 int64",
    "max_grad": "float32 float64 int32 int64 float16 bfloat16", "min_grad": "float32 float64 int32 int64 float16 bfloat16",
    "prod_grad": "float32 float64 int32 int64 float16 bfloat16",
    "fused_bias_residual_layernorm": "float32 float16 bfloat16",
    "add_grad": "float32 float64 int32 int64 float16 bfloat16", "multiply_grad": "float32 float16 float64 int32 int64 bfloat16",
}


@pytest.mark.skipif(not (os.path.exists(REG) and os.path.exists(PLUG)),
                    reason="oracle/_ref registry libraries not built (oracle/build_registry.sh needs /root/reference)")
def test_pd_register_kernel_overwrites_paddles_keys():
    keys = [(n, "GPU", DT[d]) for n, ds in EXPECT.items() for d in ds.split()]
    keys += [("softmax", "GPUDNN", DT[d]) for d in EXPECT["softmax"].split()]
    keys += [("softmax_grad", "GPUDNN", DT[d]) for d in EXPECT["softmax_grad"].split()]
    r = subprocess.run([sys.executable, "-c", CHILD, REG, PLUG, json.dumps(keys)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["before"] == 0
    assert out["after"] == len(keys), "every registration landed, and nothing else"
    missing = [k for k, v in out["has"].items() if not v]
    assert not missing, missing
    # the dummies (stand-ins for the in-tree kernels) were REPLACED: a real variadic function is there now
    assert out["dummy_fn"] is None and out["sum_fn"] and out["softmax_dnn_fn"]
    # what the registrar parsed out of our C++ signatures == the in-tree signatures (inputs, attributes, outputs)
    assert out["arity"] == {"add": [2, 0, 1], "divide": [2, 0, 1], "sum": [1, 3, 1], "sum_raw": [1, 4, 1], "mean": [1, 2, 1],
                            "max_raw": [1, 3, 1], "prod": [1, 3, 1], "argmax": [1, 4, 1], "softmax": [1, 1, 1],
                            "layer_norm": [3, 2, 3], "softmax_grad": [2, 1, 1], "layer_norm_grad": [6, 2, 3],
                            "less_than": [2, 0, 1], "logical_not": [1, 0, 1], "logical_xor": [2, 0, 1], "scale": [1, 3, 1],
                            "cast": [1, 1, 1], "where": [3, 0, 1], "add_grad": [3, 1, 2], "multiply_grad": [3, 1, 2], "sum_grad": [2, 3, 1],
                            "mean_grad": [2, 3, 1], "max_grad": [3, 3, 1], "prod_grad": [3, 3, 1],
                            "cross_entropy_with_softmax": [2, 5, 2]}
    # output dtypes left UNDEFINED where the in-tree registration bodies do so
    assert out["ln_out_dtypes"] == [DT["bfloat16"], 0, 0] and out["sum_out_dtype"] == 0 and out["argmax_out_dtype"] == 0
    # compare / logical outputs are BOOL, cast's is left to InferMeta, scale keeps the input dtype
    assert out["family_out_dtypes"] == [DT["bool"], DT["bool"], 0, DT["float32"]]
