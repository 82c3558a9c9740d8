"""Run ONE BASELINE configuration a few times — the short command ncu wraps (profiles/README.md).
usage: python profiles/run_case.py <case> [iters]
cases: add_bf16 add_int32 ln_bf16 softmax_fp32 softmax_bf16 sum_c1 sum_full_fp32 sum_full_int64 max_full_fp32
Smaller-than-BASELINE tensors (still >> L2) keep ncu's ~40 replays per launch short."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import paddlepaddle_b200 as paddle

F = paddle.nn.functional
case = sys.argv[1]
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
g = torch.Generator(device="cuda")
g.manual_seed(0)
if case == "add_bf16":
    x = torch.randn((64, 2048, 4096), device="cuda", generator=g).to(torch.bfloat16)
    y = torch.randn((4096,), device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: paddle.add(x, y)
elif case == "add_int32":
    x = torch.randint(-2**30, 2**30, (64, 2048, 4096), device="cuda", dtype=torch.int32, generator=g)
    y = torch.randint(-2**30, 2**30, (4096,), device="cuda", dtype=torch.int32, generator=g)
    fn = lambda: paddle.add(x, y)
elif case == "ln_bf16":
    x = torch.randn((65536, 8192), device="cuda", generator=g).to(torch.bfloat16)
    w = torch.rand(8192, device="cuda", generator=g).to(torch.bfloat16)
    b = torch.rand(8192, device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: F.layer_norm(x, [8192], w, b)
elif case in ("softmax_fp32", "softmax_bf16"):
    dt = torch.float32 if case.endswith("fp32") else torch.bfloat16
    x = (torch.randn((8, 32, 2048, 2048), device="cuda", generator=g) * 2).to(dt)  # 1/4 of config 4
    fn = lambda: F.softmax(x, axis=-1)
elif case in ("softmax_vocab_bf16", "softmax_vocab_fp32"):
    dt = torch.float32 if case.endswith("fp32") else torch.bfloat16
    x = (torch.randn((4096, 128256), device="cuda", generator=g) * 2).to(dt)  # LLM-vocabulary rows
    fn = lambda: F.softmax(x, axis=-1)
elif case in ("softmax_mask_bf16", "softmax_tri_bf16"):
    x = (torch.randn((4, 32, 2048, 2048), device="cuda", generator=g) * 2).to(torch.bfloat16)
    mk = torch.randn((4, 1, 2048, 2048), device="cuda", generator=g).to(torch.bfloat16)
    fn = (lambda: F.softmax_mask_fuse(x, mk)) if case == "softmax_mask_bf16" else (lambda: F.softmax_mask_fuse_upper_triangle(x))
elif case == "mulgrad_bf16":
    x = torch.randn((64, 2048, 4096), device="cuda", generator=g).to(torch.bfloat16)
    y = torch.randn((4096,), device="cuda", generator=g).to(torch.bfloat16)
    d = torch.randn((64, 2048, 4096), device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: paddle.multiply_grad(x, y, d)
elif case == "ew_general_bf16":
    x = torch.randn((64, 2048, 1, 2048), device="cuda", generator=g).to(torch.bfloat16)
    y = torch.randn((64, 1, 2, 2048), device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: paddle.add(x, y)
elif case == "ew_outer_bf16":
    x = torch.randn((131072, 4096), device="cuda", generator=g).to(torch.bfloat16)
    y = torch.randn((131072, 1), device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: paddle.add(x, y)
elif case == "softmax_cols_bf16":
    x = (torch.randn((64, 2048, 4096), device="cuda", generator=g) * 2).to(torch.bfloat16)
    fn = lambda: F.softmax(x, axis=1)
elif case in ("argmax_fp32", "sum_rows_fp32", "max_rows_bf16"):
    dt = torch.bfloat16 if case.endswith("bf16") else torch.float32
    x = torch.randn((65536, 8192), device="cuda", generator=g).to(dt)
    fn = {"argmax_fp32": lambda: paddle.argmax(x, axis=-1), "sum_rows_fp32": lambda: paddle.sum(x, axis=-1),
          "max_rows_bf16": lambda: paddle.max(x, axis=-1)}[case]
elif case == "softmax_long_bf16":
    x = (torch.randn((64, 1048576), device="cuda", generator=g) * 2).to(torch.bfloat16)  # 2 MB rows, few of them
    fn = lambda: F.softmax(x, axis=-1)
elif case == "ln_vocab_bf16":
    x = torch.randn((4096, 128256), device="cuda", generator=g).to(torch.bfloat16)
    w = torch.rand(128256, device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: F.layer_norm(x, [128256], w, w)
elif case == "ln_long_bf16":
    x = torch.randn((8192, 65536), device="cuda", generator=g).to(torch.bfloat16)
    w = torch.rand(65536, device="cuda", generator=g).to(torch.bfloat16)
    b = torch.rand(65536, device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: F.layer_norm(x, [65536], w, b)
elif case in ("softmax_grad_fp32", "softmax_grad_bf16"):
    dt = torch.float32 if case.endswith("fp32") else torch.bfloat16
    y = torch.rand((8, 32, 2048, 2048), device="cuda", generator=g).to(dt)
    dy = torch.randn((8, 32, 2048, 2048), device="cuda", generator=g).to(dt)
    fn = lambda: F.softmax_grad(y, dy, axis=-1)
elif case in ("ln_grad_bf16", "ln_grad_bf16_dx"):
    x = torch.randn((65536, 8192), device="cuda", generator=g).to(torch.bfloat16)
    dy = torch.randn((65536, 8192), device="cuda", generator=g).to(torch.bfloat16)
    w = torch.rand(8192, device="cuda", generator=g).to(torch.bfloat16)
    m = torch.zeros(65536, device="cuda")
    v = torch.ones(65536, device="cuda")
    if case.endswith("_dx"):
        fn = lambda: F.layer_norm_grad(x, [8192], w, m, v, dy, need_weight_grad=False, need_bias_grad=False)
    else:
        fn = lambda: F.layer_norm_grad(x, [8192], w, m, v, dy)
elif case in ("max_grad_rows_bf16", "max_grad_cols_fp32", "prod_grad_rows_bf16", "max_grad_mid_bf16"):
    dt = torch.float32 if case.endswith("fp32") else torch.bfloat16
    shape = (256, 64, 16384) if "mid" in case else (65536, 8192)
    axis = 1 if "mid" in case else (0 if "cols" in case else -1)
    x = torch.randn(shape, device="cuda", generator=g).to(dt)
    out = paddle.max(x, axis=axis)
    og = torch.randn(out.shape, device="cuda", generator=g).to(dt)
    fn = (lambda: paddle.prod_grad(x, out, og, axis=axis)) if case.startswith("prod") else (lambda: paddle.max_grad(x, out, og, axis=axis))
elif case == "sum_grad_mid_bf16":
    x = torch.empty((256, 64, 16384), device="cuda", dtype=torch.bfloat16)
    og = torch.randn((256, 16384), device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: paddle.sum_grad(x, og, axis=1)
elif case == "divide_bf16":
    x = torch.randn((64, 2048, 4096), device="cuda", generator=g).to(torch.bfloat16)
    y = (torch.rand((4096,), device="cuda", generator=g) + 0.5).to(torch.bfloat16)
    fn = lambda: paddle.divide(x, y)
elif case in ("ln_bf16_f32params", "ln_bf16_f32params_l1"):
    x = torch.randn((65536, 8192), device="cuda", generator=g).to(torch.bfloat16)
    w = torch.rand(8192, device="cuda", generator=g)
    b = torch.rand(8192, device="cuda", generator=g)
    fn = lambda: F.layer_norm(x, [8192], w, b)
elif case in ("softmax_fp32_8192", "softmax_fp32_1024", "softmax_fp32_4096", "softmax_bf16_8192"):
    D = int(case.rsplit("_", 1)[1])
    dt = torch.bfloat16 if "bf16" in case else torch.float32
    x = (torch.randn(((1 << 29) // D, D), device="cuda", generator=g) * 2).to(dt)
    fn = lambda: F.softmax(x, axis=-1)
elif case in ("ln_bf16_1024", "ln_bf16_2048", "ln_bf16_4096", "ln_fp32_4096", "ln_fp32_8192"):
    D = int(case.rsplit("_", 1)[1])
    dt = torch.bfloat16 if "bf16" in case else torch.float32
    x = torch.randn(((1 << 29) // D, D), device="cuda", generator=g).to(dt)
    w = torch.rand(D, device="cuda", generator=g).to(dt)
    fn = lambda: F.layer_norm(x, [D], w, w)
elif case in ("rms_bf16_1024", "rms_bf16_2048", "rms_bf16_4096", "rms_bf16_8192"):
    D = int(case.rsplit("_", 1)[1])
    x = torch.randn(((1 << 29) // D, D), device="cuda", generator=g).to(torch.bfloat16)
    w = torch.rand(D, device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: F.fused_rms_norm(x, w, None, 1e-6, 1)
elif case == "sum_c1":
    x = torch.rand((4096, 4096), device="cuda", generator=g)
    fn = lambda: paddle.sum(x, axis=-1)
elif case in ("sum_full_fp32", "max_full_fp32"):
    x = torch.rand((1 << 30,), device="cuda", generator=g)
    fn = (lambda: paddle.sum(x)) if case.startswith("sum") else (lambda: paddle.max(x))
elif case == "sum_full_int64":
    x = torch.randint(-2**40, 2**40, (1 << 29,), device="cuda", dtype=torch.int64, generator=g)
    fn = lambda: paddle.sum(x)
else:
    raise SystemExit(f"unknown case {case}")
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
fn()
s.record()
for _ in range(iters):
    fn()
e.record()
torch.cuda.synchronize()
print(f"{case}: {s.elapsed_time(e) / iters:.4f} ms/iter")
