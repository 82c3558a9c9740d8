"""Geometry sweep of the pipelined row kernels through the PDB200_* tuning knobs (read at every call).
usage: python profiles/sweep.py <case> "C=2,4" "CTAS=2,3,4" "STAGES=3,4,6" [iters]"""
import itertools
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import paddlepaddle_b200 as paddle

F = paddle.nn.functional
case = sys.argv[1]
grid = {}
iters = 10
for a in sys.argv[2:]:
    if "=" in a:
        k, v = a.split("=")
        grid[k] = v.split(",")
    else:
        iters = int(a)
g = torch.Generator(device="cuda")
g.manual_seed(0)
if case == "ln_bf16":
    x = torch.randn((65536, 8192), device="cuda", generator=g).to(torch.bfloat16)
    w = torch.rand(8192, device="cuda", generator=g).to(torch.bfloat16)
    b = torch.rand(8192, device="cuda", generator=g).to(torch.bfloat16)
    fn = lambda: F.layer_norm(x, [8192], w, b)
    nbytes = 2 * x.numel() * 2
    cname = "PDB200_LN_PIPE_C"
else:
    dt = torch.float32 if case.endswith("fp32") else torch.bfloat16
    x = (torch.randn((8, 32, 2048, 2048), device="cuda", generator=g) * 2).to(dt)
    fn = lambda: F.softmax(x, axis=-1)
    nbytes = 2 * x.numel() * x.element_size()
    cname = "PDB200_SOFTMAX_PIPE_C"
names = {"C": cname, "CTAS": "PDB200_PIPE_MAX_CTAS", "STAGES": "PDB200_PIPE_MAX_STAGES", "CONTIG": "PDB200_PIPE_CONTIG"}
keys = list(grid)
for combo in itertools.product(*[grid[k] for k in keys]):
    for k, v in zip(keys, combo):
        os.environ[names[k]] = v
    fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / iters
    print(case, " ".join(f"{k}={v}" for k, v in zip(keys, combo)), f"{ms:.4f} ms  {nbytes / ms / 1e6:.0f} GB/s", flush=True)
