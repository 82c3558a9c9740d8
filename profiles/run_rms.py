"""rms_norm bf16 [65536,8192] a few times (the short command ncu wraps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import paddlepaddle_b200 as paddle
F = paddle.nn.functional
g = torch.Generator(device="cuda"); g.manual_seed(0)
x = torch.randn((65536, 8192), device="cuda", generator=g).to(torch.bfloat16)
w = (torch.rand(8192, device="cuda", generator=g) + 0.5).to(torch.bfloat16)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
F.fused_rms_norm(x, w, None, 1e-6, 1); torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(n): F.fused_rms_norm(x, w, None, 1e-6, 1)
e.record(); torch.cuda.synchronize()
print("rms_norm bf16:", s.elapsed_time(e) / n, "ms/iter")
