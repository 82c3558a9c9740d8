"""less_than bf16 [64,2048,4096] vs [4096] -> bool, a few times (the short command ncu wraps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import paddlepaddle_b200 as paddle
g = torch.Generator(device="cuda"); g.manual_seed(0)
x = torch.randn((64, 2048, 4096), device="cuda", generator=g).to(torch.bfloat16)
y = torch.randn((4096,), device="cuda", generator=g).to(torch.bfloat16)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
paddle.less_than(x, y); torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(n): paddle.less_than(x, y)
e.record(); torch.cuda.synchronize()
print("less_than bf16:", s.elapsed_time(e) / n, "ms/iter")
