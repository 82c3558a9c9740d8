"""GB/s of the reduction engine on shapes that hit each route (rows_warp / rows_block / cols, splits).
usage: python profiles/reduce_bench.py [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import paddlepaddle_b200 as paddle

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 10
g = torch.Generator(device="cuda")
g.manual_seed(0)


def rnd(shape, dt):
    if dt in (torch.int64, torch.int32):
        return torch.randint(-2**20, 2**20, shape, device="cuda", dtype=dt, generator=g)
    if dt == torch.float32:
        return torch.randn(shape, device="cuda", generator=g)
    return torch.randn(shape, device="cuda", generator=g).to(dt)


CASES = [
    ("sum  fp32 [4096,4096] axis=-1 (C1)", (4096, 4096), torch.float32, lambda x: paddle.sum(x, axis=-1)),
    ("sum  fp32 [262144,4096] axis=-1", (262144, 4096), torch.float32, lambda x: paddle.sum(x, axis=-1)),
    ("sum  bf16 [524288,4096] axis=-1", (524288, 4096), torch.bfloat16, lambda x: paddle.sum(x, axis=-1)),
    ("max  fp32 [4194304,256] axis=-1", (4194304, 256), torch.float32, lambda x: paddle.max(x, axis=-1)),
    ("sum  fp32 [16777216,64] axis=-1", (16777216, 64), torch.float32, lambda x: paddle.sum(x, axis=-1)),
    ("sum  fp32 [64,16777216] axis=-1", (64, 16777216), torch.float32, lambda x: paddle.sum(x, axis=-1)),
    ("sum  fp32 [262144,4096] axis=0", (262144, 4096), torch.float32, lambda x: paddle.sum(x, axis=0)),
    ("sum  bf16 [262144,8192] axis=0", (262144, 8192), torch.bfloat16, lambda x: paddle.sum(x, axis=0)),
    ("sum  fp32 [4096,262144] axis=0", (4096, 262144), torch.float32, lambda x: paddle.sum(x, axis=0)),
    ("mean fp32 [512,4096,512] axis=1", (512, 4096, 512), torch.float32, lambda x: paddle.mean(x, axis=1)),
    ("sum  fp32 [64,512,64,512] axis=(1,3)", (64, 512, 64, 512), torch.float32, lambda x: paddle.sum(x, axis=(1, 3))),
    ("sum  fp32 2^30 all", (1 << 30,), torch.float32, lambda x: paddle.sum(x)),
    ("max  fp32 2^30 all", (1 << 30,), torch.float32, lambda x: paddle.max(x)),
    ("sum  int64 2^29 all", (1 << 29,), torch.int64, lambda x: paddle.sum(x)),
    ("prod fp32 2^30 all", (1 << 30,), torch.float32, lambda x: paddle.prod(x)),
    ("argmax fp32 [262144,4096] axis=-1", (262144, 4096), torch.float32, lambda x: paddle.argmax(x, axis=-1)),
    ("argmax fp32 [262144,4096] axis=0", (262144, 4096), torch.float32, lambda x: paddle.argmax(x, axis=0)),
    ("argmax fp32 2^30 flat", (1 << 30,), torch.float32, lambda x: paddle.argmax(x)),
    ("argmax bf16 [524288,4096] axis=-1", (524288, 4096), torch.bfloat16, lambda x: paddle.argmax(x, axis=-1)),
]
if os.environ.get("BIG"):  # the C5 shard sizes (17 GB / 8.6 GB inputs)
    CASES += [
        ("sum  int64 2^31 all (C5 shard)", (1 << 31,), torch.int64, lambda x: paddle.sum(x)),
        ("max  int64 2^31 all (C5 shard)", (1 << 31,), torch.int64, lambda x: paddle.max(x)),
        ("sum  fp32 2^31 all (C5 shard)", (1 << 31,), torch.float32, lambda x: paddle.sum(x)),
        ("max  fp32 2^31 all (C5 shard)", (1 << 31,), torch.float32, lambda x: paddle.max(x)),
        ("argmax fp32 2^31 flat", (1 << 31,), torch.float32, lambda x: paddle.argmax(x)),
    ]
only = os.environ.get("ONLY")
for name, shape, dt, fn in CASES:
    if only and not any(o in name for o in only.split(";")):
        continue
    x = rnd(shape, dt)
    out = fn(x)
    for _ in range(3):  # let the caching allocator settle (a second live output block is a cudaMalloc)
        fn(x)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn(x)
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / iters
    nbytes = x.numel() * x.element_size() + out.numel() * out.element_size()
    print(f"{name:44s} {ms:8.4f} ms {nbytes / ms / 1e6:8.0f} GB/s", flush=True)
    del x, out
    torch.cuda.empty_cache()
