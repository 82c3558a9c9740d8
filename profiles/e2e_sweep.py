"""Host-buffer end-to-end add (pinned host -> H2D -> kernel -> D2H) for one slab size / stream count
(PDB200_HOST_SLAB_MB, PDB200_HOST_STREAMS are read at import)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from paddlepaddle_b200 import host
hx = torch.empty((64, 2048, 4096), dtype=torch.bfloat16, pin_memory=True); hx.view(torch.int16).fill_(16000)
hy = torch.empty((4096,), dtype=torch.bfloat16, pin_memory=True); hy.view(torch.int16).fill_(16000)
ho = torch.empty_like(hx, pin_memory=True)
host.add(hx, hy, out=ho); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5): host.add(hx, hy, out=ho)
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / 5 * 1e3
print(f"slab={os.environ.get('PDB200_HOST_SLAB_MB','64')}MB streams={os.environ.get('PDB200_HOST_STREAMS','2')}: {ms:.2f} ms  {2147491840/ms/1e6:.1f} GB/s (each direction {1073741824/ms/1e6:.1f} GB/s)")
