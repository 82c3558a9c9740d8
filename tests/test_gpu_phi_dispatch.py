"""Every body of paddlepaddle_b200/plugin/phi_kernels.cc, CALLED on the GPU the way Paddle's generated eager API calls a
kernel (paddle/phi/api/generator/api_base.py:1327-1354):

    KernelFactory::Instance().SelectKernelOrThrowError(name, {backend, ALL_LAYOUT, dtype})
      -> kernel.GetVariadicKernelFn<in-tree signature>() -> (*kernel_fn)(dev_ctx, DenseTensor inputs, attrs, DenseTensor* outputs)

with outputs that carry only their meta (dims, dtype) and are allocated by the kernel through `dev_ctx.Alloc`, scratch
from the same allocator, everything enqueued on `dev_ctx.stream()`.  The host is oracle/_ref/libphi_registry_min.so:
Paddle's own KernelFactory / DenseTensor / DeviceContext compiled from the reference sources plus a GPUContext whose
stream is a real cudaStream_t and whose allocator is cudaMallocAsync on that stream (oracle/phi_registry_shim.cc).
Each result must equal, BIT FOR BIT, what the ctypes path (the Python mirror -> C-ABI) gives for the same inputs —
those are the results the parity tests pin to the oracle.  Runs in a subprocess (the registry must start empty and the
plugin's static initialisers run at dlopen)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REG = os.path.join(ROOT, "oracle", "_ref", "libphi_registry_min.so")
PLUG = os.path.join(ROOT, "oracle", "_ref", "libpdb200_phi.so")
CHILD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "phi_dispatch_child.py")


@pytest.mark.skipif(not (os.path.exists(REG) and os.path.exists(PLUG)),
                    reason="oracle/_ref registry libraries not built (oracle/build_registry.sh needs /root/reference)")
def test_every_plugin_body_runs_through_paddles_dispatch():
    r = subprocess.run([sys.executable, CHILD, REG, PLUG], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert not out["failed"], out["failed"]
    # one call at least for every kernel NAME the plugin registers
    assert not out["names_not_called"], out["names_not_called"]
    assert out["calls"] >= 60 and out["alloc_calls"] >= out["calls"], out
    print(f"{out['calls']} kernels called through SelectKernelOrThrowError -> GetVariadicKernelFn; "
          f"{out['alloc_calls']} dev_ctx.Alloc calls served by cudaMallocAsync")
