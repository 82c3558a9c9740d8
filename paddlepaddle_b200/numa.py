"""CPU / memory locality for the host-buffer path (paddlepaddle_b200.host, bench.py's e2e leg).

One process per GPU.  Pinned staging memory is placed by first touch, so a rank that runs on the CPUs of the socket its
GPU hangs off gets host buffers whose PCIe transfers do not cross the inter-socket link; eight ranks that all start on
socket 0 share one memory controller set and one link instead.  `bind_to_gpu_numa` restricts the calling process to the
CPUs NVML reports as local to the GPU (the same list `nvidia-smi topo -m` prints), intersected with what the container
allows; it changes nothing when NVML is missing, the list is empty, or the intersection is empty."""
import os


def gpu_cpu_affinity(index):
    """CPUs local to GPU `index` according to NVML, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            h = pynvml.nvmlDeviceGetHandleByIndex(int(index))
            ncpu = os.cpu_count() or 1
            words = (max(ncpu, 64) + 63) // 64
            mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
            cpus = {w * 64 + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
            return cpus or None
        finally:
            pynvml.nvmlShutdown()
    except Exception:
        return None


def bind_to_gpu_numa(local_rank):
    """Bind this process to the CPUs next to GPU `local_rank` (visible-device order).  Returns a small dict describing
    what happened, for logs."""
    info = {"bound": False}
    try:
        allowed = os.sched_getaffinity(0)
    except Exception:
        return info
    info["allowed_cpus"] = len(allowed)
    index = local_rank
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            index = int(vis.split(",")[local_rank])
        except Exception:
            index = local_rank
    local = gpu_cpu_affinity(index)
    if not local:
        info["why"] = "NVML reports no CPU affinity"
        return info
    want = allowed & local
    info["gpu_local_cpus"] = len(local)
    if not want:
        info["why"] = "none of the GPU-local CPUs is in this container's CPU set"
        return info
    if want != allowed:
        try:
            os.sched_setaffinity(0, want)
            info["bound"] = True
        except Exception as exc:  # noqa: BLE001
            info["why"] = f"sched_setaffinity: {exc}"
            return info
    else:
        info["why"] = "already on GPU-local CPUs only"
    info["cpus"] = len(want)
    return info
