"""Host-side placement for the host-pointer calls: one process per GPU, pinned to the CPU cores NVML reports as local
to that GPU, so that the pinned PCM / feature buffers it allocates afterwards (first touch) sit on the GPU's own NUMA
node and the PCIe copies do not cross the socket interconnect.  Matters from 4 GPUs per node upwards."""
from __future__ import annotations

import os


def bind_to_gpu(device_index: int) -> dict:
    """Restrict the calling process to the CPUs local to GPU `device_index` (intersected with the CPUs the process is
    already allowed on).  Returns what was done; never raises — without NVML or on a cpuset that excludes the GPU's
    cores the process is left as it is."""
    info = {"bound": False, "cpus": len(os.sched_getaffinity(0))}
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = device_index
            if visible:                                   # NVML ignores CUDA_VISIBLE_DEVICES
                ids = [v.strip() for v in visible.split(",") if v.strip()]
                if device_index < len(ids) and ids[device_index].isdigit():
                    idx = int(ids[device_index])
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            n_words = (os.cpu_count() + 63) // 64
            words = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
            local = {w * 64 + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1}
        finally:
            pynvml.nvmlShutdown()
        allowed = os.sched_getaffinity(0)
        cpus = local & allowed
        if cpus and cpus != allowed:
            os.sched_setaffinity(0, cpus)
            info.update(bound=True, cpus=len(cpus))
        info["gpu_local_cpus"] = len(local)
    except Exception as exc:                              # placement is an optimisation, not a requirement
        info["why_not"] = f"{type(exc).__name__}: {exc}"[:120]
    return info
