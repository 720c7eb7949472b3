"""Host-side placement for the host-buffer front end (``StepPipeline``): on a multi-socket box the pinned staging
buffers of a rank should live on the NUMA node its GPU hangs off, otherwise every PCIe copy crosses the socket
interconnect and eight ranks share one memory controller.  ``bind_to_gpu_numa`` pins the calling process to the
CPUs NVML reports as local to the GPU, so that the pinned allocations that follow are first-touched there
(what NCCL does for its proxy threads).  Plumbing only; a no-op when NVML is not available."""
from __future__ import annotations

import os
from typing import List, Optional


def gpu_local_cpus(device_index: int) -> Optional[List[int]]:
    try:
        import pynvml
        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        if visible:                                                  # torch's index -> NVML's index
            ids = [v.strip() for v in visible.split(",") if v.strip()]
            if device_index < len(ids) and ids[device_index].isdigit():
                device_index = int(ids[device_index])
        h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        ncpu = os.cpu_count() or 64
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [64 * w + b for w, mask in enumerate(words) for b in range(64) if (int(mask) >> b) & 1]
        return cpus or None
    except Exception:
        return None


def bind_to_gpu_numa(device_index: int) -> Optional[List[int]]:
    """Restrict this process to the CPUs local to GPU ``device_index`` (intersected with what it is already allowed
    to run on).  Returns the CPU list that was set, or None when nothing was changed."""
    cpus = gpu_local_cpus(device_index)
    if not cpus or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        allowed = os.sched_getaffinity(0)
        target = sorted(set(cpus) & set(allowed))
        if not target or set(target) == set(allowed):
            return None
        os.sched_setaffinity(0, target)
        return target
    except OSError:
        return None
