"""Bind a rank's host threads (and so its pinned buffers) to the CPUs next to its GPU.

With one process per GPU the host-buffer entry points (``rt_render_frame_host``) and the
turntable's pinned slots copy 25 MB per 1080p frame over PCIe.  On a two-socket box a copy
whose pinned pages sit on the far socket crosses the inter-socket link and the eight ranks
together stall near 100 GB/s; first-touch places pinned pages on the node of the allocating
thread, so binding the process to the GPU's own CPUs before anything is allocated keeps every
copy on its own socket.  NVML knows the mask (``nvmlDeviceGetCpuAffinity``).  Best effort:
containers that hide the topology, or an affinity mask that excludes the GPU's CPUs, leave the
process where it was.
"""

from __future__ import annotations

import os
from typing import Optional, Set


def device_cpus(device_index: int) -> Optional[Set[int]]:
    """CPUs NVML reports as local to ``device_index`` (None if it cannot tell)."""
    try:
        import pynvml

        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        words = (os.cpu_count() or 64) // 64 + 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
    except Exception:
        return None
    cpus = {64 * w + b for w, word in enumerate(mask) for b in range(64) if (int(word) >> b) & 1}
    return cpus or None


def bind_to_device(device_index: int) -> Optional[Set[int]]:
    """Restrict this process to the CPUs local to its GPU; returns the set in force, or None if
    nothing was changed."""
    if not hasattr(os, "sched_setaffinity"):
        return None
    local = device_cpus(device_index)
    if not local:
        return None
    allowed = os.sched_getaffinity(0)
    target = local & allowed
    if not target or target == allowed:
        return None
    try:
        os.sched_setaffinity(0, target)
    except OSError:
        return None
    return target
