"""NUMA placement of a rank's host buffers.

Pinned staging memory is first-touched by the process that allocates it.  On a two-socket host
with four GPUs per socket, a rank whose pages sit on the other socket sends every upload and
download across the socket interconnect; with eight ranks copying at once that interconnect is
what the copies queue on.  ``bind_to_gpu_numa`` restricts the calling process to the CPUs NVML
reports as local to the GPU, so that the pinned buffers allocated afterwards land on that node
(the reference leaves placement to ``mpiexec -f machines.txt``, README.md:20-24).
"""
from __future__ import annotations

import os


def bind_to_gpu_numa(device_index: int) -> list[int] | None:
    """Restrict this process to the CPUs local to ``device_index``.  Returns the CPU list, or None
    when NVML or the scheduler interface is not available (nothing is changed then)."""
    try:
        import pynvml

        pynvml.nvmlInit()
        try:
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[device_index]) if visible and visible.split(",")[device_index].isdigit() \
                else device_index
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            ncpu = os.cpu_count() or 1
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        finally:
            pynvml.nvmlShutdown()
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:  # noqa: BLE001  (no NVML, no permission, not Linux: placement stays as it is)
        return None
