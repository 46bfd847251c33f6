"""NUMA binding helper of the benchmark ranks (utils/affinity.py): without NVML (this container) or
with it, the call must never fail and never leave the process without CPUs."""
import os

from b200conv.utils.affinity import bind_to_gpu_numa


def test_binding_is_harmless_without_a_gpu():
    before = os.sched_getaffinity(0)
    cpus = bind_to_gpu_numa(0)
    after = os.sched_getaffinity(0)
    assert after, "the process must keep at least one CPU"
    if cpus is None:
        assert after == before          # nothing changed when NVML is not there
    else:
        assert set(cpus) == after and after <= before
    os.sched_setaffinity(0, before)
