"""Host-side placement logic of the e2e path (no GPU): NVML absent or silent must leave the process untouched."""
import os

from thormang_isaacgym_b200 import host_io


def test_near_gpu_is_a_no_op_without_topology(monkeypatch):
    monkeypatch.setattr(host_io, "gpu_local_cpus", lambda i: (None, None))
    before = os.sched_getaffinity(0)
    report = {}
    with host_io.near_gpu(0, report):
        assert os.sched_getaffinity(0) == before
    assert os.sched_getaffinity(0) == before
    assert report["affinity_narrowed_for_allocation"] is False and report["gpu_local_cpus"] is None


def test_near_gpu_narrows_and_restores(monkeypatch):
    before = sorted(os.sched_getaffinity(0))
    if len(before) < 2:
        return
    local = before[: len(before) // 2]
    monkeypatch.setattr(host_io, "gpu_local_cpus", lambda i: (local + [10 ** 6], 1))     # a CPU this process may not use is ignored
    report = {}
    with host_io.near_gpu(3, report):
        assert sorted(os.sched_getaffinity(0)) == local
    assert sorted(os.sched_getaffinity(0)) == before
    assert report == {"gpu": 3, "numa_node": 1, "gpu_local_cpus": len(local) + 1, "affinity_narrowed_for_allocation": True}


def test_near_gpu_restores_on_error(monkeypatch):
    before = sorted(os.sched_getaffinity(0))
    if len(before) < 2:
        return
    monkeypatch.setattr(host_io, "gpu_local_cpus", lambda i: (before[:1], 0))
    try:
        with host_io.near_gpu(0):
            raise RuntimeError("allocation failed")
    except RuntimeError:
        pass
    assert sorted(os.sched_getaffinity(0)) == before


def test_gpu_local_cpus_without_nvml_device():
    cpus, node = host_io.gpu_local_cpus(0)      # this container has no GPU: the probe must answer, not raise
    assert cpus is None or isinstance(cpus, list)
