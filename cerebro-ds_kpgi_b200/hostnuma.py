"""Host-side placement for the page feed (NUMA-pinned readers).

Eight GPUs pulling pinned pages at PCIe speed need ~440 GB/s of host memory reads; that only works when
each rank's staging buffers live on the NUMA node its GPU hangs off.  Linux places a page on the node of
the CPU that first touches it, so binding the process to the GPU-local CPUs *before* the pinned buffers are
allocated and filled is enough.  Nothing here is on the decode path; it is plumbing for `bench.py` and for
loaders that opt in (CDS_NUMA_BIND=1).
"""
import os


def _read(path):
    try:
        with open(path) as f:
            return f.read().strip()
    except OSError:
        return None


def _parse_cpulist(text):
    cpus = set()
    for part in (text or "").split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_numa_node(device_index):
    """NUMA node of a CUDA device from sysfs (None when unknown)."""
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
    except Exception:  # noqa
        return None
    node = _read("/sys/bus/pci/devices/%s/numa_node" % bdf)
    try:
        node = int(node)
    except (TypeError, ValueError):
        return None
    return node if node >= 0 else None


def bind_to_gpu_node(device_index):
    """Restrict this process to the CPUs of the GPU's NUMA node.  Returns a small report dict."""
    node = gpu_numa_node(device_index)
    info = {"gpu": int(device_index), "node": node, "bound": False}
    if node is None:
        return info
    cpus = _parse_cpulist(_read("/sys/devices/system/node/node%d/cpulist" % node))
    try:
        allowed = os.sched_getaffinity(0)
        target = (cpus & allowed) or None
        if target:
            os.sched_setaffinity(0, target)
            info.update(bound=True, cpus=len(target))
    except (AttributeError, OSError):
        pass
    return info
