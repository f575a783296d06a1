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


_TOPO = None


def topo_report():
    """What `nvidia-smi topo -m` says about CPU / NUMA affinity per GPU: {gpu index: {"cpus": "0-31", "numa": "0"}}
    (None when nvidia-smi is not there).  Asked once per process."""
    global _TOPO
    if _TOPO is None:
        _TOPO = {}
        try:
            import subprocess
            out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
            lines = [l for l in out.splitlines() if l.strip()]
            header = next((l for l in lines if "CPU Affinity" in l), None)
            if header:
                cols = [c.strip() for c in header.replace("\x1b[4m", "").replace("\x1b[0m", "").split("\t") if c.strip()]
                ci, ni = cols.index("CPU Affinity"), (cols.index("NUMA Affinity") if "NUMA Affinity" in cols else None)
                for l in lines:
                    f = [c.strip() for c in l.split("\t")]
                    if f and f[0].startswith("GPU") and f[0][3:].isdigit():
                        # the row has one leading name column that the header lacks
                        vals = [c for c in f[1:] if c != ""]
                        ent = {}
                        if ci < len(vals):
                            ent["cpus"] = vals[ci]
                        if ni is not None and ni < len(vals):
                            ent["numa"] = vals[ni]
                        _TOPO[int(f[0][3:])] = ent
        except Exception:  # noqa
            _TOPO = {}
    return _TOPO or None


def bind_to_gpu_node(device_index):
    """Restrict this process to the CPUs of the GPU's NUMA node.  Returns a small report dict that says what was
    found and what was done.  sysfs first; when it reports no node (-1: one NUMA node, or a virtualised topology) the
    CPU affinity column of `nvidia-smi topo -m` is used, and a box whose GPUs all share the same CPU set is left
    unbound (there is nothing to choose between)."""
    node = gpu_numa_node(device_index)
    info = {"gpu": int(device_index), "node": node, "bound": False}
    if node is None:
        topo = topo_report() or {}
        mine = topo.get(int(device_index), {})
        info["nvidia_smi_topo"] = mine or None
        cpu_sets = {v.get("cpus") for v in topo.values()}
        if mine.get("cpus") and len(cpu_sets) > 1:
            try:
                allowed = os.sched_getaffinity(0)
                target = (_parse_cpulist(mine["cpus"]) & allowed) or None
                if target:
                    os.sched_setaffinity(0, target)
                    info.update(bound=True, cpus=len(target), source="nvidia-smi topo -m")
            except (AttributeError, OSError, ValueError):
                pass
        else:
            info["why_unbound"] = "sysfs reports no NUMA node for this GPU and every GPU lists the same CPU affinity (%s)" % (
                mine.get("cpus") or "unknown")
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
