"""Host-side feed for the drop-in reader: relation files -> HBM and decoded values -> numpy, without
pinning gigabytes per call.

`da.input_fn` is called once per (partition, mode) and hands numpy arrays back (da.py:55-58 in the
reference).  Measured on a 1 GiB segment, page-locking a 1 GiB staging buffer (0.37 s + 0.09 s to free
it) and the pageable device-to-host copy (0.22 s) dwarf the decode itself (a few ms).  So:

* a small ring of page-locked slots (16 x 16 MiB) is allocated once per process and reused by every call;
* file chunks are read into free slots by a few threads (`readinto` releases the GIL) while earlier
  slots are on their way to the GPU (`cudaMemcpyAsync` on a copy stream);
* results come back the same way: device -> pinned slot -> the caller-owned numpy array, the last hop
  copied (and the fresh pages first-touched) by several threads.

Plumbing only: no decoding happens here.
"""
import os
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import segment as _segment

BLOCK_SIZE = 32768
# ring geometry (environment overrides are for measurements: tools/dropin_bench.py)
# 16 slots of 16 MiB and as many threads (up to the cores present) read a 1 GiB relation in 0.05 s; the first
# version's 4 x 64 MiB with 4 threads needed 0.07-0.28 s (profiles/README.md)
SLOT_BYTES = int(os.environ.get("CDS_IO_SLOT_MIB", "16")) << 20
N_SLOTS = int(os.environ.get("CDS_IO_SLOTS", "16"))
IO_THREADS = int(os.environ.get("CDS_IO_THREADS", str(max(1, min(16, os.cpu_count() or 4)))))

_lock = threading.Lock()
_use = threading.RLock()             # one transfer at a time through a device's ring (da.prefetch_device runs one in a thread)
_rings = {}
_pool = None


def _executor():
    global _pool
    if _pool is None:
        _pool = ThreadPoolExecutor(max_workers=IO_THREADS, thread_name_prefix="cds-io")
    return _pool


class _Ring:
    def __init__(self, device):
        import torch
        self.device = device
        self.slots = [_segment.PinnedBuffer(SLOT_BYTES) for _ in range(N_SLOTS)]
        self.tensors = [torch.from_numpy(s.array) for s in self.slots]
        self.events = [torch.cuda.Event() for _ in range(N_SLOTS)]
        self.busy = [False] * N_SLOTS
        self.stream = torch.cuda.Stream(device=device)


def _ring():
    import torch
    dev = torch.cuda.current_device()
    with _lock:
        if dev not in _rings:
            _rings[dev] = _Ring(dev)
        return _rings[dev]


def _plan_files(filename, relation_files):
    """(path, offset_in_file, nbytes) chunks of every whole page of the relation, and the `partial` flag
    (a trailing short page: the reference raises Exception("FATAL") there, pg_page_reader.py:375-377)."""
    chunks, total, partial = [], 0, False
    for path in relation_files(filename):
        size = os.path.getsize(path)
        whole = size - size % BLOCK_SIZE
        for off in range(0, whole, SLOT_BYTES):
            n = min(SLOT_BYTES, whole - off)
            chunks.append((path, off, n, total + off))
        total += whole
        if size != whole:
            partial = True
            break
    return chunks, total, partial


def _read_chunk(path, off, n, dst):
    with open(path, "rb", buffering=0) as f:
        f.seek(off)
        view = memoryview(dst)[:n]
        got = 0
        while got < n:
            k = f.readinto(view[got:])
            if not k:
                raise IOError("short read in %s" % path)
            got += k
    return n


def relation_to_device(filename, relation_files):
    """All whole pages of a relation as one uint8 CUDA tensor (16-byte aligned), plus the partial flag."""
    import torch
    with _use:
        return _relation_to_device(filename, relation_files)


def _relation_to_device(filename, relation_files):
    import torch
    ring = _ring()
    chunks, total, partial = _plan_files(filename, relation_files)
    out = torch.empty(max(total, 16), dtype=torch.uint8, device="cuda")
    if total == 0:
        return out[:0], partial
    ex = _executor()
    cur = torch.cuda.current_stream()
    ring.stream.wait_stream(cur)
    pending = []                     # (future, slot, chunk)
    next_chunk = 0

    def submit(slot):
        nonlocal next_chunk
        path, off, n, dst_off = chunks[next_chunk]
        ring.events[slot].synchronize()                  # the slot's previous copy has left it
        fut = ex.submit(_read_chunk, path, off, n, ring.slots[slot].array)
        pending.append((fut, slot, chunks[next_chunk]))
        next_chunk += 1

    for slot in range(min(N_SLOTS, len(chunks))):
        submit(slot)
    while pending:
        fut, slot, (path, off, n, dst_off) = pending.pop(0)
        fut.result()
        with torch.cuda.stream(ring.stream):
            out[dst_off:dst_off + n].copy_(ring.tensors[slot][:n], non_blocking=True)
            ring.events[slot].record(ring.stream)
        if next_chunk < len(chunks):
            submit(slot)
    cur.wait_stream(ring.stream)
    return out[:total], partial


def _copy_out(dst, src, n):
    np.copyto(dst[:n], src[:n])
    return n


def device_to_numpy(t):
    """A caller-owned numpy uint8 array with the bytes of CUDA tensor `t`."""
    import torch
    n = int(t.numel())
    host = np.empty(n, dtype=np.uint8)
    if n == 0:
        return host
    if n <= (8 << 20):
        torch.from_numpy(host).copy_(t)
        return host
    with _use:
        return _device_to_numpy(t, host, n)


def _device_to_numpy(t, host, n):
    import torch
    ring = _ring()
    ex = _executor()
    cur = torch.cuda.current_stream()
    ring.stream.wait_stream(cur)
    for ev in ring.events:
        ev.synchronize()
    futures = [None] * N_SLOTS
    offs = list(range(0, n, SLOT_BYTES))
    inflight = []                    # (slot, off, k)
    for i, off in enumerate(offs):
        slot = i % N_SLOTS
        if futures[slot] is not None:
            futures[slot].result()                       # the slot's previous contents reached the array
        k = min(SLOT_BYTES, n - off)
        with torch.cuda.stream(ring.stream):
            ring.tensors[slot][:k].copy_(t[off:off + k], non_blocking=True)
            ring.events[slot].record(ring.stream)
        inflight.append((slot, off, k))
        # hand the previous chunk to a copier thread as soon as its transfer has landed
        if len(inflight) > 1:
            ps, po, pk = inflight.pop(0)
            ring.events[ps].synchronize()
            futures[ps] = ex.submit(_copy_out, host[po:], ring.slots[ps].array, pk)
    while inflight:
        ps, po, pk = inflight.pop(0)
        ring.events[ps].synchronize()
        futures[ps] = ex.submit(_copy_out, host[po:], ring.slots[ps].array, pk)
    for f in futures:
        if f is not None:
            f.result()
    return host
