#!/usr/bin/env python
"""Decode-free host->HBM copy scaling: what the box's fabric gives N GPUs that all pull pinned pages at once.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29611 \
        tools/h2d_scale.py [--mib 1024] [--reps 5] [--out gpurun_out/h2d_scale_nN.json]

bench.py's end-to-end number is PCIe-bound: every step copies one segment (about 1 GiB of pages) from pinned host
memory to the GPU, and SCALE_r01 showed that copy slowing down from 55.6 GB/s per GPU (N <= 2) to 24.3 GB/s at
N = 8.  This tool separates "what the fabric allows" from "what our copies do": nothing but cudaMemcpyAsync, one
process per GPU, all ranks started together, every variant timed with CUDA events on the copy stream(s):

    portable        cudaHostAlloc(cudaHostAllocPortable)                          (what cds_host_alloc does)
    write_combined  cudaHostAlloc(Portable | WriteCombined)
    interleave      portable, pages placed round-robin over the NUMA nodes (set_mempolicy(MPOL_INTERLEAVE) before the
                    first touch)
    local_node      portable, pages bound to the NUMA node `nvidia-smi topo -m` / sysfs names for this GPU
    pinned_cores    portable, the touching thread restricted to this rank's own slice of the cores before the first touch
    chunks_x2/x4    portable, the buffer split in 2 / 4 parts copied concurrently on 2 / 4 streams
    chunked_16m     portable, one stream, one cudaMemcpyAsync per 16 MiB (hostio.py's chunk size)

set_mempolicy is called through libc's syscall(); no libnuma is needed.  One JSON line on stdout (rank 0).
"""
import argparse
import ctypes
import json
import os
import re
import subprocess
import sys
import time

import torch
import torch.distributed as dist

MPOL_DEFAULT, MPOL_BIND, MPOL_INTERLEAVE = 0, 2, 3
SYS_SET_MEMPOLICY = 238                       # x86_64
cudaHostAllocPortable, cudaHostAllocWriteCombined = 1, 4
cudaMemcpyHostToDevice = 1


def cudart():
    for name in ("libcudart.so.12", "libcudart.so"):
        for d in (os.path.join(os.path.dirname(torch.__file__), "lib"), "/usr/local/cuda/lib64", ""):
            try:
                return ctypes.CDLL(os.path.join(d, name) if d else name)
            except OSError:
                pass
    # torch wheels keep the runtime under nvidia/cuda_runtime
    import glob
    for p in glob.glob(os.path.join(os.path.dirname(os.path.dirname(torch.__file__)), "nvidia", "cuda_runtime", "lib", "libcudart.so*")):
        try:
            return ctypes.CDLL(p)
        except OSError:
            pass
    raise SystemExit("libcudart not found")


def numa_nodes():
    try:
        return sorted(int(d[4:]) for d in os.listdir("/sys/devices/system/node") if re.fullmatch(r"node\d+", d))
    except OSError:
        return [0]


def set_mempolicy(mode, nodes):
    libc = ctypes.CDLL("libc.so.6", use_errno=True)
    mask = ctypes.c_ulong(0)
    for n in nodes:
        mask.value |= 1 << n
    rc = libc.syscall(SYS_SET_MEMPOLICY, ctypes.c_int(mode), ctypes.byref(mask), ctypes.c_ulong(8 * ctypes.sizeof(mask) + 1))
    return rc == 0, (ctypes.get_errno() if rc else 0)


def gpu_node(local):
    """NUMA node of GPU `local`: sysfs first, `nvidia-smi topo -m` when sysfs says -1."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from cerebro_ds_kpgi_b200 import hostnuma
    return hostnuma.gpu_numa_node(local), getattr(hostnuma, "topo_report", lambda: None)()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29611")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(local)
    rt = cudart()
    rt.cudaHostAlloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t, ctypes.c_uint]
    rt.cudaFreeHost.argtypes = [ctypes.c_void_p]
    rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
    nbytes = args.mib << 20
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    streams = [torch.cuda.Stream() for _ in range(4)]
    nodes = numa_nodes()
    my_node, topo = gpu_node(local)
    all_cores = sorted(os.sched_getaffinity(0))
    per = max(1, len(all_cores) // world)
    my_cores = set(all_cores[rank * per:(rank + 1) * per]) or set(all_cores)

    def alloc(flags, policy=None, cores=None):
        note = ""
        if policy is not None:
            ok, err = set_mempolicy(*policy)
            note = "set_mempolicy ok" if ok else "set_mempolicy errno %d" % err
        if cores is not None:
            os.sched_setaffinity(0, cores)
        p = ctypes.c_void_p()
        rc = rt.cudaHostAlloc(ctypes.byref(p), nbytes, flags)
        if rc != 0:
            raise SystemExit("cudaHostAlloc failed: %d" % rc)
        ctypes.memset(p, 0x5A, nbytes)                                      # first touch places the pages
        if policy is not None:
            set_mempolicy(MPOL_DEFAULT, [])
        if cores is not None:
            os.sched_setaffinity(0, set(all_cores))
        return p, note

    def copy_once(p, parts, piece=None):
        """Enqueue one whole-buffer copy: `parts` concurrent streams, optionally one call per `piece` bytes."""
        step = nbytes // parts
        for k in range(parts):
            st = streams[k]
            lo, hi = k * step, (nbytes if k == parts - 1 else (k + 1) * step)
            chunk = piece or (hi - lo)
            o = lo
            while o < hi:
                n = min(chunk, hi - o)
                rc = rt.cudaMemcpyAsync(ctypes.c_void_p(dev.data_ptr() + o), ctypes.c_void_p(p.value + o), n, cudaMemcpyHostToDevice,
                                        ctypes.c_void_p(st.cuda_stream))
                assert rc == 0, rc
                o += n

    def measure(p, parts=1, piece=None):
        copy_once(p, parts, piece)
        torch.cuda.synchronize()
        dist.barrier()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(parts)]
        for k in range(parts):
            evs[k][0].record(streams[k])
        for _ in range(args.reps):
            copy_once(p, parts, piece)
        for k in range(parts):
            evs[k][1].record(streams[k])
        torch.cuda.synchronize()
        ms = max(a.elapsed_time(b) for a, b in evs)
        dist.barrier()
        return args.reps * nbytes / (ms / 1000.0) / 1e9

    variants = [
        ("portable", dict(flags=cudaHostAllocPortable), {}),
        ("write_combined", dict(flags=cudaHostAllocPortable | cudaHostAllocWriteCombined), {}),
        ("interleave", dict(flags=cudaHostAllocPortable, policy=(MPOL_INTERLEAVE, nodes)), {}),
        ("local_node", dict(flags=cudaHostAllocPortable, policy=(MPOL_BIND, [my_node if my_node is not None else nodes[rank % len(nodes)]])), {}),
        ("pinned_cores", dict(flags=cudaHostAllocPortable, cores=my_cores), {}),
        ("chunks_x2", dict(flags=cudaHostAllocPortable), dict(parts=2)),
        ("chunks_x4", dict(flags=cudaHostAllocPortable), dict(parts=4)),
        ("chunked_16m", dict(flags=cudaHostAllocPortable), dict(piece=16 << 20)),
    ]
    results, notes = {}, {}
    for name, akw, mkw in variants:
        p, note = alloc(**akw)
        gbs = measure(p, **mkw)
        rt.cudaFreeHost(p)
        t = torch.tensor([gbs], dtype=torch.float64)
        lst = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(lst, t)
        per_rank = [float(x.item()) for x in lst]
        results[name] = {"per_gpu_gb_per_s_min": min(per_rank), "per_gpu_gb_per_s_mean": sum(per_rank) / world,
                         "aggregate_gb_per_s": sum(per_rank), "per_rank": [round(x, 2) for x in per_rank]}
        if note:
            notes[name] = note
    if rank == 0:
        line = {"tool": "h2d_scale", "n_gpus": world, "mib_per_gpu": args.mib, "reps": args.reps, "numa_nodes": nodes,
                "gpu0_numa_node": my_node, "topo": topo, "cores": len(all_cores), "variants": results, "notes": notes,
                "how": "cudaMemcpyAsync pinned host -> HBM only, one process per GPU, all ranks between barriers, CUDA events on the copy streams"}
        text = json.dumps(line)
        if args.out:
            with open(args.out, "w") as f:
                f.write(text + "\n")
        print(text)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
