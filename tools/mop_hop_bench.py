#!/usr/bin/env python
"""MOP model-hopper run (BASELINE config 5) on N GPUs of one node:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/mop_hop_bench.py

Each rank keeps one decoded partition resident (cds_decode_device), starts with one model state of
ResNet-50 size (25.6 M float32 = 102 MB flat vector, the serialize_nd_weights layout), and runs epochs
of the ring schedule: train a sub-epoch (stand-in: one pass over the partition's labels + an axpy on the
state), hop the state to the next rank with NCCL send/recv, while the next decode of the partition
(ingest of the following epoch) runs on a side stream.  Prints one JSON line on rank 0.
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import bench
    bench.REAL_STDOUT = bench.protect_stdout()      # NCCL prints its version banner on fd 1: the JSON line goes to the saved descriptor
    from cerebro_ds_kpgi_b200 import mop, segment
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    seg, meta = bench.make_workload("imagenet_raw", 2018 + rank, 256)
    ds = segment.DeviceSegment(seg.heap.size // 32768, seg.toast.size // 32768, device=local)
    d_heap, d_toast = torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda()
    out_ind = torch.empty(meta["ind_bytes"], dtype=torch.uint8, device="cuda")
    out_dep = torch.empty(meta["dep_bytes"], dtype=torch.uint8, device="cuda")
    s = ds.decode_device(d_heap, d_toast, out_ind, out_dep, meta["isb"], meta["dsb"])
    assert s.status == 0
    labels = segment.onehot_argmax(out_dep.view(torch.int16).view(-1, meta["classes"]))
    numel = 25_557_032                                   # ResNet-50 parameters
    hop = mop.ModelHopper(rank, world, numel, device="cuda")
    hop.add_local_model(rank, torch.full((numel,), float(rank), device="cuda"))
    ingest_stream = torch.cuda.Stream()
    out_ind2, out_dep2 = torch.empty_like(out_ind), torch.empty_like(out_dep)

    def train(model, flat, rnd):
        flat.add_(float(labels.float().mean()) * 1e-6)   # stand-in for a sub-epoch over the resident partition

    hop_ms = []
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.time()
    for epoch in range(6):                                # epoch 0 warms NCCL's point-to-point channels up
        with torch.cuda.stream(ingest_stream):            # ingest for the next epoch overlaps training + hops
            ds.decode_device(d_heap, d_toast, out_ind2, out_dep2, meta["isb"], meta["dsb"], stream=ingest_stream, sync=False)
        for r in range(len(hop.schedule)):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            m = hop.hop_to(r)
            e1.record()
            train(m, hop.states[m], r)
            torch.cuda.synchronize()
            if epoch > 0:
                hop_ms.append(e0.elapsed_time(e1))
        ingest_stream.synchronize()
        assert ds.finish(ingest_stream).status == 0 and torch.equal(out_ind, out_ind2)
    dist.barrier(); torch.cuda.synchronize()
    wall = time.time() - t0
    t = torch.tensor([sum(hop_ms) / max(1, len(hop_ms))], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = float(t.item())
        bench.emit({"workload": "mop_hop", "n_gpus": world, "state_bytes": 4 * numel, "hop_ms": ms,
                          "hop_gb_per_s_per_gpu": (4 * numel / (ms / 1000.0) / 1e9) if world > 1 and ms > 0 else None,
                          "epochs": 5, "rounds_per_epoch": len(hop.schedule), "wall_s": wall,
                          "ingest_overlapped": True, "backend": "nccl send/recv (batch_isend_irecv)"})
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
