"""The model hopper on real GPUs: two ranks, two GPUs, both transports.  Skipped on a one-GPU box
(the hop leg of `bench.py --gpus N` covers the same ground in the driver's scaling run)."""
import os
import socket

import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir, transport):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from cerebro_ds_kpgi_b200 import mop
    numel = {0: 1_000_003, 1: 2_500_001, 2: 777}                 # three models on two partitions: one is always idle somewhere
    sched = mop.greedy_schedule(3, 2)                            # CTQ's rule (ctq.py:448-508)
    mop.check_schedule(sched, 3, 2)
    hop = mop.ModelHopper(rank, world, numel, device="cuda", schedule=sched, transport=transport)
    for m in numel:
        if hop.location[m] == rank:
            hop.add_local_model(m, torch.full((numel[m],), float(m + 1), device="cuda"))
    visits = []

    def train(model, flat, rnd):
        visits.append((rnd, model))
        flat += 10.0 ** rank                                     # every partition leaves its mark on the model it hosts

    for _ in range(2):
        hop.run_epoch(train)
    torch.cuda.synchronize()
    held = sorted(hop.states)
    ok = all(bool((hop.states[m] == hop.states[m][0]).all().item()) for m in held)
    torch.save({"visits": visits, "held": held, "vals": {m: float(hop.states[m][0]) for m in held}, "uniform": ok,
                "sizes": {m: hop.states[m].numel() for m in held}, "transport": hop.transport, "note": hop.transport_note},
               os.path.join(out_dir, "r%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("transport", ["symm", "p2p"])
def test_model_hopper_two_gpus(transport, tmp_path):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path), transport), nprocs=2, join=True)
    r = [torch.load(tmp_path / ("r%d.pt" % k)) for k in range(2)]
    assert r[0]["transport"] == r[1]["transport"] == transport
    # every model was trained twice on each partition: + 2 * (1 + 10); each state is one value throughout
    vals = {}
    for x in r:
        assert x["uniform"]
        vals.update(x["vals"])
        for m, n in x["sizes"].items():
            assert n == {0: 1_000_003, 1: 2_500_001, 2: 777}[m]
    assert vals == {0: 1.0 + 22.0, 1: 2.0 + 22.0, 2: 3.0 + 22.0}
    assert sorted(r[0]["held"] + r[1]["held"]) == [0, 1, 2]
    # every (round, model) visit happened exactly once per epoch over the two ranks
    per_epoch = sorted(v for x in r for v in x["visits"])
    assert len(per_epoch) == 2 * 3 * 2
