"""N > 1 host logic on the CPU: the model hopper over gloo, world_size 2 (and the schedules, no process group)."""
import os
import socket

import numpy as np
import pytest


def test_schedules_visit_every_pair_once():
    from cerebro_ds_kpgi_b200 import mop
    for n in (1, 2, 3, 8):
        assert mop.check_schedule(mop.ring_schedule(n), n, n)
        assert len(mop.ring_schedule(n)) == n
    for nm, np_ in ((8, 8), (3, 5), (5, 3), (16, 8)):
        rounds = mop.greedy_schedule(nm, np_)
        assert mop.check_schedule(rounds, nm, np_)
        assert len(rounds) >= max(nm, np_)
        shuffled = mop.greedy_schedule(nm, np_, shuffle_seed=2018)
        assert mop.check_schedule(shuffled, nm, np_)
    # CTQ's rule with equal durations and no shuffle degenerates to a Latin square for n x n
    assert len(mop.greedy_schedule(8, 8)) == 8


def test_moves_between_rounds():
    from cerebro_ds_kpgi_b200 import mop
    rounds = mop.ring_schedule(4)
    mv = mop.moves_between(rounds[0], rounds[1])
    assert sorted(mv) == [(0, 0, 1), (1, 1, 2), (2, 2, 3), (3, 3, 0)]


def test_flat_state_round_trip():
    import torch
    from cerebro_ds_kpgi_b200 import mop
    ws = [torch.randn(3, 4), torch.randn(7), torch.randn(2, 2, 2).double()]
    flat = mop.flatten_state(ws)
    assert flat.dtype == torch.float32 and flat.numel() == 27
    # same bytes as serialize_nd_weights: np.float32(np.concatenate([w.flatten() ...])) (madlib_keras_wrapper.py:129-131)
    want = np.float32(np.concatenate([w.numpy().flatten() for w in ws]))
    assert flat.numpy().tobytes() == want.tobytes()
    back = mop.unflatten_state(flat, [(3, 4), (7,), (2, 2, 2)])
    assert [tuple(b.shape) for b in back] == [(3, 4), (7,), (2, 2, 2)]
    assert torch.equal(back[0], ws[0]) and back[1].data_ptr() == flat[12:].data_ptr()
    with pytest.raises(AssertionError):
        mop.unflatten_state(flat, [(3, 4), (7,)])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cerebro_ds_kpgi_b200 import mop
    numel = {0: 1000, 1: 1777}
    hop = mop.ModelHopper(rank, world, numel)
    mine = torch.full((numel[rank],), float(rank + 1))
    hop.add_local_model(rank, mine)
    visits = []

    def train(model, flat, rnd):
        # a sub-epoch: every partition leaves its mark on the model it hosts
        visits.append((rnd, model))
        flat += 10.0 ** rank

    for _ in range(2):                       # two epochs: models come home after each
        hop.run_epoch(train)
    held = sorted(hop.states)
    torch.save({"visits": visits, "held": held, "vals": {m: float(hop.states[m][0]) for m in held},
                "sizes": {m: hop.states[m].numel() for m in held}, "moved": hop.bytes_moved},
               os.path.join(out_dir, "r%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


def test_model_hopper_world_size_2_gloo(tmp_path):
    import torch
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = torch.load(tmp_path / "r0.pt"), torch.load(tmp_path / "r1.pt")
    # ring schedule on 2 ranks: round 0 model == rank, round 1 the other model; twice
    assert r0["visits"] == [(0, 0), (1, 1), (0, 0), (1, 1)]
    assert r1["visits"] == [(0, 1), (1, 0), (0, 1), (1, 0)]
    # after the last round (s = 1) model m sits on rank (m + 1) % 2, with its own size
    assert r0["held"] == [1] and r1["held"] == [0]
    assert r0["sizes"] == {1: 1777} and r1["sizes"] == {0: 1000}
    # every model was trained twice on each partition: +2*1 (rank 0) + 2*10 (rank 1)
    assert r1["vals"][0] == 1.0 + 22.0 and r0["vals"][1] == 2.0 + 22.0
    assert r0["moved"] > 0 and r1["moved"] > 0
