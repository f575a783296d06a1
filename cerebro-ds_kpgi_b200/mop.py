"""Model hopper (MOP) over resident decoded partitions — SURVEY.md §8 f-1, BASELINE config 5.

In the reference the hopping logic lives in external Cerebro / MADlib code; what is in the tree is
* the greedy `(model, partition)` scheduler of CTQ (/root/reference/cerebro_gpdb/ctq.py:247-261, 448-508):
  an idle partition takes the first remaining pair whose model is idle; a model trains on one
  partition at a time and visits every partition once per epoch;
* the model state format: one flat float32 vector, tensors flattened and concatenated in order
  (/root/reference/cerebro_gpdb/madlib_keras_wrapper.py:119-131 `serialize_nd_weights`,
  :134-160 `deserialize_as_nd_weights`), which CTQ moves between workers as files on NFS
  (ctq.py:330-333, 404-405).

Here the partitions stay decoded in HBM (one per GPU, `da.input_fn_device`), and the model state
moves instead: `torch.distributed` point-to-point send/recv of the flat buffer — NCCL over
NVLink/NVSwitch on the GPU box, gloo in the CPU tests.  The decode path itself has no
collective; this module is the only place ranks talk to each other.
"""
import random

import numpy as np


# ---- model state: flat float32 (madlib_keras_wrapper.py:119-160) ---------------------------------

def flatten_state(tensors, out=None):
    """`serialize_nd_weights`: flatten every tensor, concatenate in order, cast to float32."""
    import torch
    n = sum(int(t.numel()) for t in tensors)
    if out is None:
        dev = tensors[0].device if tensors else "cpu"
        out = torch.empty(n, dtype=torch.float32, device=dev)
    assert out.numel() == n and out.dtype == torch.float32
    o = 0
    for t in tensors:
        k = int(t.numel())
        out[o:o + k].copy_(t.reshape(-1))
        o += k
    return out


def unflatten_state(flat, shapes):
    """`deserialize_as_nd_weights`: views of the flat vector with the given shapes (no copy)."""
    total = sum(int(np.prod(s)) for s in shapes)
    if total != int(flat.numel()):
        raise AssertionError("Number of elements in model weights({0}) doesn't match model({1}).".format(
            int(flat.numel()), total))
    out, o = [], 0
    for s in shapes:
        k = int(np.prod(s))
        out.append(flat[o:o + k].view(*s))
        o += k
    return out


# ---- schedule -----------------------------------------------------------------------------------

def greedy_schedule(n_models, n_partitions, shuffle_seed=None):
    """Rounds of an epoch under CTQ's rule (ctq.py:448-456, 489-508) with equal sub-epoch durations.

    Returns a list of rounds; each round is {partition: model}.  Every (model, partition) pair occurs
    exactly once; within a round a model is on at most one partition and vice versa.
    """
    pairs = [(m, p) for m in range(n_models) for p in range(n_partitions)]          # ctq.py:250-251
    if shuffle_seed is not None:
        random.Random(shuffle_seed).shuffle(pairs)                                    # ctq.py:252-253
    rounds = []
    while pairs:
        busy_models, this_round = set(), {}
        for p in range(n_partitions):                                                 # ctq.py:490-491
            for m, q in pairs:                                                        # _get_runnable_model, ctq.py:448-454
                if q == p and m not in busy_models:
                    this_round[p] = m
                    busy_models.add(m)
                    break
        for p, m in this_round.items():
            pairs.remove((m, p))                                                      # peek_job, ctq.py:476
        rounds.append(this_round)
    return rounds


def ring_schedule(n):
    """The conflict-free schedule for n models on n equal partitions: in round s model m trains on
    partition (m + s) mod n, so every hop is 'send to the right neighbour'."""
    return [{(m + s) % n: m for m in range(n)} for s in range(n)]


def moves_between(prev_round, next_round):
    """Model transfers needed between two rounds: list of (model, src_partition, dst_partition)."""
    where = {m: p for p, m in prev_round.items()}
    out = []
    for p, m in sorted(next_round.items()):
        src = where.get(m)
        if src is not None and src != p:
            out.append((m, src, p))
    return out


# ---- the hopper -----------------------------------------------------------------------------------

class ModelHopper:
    """Moves flat model states between ranks according to a schedule.

    One rank = one GPU = one resident partition.  `home[m]` is the rank that holds model m before the
    first round (rank m by default).  `states` maps model -> flat float32 tensor for the models this rank
    currently holds; buffers of arriving models are allocated once per model size and reused.
    """

    def __init__(self, rank, world_size, state_numel, device=None, schedule=None, group=None):
        import torch
        self.rank, self.world = rank, world_size
        self.numel = dict(state_numel) if isinstance(state_numel, dict) else {m: int(state_numel) for m in range(world_size)}
        self.device = device if device is not None else "cpu"
        self.schedule = schedule if schedule is not None else ring_schedule(world_size)
        self.group = group
        self.states = {}
        self.location = {m: m % world_size for m in self.numel}          # model -> rank, before round 0
        self.round = -1
        self.bytes_moved = 0
        self._torch = torch

    def add_local_model(self, model, flat):
        assert self.location[model] == self.rank and flat.dtype == self._torch.float32
        self.states[model] = flat

    def model_for_round(self, r):
        """Model this rank trains in round r (None: idle)."""
        return self.schedule[r].get(self.rank)

    def hop_to(self, r):
        """Bring every model to the rank round r needs it on.  Returns the model this rank now trains."""
        import torch.distributed as dist
        torch = self._torch
        target = {m: p for p, m in self.schedule[r].items()}
        ops, arriving, leaving = [], {}, []
        for m in sorted(target):
            src, dst = self.location[m], target[m]
            if src == dst:
                continue
            if src == self.rank:
                ops.append(dist.P2POp(dist.isend, self.states[m], dst, self.group))
                leaving.append(m)
            elif dst == self.rank:
                buf = torch.empty(self.numel[m], dtype=torch.float32, device=self.device)
                arriving[m] = buf
                ops.append(dist.P2POp(dist.irecv, buf, src, self.group))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        for m in leaving:
            self.bytes_moved += 4 * self.numel[m]
            del self.states[m]
        self.states.update(arriving)
        for m, p in target.items():
            self.location[m] = p
        self.round = r
        return self.schedule[r].get(self.rank)

    def run_epoch(self, train_fn):
        """`train_fn(model, flat_state, round)` is called once per round on the model resident here;
        it trains in place on this rank's partition (one sub-epoch)."""
        for r in range(len(self.schedule)):
            m = self.hop_to(r)
            if m is not None:
                train_fn(m, self.states[m], r)


def check_schedule(rounds, n_models, n_partitions):
    """Invariants of an epoch (used by the tests and by callers that bring their own schedule)."""
    seen = set()
    for rd in rounds:
        models = list(rd.values())
        assert len(models) == len(set(models)), "a model is on two partitions in one round"
        for p, m in rd.items():
            assert 0 <= p < n_partitions and 0 <= m < n_models
            assert (m, p) not in seen, "a (model, partition) pair is visited twice"
            seen.add((m, p))
    assert len(seen) == n_models * n_partitions, "an epoch must visit every (model, partition) pair"
    return True
