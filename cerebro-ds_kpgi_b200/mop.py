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
import ctypes
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

    One rank = one GPU = one resident partition.  `location[m]` is the rank that holds model m (rank m % world
    before the first round).  `states` maps model -> flat float32 tensor for the models this rank currently holds.

    Transports:
      "p2p"   `torch.distributed` batch_isend_irecv of the flat buffer (NCCL on GPUs, gloo in the CPU tests);
              receive buffers come from a per-size free list, so a steady run allocates nothing.
      "symm"  GPUs of one node: every rank owns an inbox of `n_models + 1` state-sized slots in symmetric memory
              (`torch.distributed._symmetric_memory`), mapped into every peer.  A hop is ONE copy kernel
              (libcds_da `cds_peer_copy`: plain 16-byte stores over NVLink into the destination's free slot)
              followed by the symmetric-memory barrier; the model then trains in place in that slot.  Which slot is
              free is bookkeeping every rank replays identically from the schedule, so no message is needed.
      "auto"  "symm" on CUDA devices when the rendezvous succeeds, else "p2p" (`transport_note` says why).
    """

    def __init__(self, rank, world_size, state_numel, device=None, schedule=None, group=None, transport="auto", p2p_chunks=1):
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
        self.p2p_chunks = max(1, int(p2p_chunks))
        self._torch = torch
        self._free = {}                                                   # p2p: numel -> [spare receive buffers]
        self.transport, self.transport_note = "p2p", ""
        is_cuda = str(self.device).startswith("cuda")
        if transport == "symm" or (transport == "auto" and is_cuda and world_size > 1):
            try:
                self._init_symm()
                self.transport = "symm"
            except Exception as e:  # noqa
                if transport == "symm":
                    raise
                self.transport_note = "symmetric memory unavailable (%s): torch.distributed send/recv" % str(e).splitlines()[0][:160]

    # ---- symmetric-memory transport ---------------------------------------------------------------
    def _init_symm(self):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        torch = self._torch
        self._L = _lib.require_device()
        self._slot_numel = (max(self.numel.values()) + 63) & ~63          # slots stay 256-byte aligned
        n_slots = len(self.numel) + 1
        inbox = symm.empty(n_slots * self._slot_numel, dtype=torch.float32, device=self.device)
        group = self.group if self.group is not None else dist.group.WORLD
        try:
            hdl = symm.rendezvous(inbox, group)
        except Exception:  # noqa  (older releases want the group enabled first)
            symm.enable_symm_mem_for_group(group.group_name)
            hdl = symm.rendezvous(inbox, group)
        self._inbox, self._hdl = inbox, hdl
        self._peer_ptr = [int(p) for p in hdl.buffer_ptrs]
        # slot bookkeeping, replayed identically on every rank
        self._slot_of = {}
        self._free_slots = {r: list(range(n_slots)) for r in range(self.world)}
        for m in sorted(self.numel):
            self._slot_of[m] = self._free_slots[self.location[m]].pop(0)

    def _slot_view(self, slot, numel):
        o = slot * self._slot_numel
        return self._inbox[o:o + numel]

    def add_local_model(self, model, flat):
        assert self.location[model] == self.rank and flat.dtype == self._torch.float32
        if self.transport == "symm":
            view = self._slot_view(self._slot_of[model], self.numel[model])
            view.copy_(flat)
            flat = view
        self.states[model] = flat

    def model_for_round(self, r):
        """Model this rank trains in round r (None: idle)."""
        return self.schedule[r].get(self.rank)

    def _moves(self, r):
        target = {m: p for p, m in self.schedule[r].items()}
        return target, [(m, self.location[m], target[m]) for m in sorted(target) if self.location[m] != target[m]]

    def hop_to(self, r):
        """Bring every model to the rank round r needs it on.  Returns the model this rank now trains."""
        target, moves = self._moves(r)
        if moves:
            if self.transport == "symm":
                self._hop_symm(moves)
            else:
                self._hop_p2p(moves)
        for m, p in target.items():
            self.location[m] = p
        self.round = r
        return self.schedule[r].get(self.rank)

    def _hop_symm(self, moves):
        from . import _lib
        torch = self._torch
        stream = torch.cuda.current_stream()
        sp = ctypes.c_void_p(stream.cuda_stream)
        freed, arrived = [], []
        for m, src, dst in moves:
            slot = self._free_slots[dst].pop(0)
            if src == self.rank:
                cur = self.states.pop(m)
                _lib.check(self._L.cds_peer_copy(ctypes.c_void_p(cur.data_ptr()),
                                                 ctypes.c_void_p(self._peer_ptr[dst] + 4 * slot * self._slot_numel),
                                                 4 * self.numel[m], sp))
                self.bytes_moved += 4 * self.numel[m]
            if dst == self.rank:
                arrived.append((m, slot))
            freed.append((src, self._slot_of[m]))
            self._slot_of[m] = slot
        self._hdl.barrier(channel=0)                                      # every copy of this hop has landed
        for src, slot in freed:
            self._free_slots[src].append(slot)                            # reusable from the next hop on
        for m, slot in arrived:
            self.states[m] = self._slot_view(slot, self.numel[m])

    def _hop_p2p(self, moves):
        import torch.distributed as dist
        torch = self._torch
        ops, arriving, leaving = [], {}, []
        for m, src, dst in moves:
            if src == self.rank:
                buf = self.states[m]
                leaving.append(m)
                fn, peer = dist.isend, dst
            elif dst == self.rank:
                spare = self._free.get(self.numel[m])
                buf = spare.pop() if spare else torch.empty(self.numel[m], dtype=torch.float32, device=self.device)
                arriving[m] = buf
                fn, peer = dist.irecv, src
            else:
                continue
            k = min(self.p2p_chunks, max(1, buf.numel()))
            step = -(-buf.numel() // k)
            for c in range(k):
                part = buf[c * step:(c + 1) * step]
                if part.numel():
                    ops.append(dist.P2POp(fn, part, peer, self.group))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        for m in leaving:
            self.bytes_moved += 4 * self.numel[m]
            gone = self.states.pop(m)
            self._free.setdefault(gone.numel(), []).append(gone)          # becomes a receive buffer of a later hop
        self.states.update(arriving)

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
