"""The allocation-free pglz entry point (cds_pglz_decompress_batch_static) against the regular one.

Who needs it: callers that may not touch the CUDA allocator while decoding — a loader thread next to
a training framework that owns the device memory pool, a latency-sensitive ingest loop, a process
with a fixed memory budget.  The entry point keeps its work area in an array compiled into the
library, so it is limited to CDS_STATIC_MAX_DATUMS datums per call and all calls share that array.

This file is named test_zz_* so that pytest collects it last: a run that stops at the first failure has
then already run every other test.  It must pass alone in a fresh process and after the other tests.
"""
import ctypes
import os
import struct

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

MAX_DATUMS = 4096            # CDS_STATIC_MAX_DATUMS in include/cds_da.h


def _pack(datums):
    """Datums back to back (exact source offsets), outputs in 16-byte slots."""
    src_off = np.zeros(len(datums) + 1, dtype=np.uint64)
    dst_off = np.zeros(len(datums) + 1, dtype=np.uint64)
    for i, d in enumerate(datums):
        raw = struct.unpack_from("<I", d, 4)[0] if len(d) >= 8 else 0
        src_off[i + 1] = src_off[i] + len(d)
        dst_off[i + 1] = dst_off[i] + ((raw + 15) & ~15)
    blob = np.frombuffer(b"".join(datums) + b"\0" * 16, dtype=np.uint8).copy()
    return blob, src_off, dst_off


def _run_both(torch, L, datums):
    """Decode `datums` with the allocating entry point and with the static one; return both results."""
    from cerebro_ds_kpgi_b200 import _lib
    p = lambda t: ctypes.c_void_p(t.data_ptr())            # noqa: E731
    blob, src_off, dst_off = _pack(datums)
    n = len(datums)
    d_src = torch.from_numpy(blob).cuda()
    total = int(dst_off[-1]) + 16
    # reference result: existing entry point (host offset arrays, allocates its work arrays itself)
    d_dst_a = torch.full((total,), 0xAA, dtype=torch.uint8, device="cuda")
    st_a = np.full(n, -1, dtype=np.int32)
    rc = L.cds_pglz_decompress_batch(0, p(d_src), src_off.ctypes.data, p(d_dst_a), dst_off.ctypes.data, n,
                                     st_a.ctypes.data, None)
    assert rc == 0
    # static entry point: every argument is a device buffer owned by the caller
    d_dst_b = torch.full((total,), 0xAA, dtype=torch.uint8, device="cuda")
    d_so = torch.from_numpy(src_off.view(np.int64)).cuda()
    d_do = torch.from_numpy(dst_off.view(np.int64)).cuda()
    d_status = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    d_first = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.cds_pglz_decompress_batch_static(0, p(d_src), p(d_so), p(d_dst_b), p(d_do), n, p(d_status), p(d_first),
                                            ctypes.c_void_p(stream))
    _lib.check(rc)
    torch.cuda.synchronize()
    return st_a, d_dst_a.cpu().numpy(), d_status.cpu().numpy(), d_dst_b.cpu().numpy(), int(d_first.item()), dst_off


def test_static_entry_point_matches_regular_one_on_two_calls_in_a_row():
    import torch
    from cerebro_ds_kpgi_b200 import _lib, synth
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    torch.cuda.set_device(0)
    L = _lib.require_device()

    z = np.load(os.path.join(GOLDEN, "pglz_kat.npz"))
    so = z["src_off"]
    kat = [z["src"][so[i]:so[i + 1]].tobytes() for i in range(len(so) - 1)]
    rng = np.random.default_rng(11)

    # ---- call 1: known-answer vectors with damaged streams mixed in
    first = list(kat)
    for k in (5, 60, 123, 200):
        b = bytearray(kat[k] if len(kat[k]) > 12 else kat[150])
        b[len(b) // 2] ^= 0x5A
        first.insert(k, bytes(b))
    first.insert(3, b"\x40\x00\x00\x05\x01")                              # shorter than a header
    st_a, out_a, st_b, out_b, first_bad, dst_off = _run_both(torch, L, first)
    assert np.array_equal(st_a, st_b)
    assert (st_a != 0).any() and (st_a == 0).any()
    for i in np.nonzero(st_a == 0)[0]:
        raw = struct.unpack_from("<I", first[i], 4)[0]
        o = int(dst_off[i])
        assert out_a[o:o + raw].tobytes() == out_b[o:o + raw].tobytes(), i
    assert first_bad == int(np.nonzero(st_a != 0)[0][0])

    # ---- call 2, straight after: fewer datums, all intact, different bytes.  Nothing of call 1 may
    # show through: the error word has to read "none" again and the job slots beyond n are ignored.
    second = [synth.pglz_varlena(rng.integers(0, 5, size=int(m), dtype=np.uint8), "lz") for m in rng.integers(1, 20000, 37)]
    st_a, out_a, st_b, out_b, first_bad, dst_off = _run_both(torch, L, second)
    assert np.array_equal(st_a, st_b) and not st_b.any()
    assert out_a.tobytes() == out_b.tobytes()
    assert first_bad == -1

    # ---- call 3: the known-answer vectors alone, against the bytes BOTH reference decoders produced for them
    # (tests/golden/pglz_kat.npz `raw`), not only against the other entry point
    st_a, out_a, st_b, out_b, first_bad, dst_off = _run_both(torch, L, kat)
    assert not st_b.any() and first_bad == -1
    ro = z["raw_off"]
    for i in range(len(kat)):
        want = z["raw"][ro[i]:ro[i + 1]].tobytes()
        o = int(dst_off[i])
        assert out_b[o:o + len(want)].tobytes() == want, i

    # ---- size limit: refused before anything is enqueued, with a message that names the limit
    dummy = torch.zeros(64, dtype=torch.uint8, device="cuda")
    offs = torch.zeros(MAX_DATUMS + 2, dtype=torch.int64, device="cuda")
    stat = torch.zeros(MAX_DATUMS + 1, dtype=torch.int32, device="cuda")
    p = lambda t: ctypes.c_void_p(t.data_ptr())            # noqa: E731
    rc = L.cds_pglz_decompress_batch_static(0, p(dummy), p(offs), p(dummy), p(offs), MAX_DATUMS + 1, p(stat), None, None)
    assert rc == _lib.CDS_ERR_CAPACITY
    msg = L.cds_last_error().decode()
    assert str(MAX_DATUMS) in msg and "CDS_STATIC_MAX_DATUMS" in msg
    # exactly the maximum is accepted (empty datums: every status is "corrupt", nothing is written)
    rc = L.cds_pglz_decompress_batch_static(0, p(dummy), p(offs), p(dummy), p(offs), MAX_DATUMS, p(stat), None, None)
    assert rc == 0
    torch.cuda.synchronize()
    assert int((stat[:MAX_DATUMS] == _lib.CDS_ERR_PGLZ_CORRUPT).sum()) == MAX_DATUMS
