"""Host-side pieces of bench.py: the checksum the correctness gate compares against, and the split of a workload's
rows over the reference arm's worker processes."""
import numpy as np
import pytest


def direct_sum(buf):
    """sum of w_i * (2 i + 1) * SUM_MUL mod 2^64 over the 64-bit little-endian words of `buf` (tail zero-padded)."""
    import bench
    b = bytes(buf) + b"\0" * (-len(buf) % 8)
    total = 0
    for i in range(len(b) // 8):
        w = int.from_bytes(b[8 * i:8 * i + 8], "little")
        total = (total + w * ((2 * i + 1) * bench.SUM_MUL)) & bench.MASK64
    return total


def test_expected_sum_is_position_linear():
    import bench
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, 4096, dtype=np.uint8)
    b = rng.standard_normal(300).astype(np.float32)
    c = rng.integers(-5, 5, (7, 12), dtype=np.int16)
    es = bench.ExpectedSum()
    seq = [a, b, a, c, b, a]
    for x in seq:
        es.add(x)
    flat = b"".join(np.ascontiguousarray(x).tobytes() for x in seq)
    assert es.total == direct_sum(flat)
    assert es.words * 8 == len(flat)


def test_split_units_sums_exactly():
    import bench
    for units, procs in ((6971, 16), (8016, 32), (3, 16), (1024, 7), (0, 4)):
        parts = bench.split_units(units, procs)
        assert sum(parts) == units and len(parts) == procs and max(parts) - min(parts) <= 1


def test_workload_meta_accounts_for_every_byte():
    import bench
    seg, meta = bench.make_workload("imagenet_pglz", 2018, 8)
    assert meta["gather_out_bytes"] + meta["pglz_out_bytes"] == meta["ind_bytes"] + meta["dep_bytes"]
    assert 0 < meta["pglz_in_bytes"] < meta["pglz_out_bytes"]
    assert meta["labels"].shape == (meta["rows"],)
    seg, meta = bench.make_workload("imagenet_raw", 2018, 8)
    assert meta["pglz_out_bytes"] == 0 and meta["pglz_in_bytes"] == 0


def test_every_configuration_rides_along_at_every_n():
    """The default run carries configs[2], configs[3] and the packed layout beside the headline workload, at N = 1 and N = 8 alike."""
    import argparse
    import bench
    for gpus in (1, 2, 8):
        got = bench.extra_workloads(argparse.Namespace(gpus=gpus, workload="imagenet_raw"))
        assert got == ["imagenet_pglz", "criteo", "imagenet_packed"]
    assert "criteo" not in bench.extra_workloads(argparse.Namespace(gpus=1, workload="criteo"))


@pytest.mark.gpu
def test_device_checksum_and_peer_copy_match_the_host():
    import ctypes
    import torch
    import bench
    from cerebro_ds_kpgi_b200 import _lib
    L = _lib.require_device()
    rng = np.random.default_rng(3)
    p = lambda t: ctypes.c_void_p(t.data_ptr())            # noqa: E731
    out = torch.zeros(1, dtype=torch.int64, device="cuda")
    for n in (0, 1, 7, 8, 9, 4096, 150528, 1000003):
        host = rng.integers(0, 256, n, dtype=np.uint8)
        dev = torch.from_numpy(host).cuda() if n else torch.zeros(1, dtype=torch.uint8, device="cuda")
        _lib.check(L.cds_checksum64(p(dev), n, p(out), None))
        torch.cuda.synchronize()
        assert (int(out.item()) & bench.MASK64) == direct_sum(host.tobytes()), n
        # unaligned source and destination, same and different alignment
        for so, do in ((0, 0), (3, 3), (5, 12)):
            if n <= 16:
                continue
            dst = torch.full((n + 64,), 0xEE, dtype=torch.uint8, device="cuda")
            _lib.check(L.cds_peer_copy(ctypes.c_void_p(dev.data_ptr() + so), ctypes.c_void_p(dst.data_ptr() + do), n - so, None))
            torch.cuda.synchronize()
            got = dst.cpu().numpy()
            assert got[do:do + n - so].tobytes() == host[so:].tobytes() and (got[:do] == 0xEE).all() and (got[do + n - so:] == 0xEE).all()
