"""GPU edge cases and full-size properties (empty / ragged inputs, maximum sizes, BASELINE sizes)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    return torch


def decode(torch, heap, toast, **kw):
    from cerebro_ds_kpgi_b200 import segment
    h = torch.from_numpy(np.ascontiguousarray(heap)).cuda() if heap.size else torch.zeros(0, dtype=torch.uint8, device="cuda")
    t = torch.from_numpy(np.ascontiguousarray(toast)).cuda() if toast.size else torch.zeros(0, dtype=torch.uint8, device="cuda")
    return segment.decode_pages(h, t, **kw)


def test_empty_relations(torch_cuda):
    from cerebro_ds_kpgi_b200 import segment
    torch = torch_cuda
    ds = segment.DeviceSegment(1, 1)
    empty = torch.zeros(16, dtype=torch.uint8, device="cuda")
    out = torch.zeros(16, dtype=torch.uint8, device="cuda")
    # zero pages of either relation (numel // 32768 == 0)
    s = ds.decode_device(empty, empty, out, out)
    assert s.status == 0 and s.n_rows == 0 and s.n_chunks == 0 and s.n_positions == 0 and s.ind_bytes == 0
    # a page with a valid header and no tuples
    from cerebro_ds_kpgi_b200 import synth
    blank = synth.RelWriter(1, 3)
    blank.toast_value(1, np.zeros(10, np.uint8))           # forces page 0 to exist ...
    pg = blank.pages().copy()
    pg[12:14] = np.frombuffer(np.uint16(24).tobytes(), np.uint8)   # ... then drop its ItemIds: pd_lower = 24
    d = torch.from_numpy(pg).cuda()
    s = ds.decode_device(d, d, out, out)
    assert s.status == 0 and s.n_rows == 0 and s.n_chunks == 0
    ds.close()


def test_rows_without_chunks_are_absent_like_in_the_reference(torch_cuda, coracle):
    """TOAST file truncated to zero pages: external-only rows vanish (the reference's dict never gets them),
    rows with an inline value surface as a missing independent_var."""
    from cerebro_ds_kpgi_b200 import synth
    import da_oracle
    ind, dep = synth.imagenet_rows(4, seed=2, shape=(8, 8, 3), classes=4)
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="ext_pglz")
    status, part = decode(torch_cuda, seg.heap, seg.toast[:0])
    assert status == 0 and part.summary.n_rows == 4 and part.summary.n_positions == 0
    assert coracle.decode(seg.heap, seg.toast[:0])["order"].size == 0
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="inline_pglz")
    status, _ = decode(torch_cuda, seg.heap, seg.toast[:0])
    with pytest.raises(da_oracle.OracleError) as ei:
        coracle.decode(seg.heap, seg.toast[:0])
    assert status == ei.value.code == da_oracle.ERR_MISSING_VALUE


def test_unreferenced_and_foreign_chunks_are_ignored(torch_cuda, coracle):
    """Chunks whose chunk_id no pointer names (dead tuples, other tables' values) are skipped (ppr:516)."""
    from cerebro_ds_kpgi_b200 import synth
    ind, dep = synth.imagenet_rows(5, seed=3, shape=(20, 20, 3), classes=6, kind="normal")
    a = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="ext_raw", first_valueid=20000)
    b = synth.build_segment(ind[::-1], dep[::-1], ind_storage="ext_pglz", dep_storage="ext_raw", first_valueid=90000)
    mixed = np.concatenate([b.toast, a.toast, b.toast])
    want = coracle.decode(a.heap, mixed)
    status, part = decode(torch_cuda, a.heap, mixed)
    assert status == 0
    assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()
    # several small values share a page and the walk is by ascending address, so the result order is the
    # reference's first-seen order, not table order
    assert part.ind.cpu().numpy().tobytes() == b"".join(ind[r].tobytes() for r in want["order"])
    assert sorted(want["order"]) == list(range(5))


def test_ragged_value_sizes_and_odd_lengths(torch_cuda, coracle):
    """Buffers of different row counts, values that are not multiples of 16/4 bytes, 1-chunk and multi-chunk values,
    shuffled physical order."""
    from cerebro_ds_kpgi_b200 import synth
    rng = np.random.default_rng(12)
    ind, dep = [], []
    for i in range(40):
        b = int(rng.integers(1, 6))
        ind.append(rng.standard_normal((b, 7 + 9 * (i % 4), 11, 3)).astype(np.float32))
        dep.append(rng.integers(-3, 3, size=(b, 5 + (i % 3))).astype(np.int16))
    st_i = [["ext_raw", "ext_pglz", "ext_literal"][i % 3] for i in range(40)]
    st_d = [["ext_raw", "inline_pglz", "ext_pglz", "inline_literal"][i % 4] for i in range(40)]
    n_ext = 40 + sum(1 for s in st_d if s.startswith("ext"))
    seg = synth.build_segment(ind, dep, ind_storage=st_i, dep_storage=st_d,
                              toast_perm=list(rng.permutation(n_ext)), buffer_ids=list(rng.permutation(1000)[:40]))
    want = coracle.decode(seg.heap, seg.toast)
    status, part = decode(torch_cuda, seg.heap, seg.toast)
    assert status == 0
    assert np.array_equal(part.layout["order"], want["order"])
    assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()


def test_many_chunks_per_page_beyond_stage_cache(torch_cuda, coracle):
    """> 48 chunks on one TOAST page: the gather falls back from its shared-memory descriptor cache to global reads."""
    from cerebro_ds_kpgi_b200 import synth
    rng = np.random.default_rng(5)
    ind = [rng.standard_normal((1, 5, 5, 3)).astype(np.float32) for _ in range(400)]
    dep = [np.eye(3, dtype=np.int16)[[i % 3]] for i in range(400)]
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="ext_raw")
    assert 800 / (seg.toast.size // 32768) > 100            # far more tuples per page than the descriptor cache holds
    want = coracle.decode(seg.heap, seg.toast)
    status, part = decode(torch_cuda, seg.heap, seg.toast)
    assert status == 0
    assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()


def test_capacity_errors(torch_cuda):
    from cerebro_ds_kpgi_b200 import segment, synth, _lib
    torch = torch_cuda
    rng = np.random.default_rng(5)
    ind = [rng.standard_normal((1, 5, 5, 3)).astype(np.float32) for _ in range(300)]
    dep = [np.eye(3, dtype=np.int16)[[i % 3]] for i in range(300)]
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="ext_raw")
    h, t = torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda()
    out = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    # chunk directory too small
    ds = segment.DeviceSegment(seg.heap.size // 32768, seg.toast.size // 32768, max_chunks=50)
    with pytest.raises(_lib.CdsError) as ei:
        ds.decode_device(h, t, out, out)
    assert ei.value.code == _lib.CDS_ERR_CAPACITY
    ds.close()
    # no staging area although values are compressed
    ds = segment.DeviceSegment(seg.heap.size // 32768, seg.toast.size // 32768, scratch_bytes=0)
    with pytest.raises(_lib.CdsError) as ei:
        ds.decode_device(h, t, out, out)
    assert ei.value.code == _lib.CDS_ERR_CAPACITY
    ds.close()
    # more pages than the handle was created for: refused before anything is enqueued
    ds = segment.DeviceSegment(1, 1)
    big = torch.zeros(4 * 32768, dtype=torch.uint8, device="cuda")
    with pytest.raises(_lib.CdsError):
        ds.decode_device(h, big, out, out)
    ds.close()


def test_handle_reuse_across_different_segments(torch_cuda, coracle):
    """One handle, several decodes of different content back to back (state, scans and hash table are reset)."""
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    ds = segment.DeviceSegment(4, 64)
    for seed, st in ((1, "ext_raw"), (2, "ext_pglz"), (3, "ext_literal"), (4, "ext_raw")):
        ind, dep = synth.imagenet_rows(6, seed=seed, shape=(30, 30, 3), classes=9, kind="normal" if seed % 2 else "norm_u8")
        seg = synth.build_segment(ind, dep, ind_storage=st, dep_storage="inline_pglz" if seed % 2 else "ext_raw",
                                  first_valueid=20000 + 7 * seed)
        want = coracle.decode(seg.heap, seg.toast)
        out_i = torch.zeros(want["out_ind"].size, dtype=torch.uint8, device="cuda")
        out_d = torch.zeros(want["out_dep"].size, dtype=torch.uint8, device="cuda")
        s = ds.decode_device(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda(), out_i, out_d)
        assert s.status == 0 and s.n_rows == 6
        assert out_i.cpu().numpy().tobytes() == want["out_ind"].tobytes()
        assert out_d.cpu().numpy().tobytes() == want["out_dep"].tobytes()
    ds.close()


def test_full_size_segment_properties(torch_cuda):
    """BASELINE configs[1] at full size (1 GiB TOAST segment): the decoded tensor must be the generator's rows,
    in table order, byte for byte (for uncompressed values the output IS the concatenation of the sources);
    labels must equal argmax of the one-hot rows; decoding twice is idempotent."""
    import bench
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    n = synth.rows_for_toast_bytes(1 << 30)
    ind, dep = synth.imagenet_rows(n, seed=2018, kind="norm_u8")
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="ext_raw", keep_expected=False)
    assert seg.toast.size >= (1 << 30) * 0.97
    ds = segment.DeviceSegment(seg.heap.size // 32768, seg.toast.size // 32768)
    out_i = torch.empty(n * bench.IMAGENET_ROW_BYTES, dtype=torch.uint8, device="cuda")
    out_d = torch.empty(n * bench.LABEL_ROW_BYTES, dtype=torch.uint8, device="cuda")
    h, t = torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda()
    s = ds.decode_device(h, t, out_i, out_d, bench.IMAGENET_ROW_BYTES, bench.LABEL_ROW_BYTES)
    assert s.status == 0 and s.n_rows == n and s.n_positions == n and s.ind_bytes == n * bench.IMAGENET_ROW_BYTES
    src = torch.from_numpy(np.concatenate([x.reshape(-1) for x in ind])).cuda()
    assert torch.equal(out_i.view(torch.float32), src)
    lab = segment.onehot_argmax(out_d.view(torch.int16).view(n, 1000))
    want_lab = torch.from_numpy(np.concatenate([np.argmax(d, axis=1) for d in dep])).cuda()
    assert torch.equal(lab, want_lab)
    first = out_i.clone()
    s = ds.decode_device(h, t, out_i, out_d, bench.IMAGENET_ROW_BYTES, bench.LABEL_ROW_BYTES)
    assert s.status == 0 and torch.equal(first, out_i)
    ds.close()


def test_full_size_pglz_round_trip(torch_cuda):
    """configs[2] shape at a few hundred MiB: encode -> pages -> decode must return the source rows."""
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    a_ind, dep = synth.imagenet_rows(1500, seed=2018, kind="norm_u8")
    b_ind, _ = synth.imagenet_rows(1500, seed=2019, kind="normal")
    ind = [a_ind[i] if i % 2 else b_ind[i] for i in range(1500)]
    st = ["ext_pglz" if i % 2 else "ext_literal" for i in range(1500)]
    seg = synth.build_segment(ind, dep, ind_storage=st, dep_storage="inline_pglz", keep_expected=False, effort=2)
    status, part = decode(torch, seg.heap, seg.toast, ind_sample_bytes=150528, dep_sample_bytes=2000)
    assert status == 0
    src = torch.from_numpy(np.concatenate([x.reshape(-1) for x in ind])).cuda()
    assert torch.equal(part.ind.view(torch.float32), src)
    assert torch.equal(part.dep.view(torch.int16), torch.from_numpy(np.concatenate([d.reshape(-1) for d in dep])).cuda())


def test_repeated_chunk_seq_is_refused_not_overlaid(torch_cuda):
    """Two chunks of one value with the same chunk_seq (and the right count and byte total): the reference would join
    them in sorted order; the gather would drop both onto the same bytes, so the decode reports "unexpected chunk size"."""
    import struct
    from cerebro_ds_kpgi_b200 import synth, _lib
    ind, dep = synth.imagenet_rows(2, seed=5, shape=(60, 60, 3), classes=4, kind="normal")       # 43 200 B: 6 chunks
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="inline_pglz")
    status, _ = decode(torch_cuda, seg.heap, seg.toast)
    assert status == 0
    toast = seg.toast.copy()
    iid = struct.unpack_from("<I", toast, 24 + 4 * 2)[0]           # third inserted tuple: chunk_seq 2 of the first value
    off = iid & 0x7FFF
    assert struct.unpack_from("<I", toast, off + 28)[0] == 2
    struct.pack_into("<I", toast, off + 28, 1)                     # now chunk_seq 1 appears twice, 2 never
    status, _ = decode(torch_cuda, seg.heap, toast)
    assert status == _lib.CDS_ERR_CHUNK_SIZE


def test_one_call_decode_with_a_heap_of_more_than_1020_pages(torch_cuda, coracle):
    """The one-call decode parses heap rows in the same launch as the page indexes when the heap has at most 1020
    pages (each block computes the heap's item prefix itself); larger heaps take the separate kernels."""
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    rng = np.random.default_rng(17)
    n = 1150
    ind = [rng.standard_normal((1, 40)).astype(np.float32) for _ in range(n)]
    dep = [rng.integers(-300, 300, size=(1, 9000)).astype(np.int16) for _ in range(n)]           # ~20 KB inline: one row per page
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="inline_literal")
    assert seg.heap.size // 32768 > 1020
    want = coracle.decode(seg.heap, seg.toast)
    ds = segment.DeviceSegment(seg.heap.size // 32768, seg.toast.size // 32768)
    out_i = torch.zeros(want["out_ind"].size + 16, dtype=torch.uint8, device="cuda")
    out_d = torch.zeros(want["out_dep"].size + 16, dtype=torch.uint8, device="cuda")
    s = ds.decode_device(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda(), out_i, out_d)
    assert s.status == 0 and s.n_rows == n
    assert out_i[:want["out_ind"].size].cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert out_d[:want["out_dep"].size].cpu().numpy().tobytes() == want["out_dep"].tobytes()
    # and a heap of exactly the limit or just below goes through the fused launch with the same result
    m = 1000
    seg2 = synth.build_segment(ind[:m], dep[:m], ind_storage="ext_raw", dep_storage="inline_literal")
    assert seg2.heap.size // 32768 <= 1020
    want2 = coracle.decode(seg2.heap, seg2.toast)
    s2 = ds.decode_device(torch.from_numpy(seg2.heap).cuda(), torch.from_numpy(seg2.toast).cuda(), out_i, out_d)
    assert s2.status == 0 and s2.n_rows == m
    assert out_d[:want2["out_dep"].size].cpu().numpy().tobytes() == want2["out_dep"].tobytes()
    ds.close()


def test_toast_page_claiming_more_items_than_it_holds_is_refused(torch_cuda, coracle):
    """pd_lower of a TOAST page inflated: the physical walk (ppr:391-414) runs off the end of the page after the real
    tuples and the reference's header unpack fails; found by the seed-shifted corruption soak (a silent pass before)."""
    import da_oracle
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    ind, dep = synth.imagenet_rows(6, seed=3, shape=(36, 36, 3), classes=12, images_per_row=2)
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="inline_pglz")
    for extra in (4, 52 * 256):
        toast = seg.toast.copy()
        page = 32768
        lower = int(toast[page + 12]) | (int(toast[page + 13]) << 8)
        lower += extra                                                      # still a multiple of 4 past the header
        toast[page + 12], toast[page + 13] = lower & 0xFF, lower >> 8
        with pytest.raises(da_oracle.OracleError) as ei:
            coracle.decode(seg.heap, toast)
        st, part = segment.decode_pages(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(toast).cuda())
        assert st == ei.value.code == 4
