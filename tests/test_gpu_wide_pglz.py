"""Intra-value parallel pglz (csrc/pglz_wide.cuh) against the oracle: the same streams the
warp decoder is tested with, pushed through the wide path by lowering its threshold, plus values of many
output blocks, corrupt streams, and a segment whose values are packed buffers of many images."""
import struct

import numpy as np
import pytest

from test_gpu_fuzz import structured_bytes

pytestmark = pytest.mark.gpu

# CDS_FUZZ_SEED=k shifts every generator seed: soak runs over other random cases (0: the committed, reproducible ones)
SEED_SHIFT = int(__import__("os").environ.get("CDS_FUZZ_SEED", "0"))
BLOCK = 12 * 4096          # kWideBlock of csrc/pglz_wide.cuh: output bytes per block of the wide decoder


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    return torch


NO_LIMIT = 2 ** 64 - 1
CURRENT_SHAPE = ["stored_blocks"]


@pytest.fixture(params=["stored_blocks", "two_parses"])
def wide_threshold(request):
    """Lower the wide-decode threshold for one test, restore the default afterwards.  Every test runs twice: with the store
    of resolved blocks (each block parsed once) and without it (cds_set_wide_store_limit(0): each block parsed twice)."""
    from cerebro_ds_kpgi_b200 import _lib
    L = _lib.require_device()
    CURRENT_SHAPE[0] = request.param
    _lib.check(L.cds_set_wide_store_limit(NO_LIMIT if request.param == "stored_blocks" else 0))

    def set_to(n, force=1):
        _lib.check(L.cds_set_wide_pglz_threshold(int(n), int(force)))
    yield set_to
    _lib.check(L.cds_set_wide_pglz_threshold(1 << 20, 0))
    _lib.check(L.cds_set_wide_store_limit(NO_LIMIT))


def wide_count():
    from cerebro_ds_kpgi_b200 import _lib
    return int(_lib.load().cds_wide_decoded_count())


def image_like(rng, n):
    """float32 bytes with 256 distinct patterns per channel (normalised uint8 pixels): short matches everywhere."""
    lut = ((np.arange(256)[None, :] / 255.0 - np.array([0.485, 0.456, 0.406])[:, None]) /
           np.array([0.229, 0.224, 0.225])[:, None]).astype(np.float32)
    k = n // 4 + 1
    px = rng.integers(0, 256, k)
    vals = lut[np.arange(k) % 3, px]
    return vals.view(np.uint8)[:n].copy()


def sparse_like(rng, n):
    """Criteo-like: float32 zeros with a few ones (long offset-1 runs chained through earlier runs)."""
    x = np.zeros(n // 4 + 1, dtype=np.float32)
    x[rng.integers(0, x.size, max(1, x.size // 180))] = 1.0
    return x.view(np.uint8)[:n].copy()


def test_wide_matches_oracle_on_structured_streams(torch_cuda, coracle, wide_threshold):
    from cerebro_ds_kpgi_b200 import segment, synth
    wide_threshold(1024)
    rng = np.random.default_rng(4242 + SEED_SHIFT)
    datums, raws = [], []
    sizes = [1024, 1025, 4095, 4096, 4097, BLOCK - 1, BLOCK, BLOCK + 1, BLOCK + 4095, 2 * BLOCK, 2 * BLOCK + 1, 3 * BLOCK - 1, 57344, 300001]
    for t, n in enumerate(sizes + [int(rng.integers(1024, 400000)) for _ in range(24)]):
        kind = t % 4
        raw = (structured_bytes(rng, n) if kind == 0 else image_like(rng, n) if kind == 1 else
               sparse_like(rng, n) if kind == 2 else rng.integers(0, 256, n, dtype=np.uint8))
        mode = "literal" if t % 7 == 3 else "lz"
        datums.append(synth.pglz_varlena(raw, mode, effort=int(rng.integers(1, 40))))
        raws.append(raw.tobytes())
    # small values in the same batch keep going through the warp decoder
    for n in (0, 1, 9, 500, 1023):
        raw = structured_bytes(rng, n)
        datums.append(synth.pglz_varlena(raw, "lz"))
        raws.append(raw.tobytes())
    before = wide_count()
    out, codes = segment.pglz_decompress_batch(datums)
    assert wide_count() - before == len(datums) - 5              # every value of >= 1024 bytes took the wide path
    assert codes == [0] * len(datums)
    for i, (got, want) in enumerate(zip(out, raws)):
        assert got == want, "datum %d (%d bytes)" % (i, len(want))
    for d, want in zip(datums[:6], raws[:6]):
        assert bytes(coracle.pglz_decompress(d)) == want


def test_wide_large_values(torch_cuda, coracle, wide_threshold):
    """Values of hundreds of output blocks with the default threshold (1 MiB)."""
    from cerebro_ds_kpgi_b200 import segment, synth
    rng = np.random.default_rng(99 + SEED_SHIFT)
    raws = [image_like(rng, 9 * 1024 * 1024 + 13), sparse_like(rng, 7306 * 4 * 300), structured_bytes(rng, 3_000_000),
            rng.integers(0, 256, 2_500_000, dtype=np.uint8)]
    modes = ["lz", "lz", "lz", "literal"]
    datums = [synth.pglz_varlena(r, m) for r, m in zip(raws, modes)]
    before = wide_count()
    out, codes = segment.pglz_decompress_batch(datums)
    assert wide_count() - before == 4
    assert codes == [0, 0, 0, 0]
    for got, want, d in zip(out, raws, datums):
        assert got == want.tobytes()
        assert bytes(coracle.pglz_decompress(d)) == want.tobytes()


def test_cost_model_keeps_many_medium_values_on_their_warps(torch_cuda, wide_threshold):
    """A thousand 7 MB Criteo buffers decode faster with one warp each; only few, huge values go wide."""
    from cerebro_ds_kpgi_b200 import segment, synth
    rng = np.random.default_rng(3 + SEED_SHIFT)
    raw = sparse_like(rng, 1_100_000)
    datum = synth.pglz_varlena(raw, "lz")
    before = wide_count()
    out, codes = segment.pglz_decompress_batch([datum] * 120)
    assert wide_count() == before                               # not taken: 120 run-dominated values keep 120 warps busy
    assert codes == [0] * 120 and all(o == raw.tobytes() for o in out)
    out, codes = segment.pglz_decompress_batch([datum] * 3)
    assert wide_count() - before == 3                           # three of them alone would leave the GPU idle
    assert codes == [0] * 3 and all(o == raw.tobytes() for o in out)


def test_wide_corrupt_streams_agree_with_oracle(torch_cuda, coracle, wide_threshold):
    """Damage streams that take the wide path: wherever the oracle refuses, the wide decoder refuses; wherever it
    decodes, the bytes are identical."""
    from cerebro_ds_kpgi_b200 import segment, synth
    wide_threshold(1024)
    rng = np.random.default_rng(5150 + SEED_SHIFT)
    base = [synth.pglz_varlena(structured_bytes(rng, 150000), "lz"), synth.pglz_varlena(image_like(rng, 200000), "lz"),
            synth.pglz_varlena(sparse_like(rng, 180000), "lz")]
    datums = []
    for t in range(90):
        d = bytearray(base[t % 3])
        kind = t % 6
        if kind == 0:                                   # flip one stream byte
            d[int(rng.integers(8, len(d)))] ^= int(rng.integers(1, 256))
        elif kind == 1:                                 # early flip: offsets before the output start are likely
            d[int(rng.integers(8, 40))] ^= int(rng.integers(1, 256))
        elif kind == 2:                                 # truncate the stream (header adjusted)
            cut = int(rng.integers(9, len(d)))
            d = d[:cut]
            d[0:4] = struct.pack(">I", 0x40000000 | cut)
        elif kind == 3:                                 # rawsize off by a little
            raw = struct.unpack_from("<I", d, 4)[0]
            d[4:8] = struct.pack("<I", max(1024, raw + int(rng.integers(-3, 4))))
        elif kind == 4:                                 # append junk (header adjusted)
            d += bytes(rng.integers(0, 256, int(rng.integers(1, 30)), dtype=np.uint8))
            d[0:4] = struct.pack(">I", 0x40000000 | len(d))
        else:                                           # zero an offset byte somewhere
            d[int(rng.integers(8, len(d)))] = 0
        datums.append(bytes(d))
    before = wide_count()
    out, codes = segment.pglz_decompress_batch(datums)
    assert wide_count() - before >= 60            # (a damaged header can push a stream back to the warp decoder)
    n_bad = 0
    for i, d in enumerate(datums):
        try:
            want = bytes(coracle.pglz_decompress(d))
        except Exception:
            want = None
        if want is None:
            n_bad += 1
            assert codes[i] == 11, i
        else:
            assert codes[i] == 0 and out[i] == want, i
    assert 10 < n_bad < 90


def test_packed_buffer_segment_takes_the_wide_path(torch_cuda, coracle, wide_threshold):
    """B images per buffer (the paper's layout, B = 3210 there): few huge external values per segment."""
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    ind, dep = synth.imagenet_rows(3, seed=11, shape=(112, 112, 3), classes=1000, images_per_row=24)     # 3.6 MB per value
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="ext_pglz")
    want = coracle.decode(seg.heap, seg.toast)
    before = wide_count()
    st, part = segment.decode_pages(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda())
    assert st == 0
    assert part.summary.n_wide_values == 3 and wide_count() - before == 3
    assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()
    # the same segment with the wide decoder switched off gives the same bytes
    wide_threshold(0)
    st2, part2 = segment.decode_pages(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda())
    assert st2 == 0 and part2.summary.n_wide_values == 0
    assert torch.equal(part2.ind, part.ind) and torch.equal(part2.dep, part.dep)


def test_packed_buffers_bf16_and_uint8_outputs_take_the_wide_path(torch_cuda, coracle, wide_threshold):
    """The output conversions of the warp decoder (float32 -> bfloat16, uint8 -> float32 through the table) on the
    wide path: same results as converting the oracle's bytes."""
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    ind, dep = synth.imagenet_rows(3, seed=12, shape=(112, 112, 3), classes=1000, images_per_row=20)     # 3 MB per value
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="ext_pglz")
    want = coracle.decode(seg.heap, seg.toast)
    before = wide_count()
    st, part = segment.decode_pages(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda(), ind_bf16=True)
    assert st == 0 and wide_count() - before == 3
    ref = torch.from_numpy(want["out_ind"].view(np.float32).copy()).to(torch.bfloat16)
    got = part.ind.view(torch.bfloat16).cpu()
    assert got.shape == ref.shape and torch.equal(got.view(torch.int16), ref.view(torch.int16))
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()

    rng = np.random.default_rng(77 + SEED_SHIFT)
    px = []
    for i in range(3):
        x = rng.integers(0, 256, size=(30, 112, 112, 3), dtype=np.uint8)        # 1.1 MB per value
        x[:, 30:80] = x[:, 30:31]                                                # repeated rows: real matches
        px.append(x)
    seg = synth.build_segment(px, dep, ind_storage="ext_pglz", dep_storage="ext_pglz")
    lut = synth.norm_lut()
    flat = np.concatenate([p.reshape(-1, 3) for p in px])
    want_f = lut[np.arange(3)[None, :], flat].astype(np.float32).reshape(-1)
    before = wide_count()
    st, part = segment.decode_pages(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda(), u8_normalize=lut)
    assert st == 0 and wide_count() - before == 3
    got = part.ind.view(torch.float32).cpu().numpy()
    assert got.shape == want_f.shape and got.tobytes() == want_f.tobytes()


def test_wide_periodic_streams(torch_cuda, coracle, wide_threshold):
    """Streams whose bytes repeat with the period of a control group: a parse that starts out of phase stays out
    of phase for ever, so guessing group boundaries would fail; the entry maps must still find them."""
    from cerebro_ds_kpgi_b200 import segment, synth
    wide_threshold(1024)
    raws = [np.zeros(3_000_000, dtype=np.uint8),                                    # ctrl FF + 8 x (0F 01 FF): period 25
            np.full(700_001, 0xFF, dtype=np.uint8),
            np.resize(np.array([1, 2, 3], dtype=np.uint8), 1_000_003),              # offset-3 matches of maximum length
            np.resize(np.arange(256, dtype=np.uint8), 900_000),
            np.resize(np.frombuffer(b"\x00\x00\x80\x3f" + b"\x00" * 60, dtype=np.uint8), 1_500_000)]
    datums = [synth.pglz_varlena(r, "lz") for r in raws] + [synth.pglz_varlena(raws[0][:200_000], "literal")]
    raws.append(raws[0][:200_000])
    before = wide_count()
    out, codes = segment.pglz_decompress_batch(datums)
    assert wide_count() - before == len(datums) and codes == [0] * len(datums)
    for got, want, d in zip(out, raws, datums):
        assert got == want.tobytes()
        assert bytes(coracle.pglz_decompress(d)) == want.tobytes()


def test_corrupt_packed_buffer_in_a_segment_is_refused_like_the_oracle(torch_cuda, coracle, wide_threshold):
    """Damage the stream of one large external value inside the TOAST file: same status and same first-error position
    as with the wide decoder switched off, and the oracle refuses too."""
    import da_oracle
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    ind, dep = synth.imagenet_rows(3, seed=21, shape=(112, 112, 3), classes=1000, images_per_row=12)     # 1.8 MB per value
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="inline_pglz")
    rng = np.random.default_rng(1 + SEED_SHIFT)
    for trial in range(6):
        toast = seg.toast.copy()
        # a payload byte somewhere in the middle pages (page header 24 B + line pointers, tuples at the end of the page)
        page = int(rng.integers(1, toast.size // 32768 - 1))
        pos = page * 32768 + int(rng.integers(2000, 30000))
        toast[pos] ^= int(rng.integers(1, 256))
        try:
            want = coracle.decode(seg.heap, toast)
        except da_oracle.OracleError as e:
            want = e.code
        results = []
        for thr in (1 << 20, 0):
            wide_threshold(thr, force=0)
            before = wide_count()
            st, part = segment.decode_pages(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(toast).cuda())
            results.append((st, int(part.summary.err_where), int(part.summary.err_stage)))
            if thr:
                assert wide_count() - before == 3
            if st == 0:
                assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
        assert results[0] == results[1], (trial, results)
        assert (results[0][0] == 0) == isinstance(want, dict), (trial, results, want if not isinstance(want, dict) else 0)
        if not isinstance(want, dict):
            assert results[0][0] == want


def test_more_wide_candidates_than_the_work_area_holds(torch_cuda, coracle, wide_threshold):
    """A handle's work area lists at most 1024 values; the rest stay with the warp decoder and the result is the same."""
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    wide_threshold(1024, force=1)
    rng = np.random.default_rng(12 + SEED_SHIFT)
    n = 1300
    ind = [np.resize(rng.standard_normal(int(rng.integers(20, 60))).astype(np.float32), (1, int(rng.integers(300, 900)))) for _ in range(n)]
    dep = [rng.integers(0, 2, size=(1, 8)).astype(np.int16) for _ in range(n)]
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="inline_pglz")
    want = coracle.decode(seg.heap, seg.toast)
    before = wide_count()
    st, part = segment.decode_pages(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda())
    assert st == 0
    assert part.summary.n_wide_values == n and wide_count() - before == 1024
    assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()


def test_wide_streams_ending_on_segment_and_block_boundaries(torch_cuda, coracle, wide_threshold):
    """Literal-only streams of 4096 m bytes are 4608 m = 9 * 512 m bytes long: the stream ends exactly on a source
    segment boundary; multiples of the output block (49152 bytes) end exactly on a block boundary."""
    from cerebro_ds_kpgi_b200 import segment, synth
    wide_threshold(1024)
    rng = np.random.default_rng(8 + SEED_SHIFT)
    sizes = [4096, 4096 * 2, 4096 * 11, 4096 * 12, 4096 * 13, 4096 * 14, 4096 * 24, 4096 * 25, BLOCK * 3]
    raws = [rng.integers(0, 256, n, dtype=np.uint8) for n in sizes]
    datums = [synth.pglz_varlena(r, "literal") for r in raws]
    for d, n in zip(datums, sizes):
        assert (len(d) - 8) % 512 == 0, n
    raws += [structured_bytes(rng, n) for n in (BLOCK, BLOCK * 2, BLOCK * 4, 57344)]
    datums += [synth.pglz_varlena(r, "lz") for r in raws[len(sizes):]]
    out, codes = segment.pglz_decompress_batch(datums)
    assert codes == [0] * len(datums)
    for got, want in zip(out, raws):
        assert got == want.tobytes()
    # one byte short / one byte long at those boundaries is refused
    bad = []
    for d in datums[:4]:
        t = bytearray(d[:-1]); t[0:4] = struct.pack(">I", 0x40000000 | len(t)); bad.append(bytes(t))
        t = bytearray(d + b"\x00"); t[0:4] = struct.pack(">I", 0x40000000 | len(t)); bad.append(bytes(t))
    out, codes = segment.pglz_decompress_batch(bad)
    for d, c in zip(bad, codes):
        with pytest.raises(Exception):
            coracle.pglz_decompress(d)
        assert c == 11


@pytest.mark.timeout(120)
def test_wide_refuses_offset_zero_and_offsets_before_the_start_without_spinning(torch_cuda, coracle, wide_threshold):
    """Streams whose chain of control groups is well-formed (so they reach the block kernels) but that hold a match with
    offset 0, or one that reaches before the start of the output.  (An item that is not entered into the block's
    forest once left stale pointers behind, and the pointer-jumping loop never ended: found by tools/soak_fuzz.py.)"""
    from cerebro_ds_kpgi_b200 import segment
    wide_threshold(1024)
    rng = np.random.default_rng(1 + SEED_SHIFT)

    def literal_groups(n_groups):
        return b"".join(b"\x00" + bytes(rng.integers(0, 256, 8, dtype=np.uint8)) for _ in range(n_groups))

    def varlena(stream, rawsize):
        return struct.pack(">I", 0x40000000 | (8 + len(stream))) + struct.pack("<I", rawsize) + stream

    datums = []
    for n_before, n_after in ((137, 0), (137, 9000), (8000, 300), (0, 200)):
        for off, length in ((0, 8), (0, 3), (0, 200)):
            if length <= 17:
                tag = bytes([(length - 3) | ((off >> 8) << 4), off & 0xFF])
            else:
                tag = bytes([0x0F | ((off >> 8) << 4), off & 0xFF, length - 18])
            stream = literal_groups(n_before) + b"\x01" + tag + bytes(rng.integers(0, 256, 7, dtype=np.uint8)) + literal_groups(n_after)
            datums.append(varlena(stream, 8 * n_before + length + 7 + 8 * n_after))
    for off in (1, 100, 4095):                                              # first item of the value: nothing to copy from yet
        tag = bytes([0x0F | ((off >> 8) << 4), off & 0xFF, 255])
        stream = b"\x01" + tag + bytes(rng.integers(0, 256, 7, dtype=np.uint8)) + literal_groups(400)
        datums.append(varlena(stream, 273 + 7 + 3200))
    good = varlena(literal_groups(500), 4000)
    before = wide_count()
    out, codes = segment.pglz_decompress_batch(datums + [good])
    assert wide_count() - before == len(datums) + 1
    assert codes == [11] * len(datums) + [0]
    for d in datums:
        with pytest.raises(Exception):
            coracle.pglz_decompress(d)
    assert out[-1] == bytes(coracle.pglz_decompress(good))


def test_a_handle_gets_its_store_after_the_first_wide_decode(torch_cuda, coracle, wide_threshold):
    """One handle, three segments of packed buffers in a row: the first decode parses every block twice (the handle has no
    store of resolved blocks yet), the later ones once -- also a larger segment, which the store does not hold yet, and a
    smaller one after it.  Same bytes as the oracle every time, in all three output formats."""
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    wide_threshold(1 << 16)
    segs = []
    for seed, per_row in ((21, 12), (22, 12), (23, 30), (24, 5)):
        ind, dep = synth.imagenet_rows(3, seed=seed + SEED_SHIFT, shape=(112, 112, 3), classes=1000, images_per_row=per_row)
        sg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="ext_pglz")
        segs.append((sg, coracle.decode(sg.heap, sg.toast)))
    max_heap = max(sg.heap.size for sg, _ in segs) // 32768
    max_toast = max(sg.toast.size for sg, _ in segs) // 32768
    for bf16 in (False, True):
        ds = segment.DeviceSegment(max_heap, max_toast)
        ds.set_output_format(ind_bf16=bf16)
        try:
            for sg, want in segs:
                ind_bytes = want["out_ind"].size // (2 if bf16 else 1)
                out_ind = torch.full((ind_bytes + 64,), 0xA5, dtype=torch.uint8, device="cuda")
                out_dep = torch.full((want["out_dep"].size + 64,), 0xA5, dtype=torch.uint8, device="cuda")
                before = wide_count()
                s = ds.decode_device(torch.from_numpy(sg.heap).cuda(), torch.from_numpy(sg.toast).cuda(), out_ind, out_dep)
                assert s.status == 0 and wide_count() - before == 3
                if bf16:
                    ref = torch.from_numpy(want["out_ind"].view(np.float32).copy()).to(torch.bfloat16).view(torch.int16)
                    assert torch.equal(out_ind[:ind_bytes].view(torch.int16).cpu(), ref)
                else:
                    assert out_ind[:ind_bytes].cpu().numpy().tobytes() == want["out_ind"].tobytes()
                assert out_dep[:want["out_dep"].size].cpu().numpy().tobytes() == want["out_dep"].tobytes()
                assert bool((out_ind[ind_bytes:] == 0xA5).all()) and bool((out_dep[want["out_dep"].size:] == 0xA5).all())
                if sg is segs[-1][0]:
                    # a limit lowered afterwards also keeps a handle from using the store it holds: same bytes again
                    from cerebro_ds_kpgi_b200 import _lib
                    _lib.check(_lib.load().cds_set_wide_store_limit(0))
                    again = torch.full_like(out_ind, 0xA5)
                    s = ds.decode_device(torch.from_numpy(sg.heap).cuda(), torch.from_numpy(sg.toast).cuda(), again, out_dep)
                    assert s.status == 0 and torch.equal(again, out_ind)
                    _lib.check(_lib.load().cds_set_wide_store_limit(NO_LIMIT if "stored" in CURRENT_SHAPE[0] else 0))
        finally:
            ds.close()
