"""Boundary tests: the reference's own FFI symbol bound the reference's way, the caller's CUDA
device left as it was, no state carried from one decode to the next on a handle, and literal-only pglz streams
(unpacked by the gather kernel itself) against the oracle."""
import ctypes
import os
import struct
import time

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    torch.cuda.set_device(0)
    return torch


def wrap_function(lib, funcname, restype, argtypes):
    """The reference's binding helper, pg_page_reader.py:51-57."""
    func = lib.__getattr__(funcname)
    func.restype = restype
    func.argtypes = argtypes
    return func


def test_reference_ffi_symbol_bound_the_reference_way(torch_cuda, capfd):
    """pglz_decompress exactly as pg_page_reader.py:47-49 binds it: wrap_function(lib, 'pglz_decompress', None,
    [c_char_p, c_char_p]) with dest = ctypes.create_string_buffer(rawsize) (ppr:247-248), over the 280 vectors that
    both reference decoders produced, then one corrupt stream (the C reference printf()s, pg_lzcompress.c:138-139)."""
    from cerebro_ds_kpgi_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    pglz_decompress_c = wrap_function(lib, 'pglz_decompress', None, [ctypes.c_char_p, ctypes.c_char_p])
    z = np.load(os.path.join(GOLDEN, "pglz_kat.npz"))
    so, ro = z["src_off"], z["raw_off"]
    assert len(so) - 1 >= 280
    for i in range(len(so) - 1):
        source = z["src"][so[i]:so[i + 1]].tobytes()
        rawsize = struct.unpack_from("<I", source, 4)[0]
        dest = ctypes.create_string_buffer(rawsize)
        pglz_decompress_c(source, dest)
        assert dest.raw == z["raw"][ro[i]:ro[i + 1]].tobytes(), i
    out = capfd.readouterr().out
    assert "corrupt" not in out
    # a stream cut one byte short (its header says so too): decoding ends before the output is full (ppr:229-230)
    k = max(range(len(so) - 1), key=lambda i: so[i + 1] - so[i])
    good = z["src"][so[k]:so[k + 1]].tobytes()
    bad = struct.pack(">I", 0x40000000 | (len(good) - 1)) + good[4:-1]
    dest = ctypes.create_string_buffer(struct.unpack_from("<I", bad, 4)[0])
    pglz_decompress_c(bad, dest)
    assert capfd.readouterr().out == "compressed data is corrupt"


def test_entry_points_restore_the_callers_device(torch_cuda, coracle):
    """A library call must not leave another CUDA device current in the calling thread (needs two GPUs: the
    current device is per thread and shared through the driver, so torch sees what the library leaves behind)."""
    from cerebro_ds_kpgi_b200 import _lib, segment, synth
    if torch_cuda.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    L = _lib.require_device()
    ind, dep = synth.imagenet_rows(4, seed=1, shape=(8, 8, 3), classes=10)
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="inline_pglz")
    want = coracle.decode(seg.heap, seg.toast)
    with torch_cuda.cuda.device(1):
        h = torch_cuda.from_numpy(seg.heap).cuda()
        t = torch_cuda.from_numpy(seg.toast).cuda()
        out_ind = torch_cuda.zeros(want["out_ind"].size + 64, dtype=torch_cuda.uint8, device="cuda")
        out_dep = torch_cuda.zeros(want["out_dep"].size + 64, dtype=torch_cuda.uint8, device="cuda")
        stream = torch_cuda.cuda.current_stream()
    torch_cuda.cuda.set_device(0)
    ds = segment.DeviceSegment(h.numel() // 32768, t.numel() // 32768, device=1)
    assert torch_cuda.cuda.current_device() == 0
    ds.decode_device(h, t, out_ind, out_dep, stream=stream, sync=False)
    assert torch_cuda.cuda.current_device() == 0
    s = ds.finish(stream=stream)
    assert torch_cuda.cuda.current_device() == 0
    assert s.status == 0 and out_ind[:want["out_ind"].size].cpu().numpy().tobytes() == want["out_ind"].tobytes()
    ds.close()
    assert torch_cuda.cuda.current_device() == 0


def test_small_values_then_packed_buffers_on_one_handle(torch_cuda, coracle):
    """The pglz stage is decided on the device for each decode: a handle that has just decoded a segment of small
    values decodes a segment of packed buffers (1.5 MB values) through the wide decoder at once — nothing is carried
    over from the previous summary (round 1: seconds, one warp per value)."""
    from cerebro_ds_kpgi_b200 import _lib, segment, synth
    L = _lib.require_device()
    s_ind, s_dep = synth.imagenet_rows(12, seed=3, kind="norm_u8")
    small = synth.build_segment(s_ind, s_dep, ind_storage="ext_pglz", dep_storage="inline_pglz")
    p_ind, p_dep = synth.imagenet_rows(3, seed=4, kind="norm_u8", images_per_row=40)      # 6 MB per value
    packed = synth.build_segment(p_ind, p_dep, ind_storage="ext_pglz", dep_storage="inline_pglz")
    hp = max(small.heap.size, packed.heap.size) // 32768
    tp = max(small.toast.size, packed.toast.size) // 32768
    ds = segment.DeviceSegment(hp, tp, device=0)

    def run(seg):
        want = coracle.decode(seg.heap, seg.toast)
        h, t = torch_cuda.from_numpy(seg.heap).cuda(), torch_cuda.from_numpy(seg.toast).cuda()
        out_ind = torch_cuda.full((want["out_ind"].size + 64,), 0xEE, dtype=torch_cuda.uint8, device="cuda")
        out_dep = torch_cuda.full((want["out_dep"].size + 64,), 0xEE, dtype=torch_cuda.uint8, device="cuda")
        torch_cuda.cuda.synchronize()
        t0 = time.perf_counter()
        ds.decode_device(h, t, out_ind, out_dep, sync=False)
        s = ds.finish()
        dt = time.perf_counter() - t0
        assert s.status == 0
        assert out_ind[:want["out_ind"].size].cpu().numpy().tobytes() == want["out_ind"].tobytes()
        assert out_dep[:want["out_dep"].size].cpu().numpy().tobytes() == want["out_dep"].tobytes()
        return s, dt

    before = int(L.cds_wide_decoded_count())
    s, _ = run(small)
    assert s.n_wide_values == 0 and int(L.cds_wide_decoded_count()) == before
    run(small)
    s, dt = run(packed)
    assert s.n_wide_values == 3 and int(L.cds_wide_decoded_count()) - before == 3
    assert dt < 0.05, "packed buffers after small values took %.1f ms" % (dt * 1000.0)
    s, _ = run(small)
    assert int(L.cds_wide_decoded_count()) - before == 3
    ds.close()


def _lit_varlena(raw, ctrl_of_last=0, damage_at=None):
    """A literal-only pglz varlena built by hand: groups of one control byte + up to 8 literals."""
    raw = bytes(raw)
    body = bytearray()
    n_groups = (len(raw) + 7) // 8
    for g in range(n_groups):
        body.append(ctrl_of_last if g == n_groups - 1 else 0)
        body += raw[8 * g:8 * g + 8]
    if damage_at is not None:
        body[damage_at] = 0x10
    return struct.pack(">I", 0x40000000 | (8 + len(body))) + struct.pack("<I", len(raw)) + bytes(body)


@pytest.mark.parametrize("shape", [(16, 16, 3), (5, 7, 3), (1, 3, 1), (37, 41, 3)])
def test_literal_only_external_values_vs_oracle(shape, torch_cuda, coracle):
    """External literal-only streams of many sizes and output alignments (value sizes that are not multiples of 8
    put the following values at odd offsets), mixed with compressed and uncompressed ones."""
    from cerebro_ds_kpgi_b200 import segment, synth
    n = 23
    ind, dep = synth.imagenet_rows(n, seed=shape[0] * 100 + shape[1], shape=shape, classes=7, kind="normal", images_per_row=3)
    ind_st = ["ext_literal", "ext_pglz", "ext_literal", "ext_raw"] * 6
    dep_st = ["ext_literal", "inline_pglz", "inline_literal", "ext_literal", "ext_raw"] * 5
    seg = synth.build_segment(ind, dep, ind_storage=ind_st[:n], dep_storage=dep_st[:n])
    want = coracle.decode(seg.heap, seg.toast)
    h, t = torch_cuda.from_numpy(seg.heap).cuda(), torch_cuda.from_numpy(seg.toast).cuda()
    status, part = segment.decode_pages(h, t)
    assert status == 0
    assert np.array_equal(part.layout["order"], want["order"])
    assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()


def test_literal_only_control_bytes_are_checked(torch_cuda, coracle):
    """A literal-only stream whose control byte announces a tag is corrupt (ppr:229-230); unused bits of the FINAL
    group's control byte are never looked at (ppr:203) — same verdicts and bytes as the oracle, through the segment
    path (the gather's literal unpacking) and through the batch path (warp decoder)."""
    import da_oracle
    from cerebro_ds_kpgi_b200 import segment, synth
    rng = np.random.default_rng(77)
    raw = rng.integers(0, 256, 20003, dtype=np.uint8).tobytes()          # 3 chunks, final group of 3 literals
    cases = {
        "intact": _lit_varlena(raw),
        "unused_bits_set": _lit_varlena(raw, ctrl_of_last=0xF8),          # bits 3..7 of the last control byte: no item there
        "used_bit_set": _lit_varlena(raw, ctrl_of_last=0x04),             # bit 2 belongs to the third literal
        "tag_in_first_chunk": _lit_varlena(raw, damage_at=9 * 100),
        "tag_in_second_chunk": _lit_varlena(raw, damage_at=9 * 1000),
        "tag_in_last_group_but_one": _lit_varlena(raw, damage_at=9 * ((len(raw) + 7) // 8 - 2)),
    }
    out, codes = segment.pglz_decompress_batch(list(cases.values()))
    for (name, datum), got, code in zip(cases.items(), out, codes):
        try:
            want = bytes(coracle.pglz_decompress(datum))
            assert code == 0 and got == want, name
        except da_oracle.OracleError:
            assert code == da_oracle.ERR_PGLZ_CORRUPT, name
    assert codes[0] == 0 and codes[1] == 0 and codes[2] != 0 and codes[3] != 0

    # the same streams as external values of a segment: build an all-literal segment, then swap payloads in place
    for name, datum in cases.items():
        ind = [np.frombuffer(raw, dtype=np.uint8).copy()]
        dep = [np.zeros((1, 4), dtype=np.int16)]
        seg = synth.build_segment(ind, dep, ind_storage="ext_literal", dep_storage="inline_pglz")
        stream = datum[8:]
        toast = seg.toast.copy()
        good = _lit_varlena(raw)[8:]
        # locate every byte where the damaged stream differs and patch the page image (chunks are in order)
        diff = [i for i in range(len(stream)) if stream[i] != good[i]]
        flat = toast.tobytes()
        for i in diff:
            # stream byte i lives in chunk (i + 4) // 8140 at payload offset (i + 4) % 8140; find it by its neighbourhood
            lo = max(0, i - 6)
            hi = min(len(good), i + 7)
            # keep the window inside one chunk
            c = (i + 4) // 8140
            lo = max(lo, c * 8140 - 4)
            hi = min(hi, (c + 1) * 8140 - 4)
            at = flat.find(good[lo:hi])
            assert at >= 0 and flat.find(good[lo:hi], at + 1) < 0, "window not unique"
            toast[at + (i - lo)] = stream[i]
        try:
            want = coracle.decode(seg.heap, toast)
            want_code = 0
        except da_oracle.OracleError as e:
            want, want_code = None, e.code
        h, t = torch_cuda.from_numpy(seg.heap).cuda(), torch_cuda.from_numpy(toast).cuda()
        status, part = segment.decode_pages(h, t)
        assert status == want_code, name
        if want is not None:
            assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes(), name


def test_helper_work_memory_is_kept_and_can_be_trimmed(torch_cuda):
    """cds_pglz_decompress_batch keeps its work memory between calls (no cudaMalloc per call); cds_trim_work_memory
    frees it and the next call simply allocates again."""
    from cerebro_ds_kpgi_b200 import _lib, segment, synth
    L = _lib.require_device()
    rng = np.random.default_rng(2)
    datums = [synth.pglz_varlena(rng.integers(0, 4, size=5000, dtype=np.uint8), "lz") for _ in range(20)]
    want, codes = segment.pglz_decompress_batch(datums)
    assert codes == [0] * 20
    free0 = torch_cuda.cuda.mem_get_info()[0]
    for _ in range(5):
        got, codes = segment.pglz_decompress_batch(datums)
        assert got == want
    assert torch_cuda.cuda.mem_get_info()[0] == free0          # nothing was allocated or freed on the way
    assert L.cds_trim_work_memory(0) == 0
    got, codes = segment.pglz_decompress_batch(datums)
    assert got == want and codes == [0] * 20
