"""GPU parity tests: the CUDA path through the C ABI vs. reference outputs (tests/golden) and the
oracle (oracle/) on identical synthetic segments.  Bit-exact: this path moves bytes, it does no
arithmetic on them, so every comparison is on raw bytes."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, SEGMENT_CASES, load_golden_segment

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device (there is no CPU fallback to test)")
    return torch


def gpu_decode(torch, heap, toast, **kw):
    from cerebro_ds_kpgi_b200 import segment
    h = torch.from_numpy(np.ascontiguousarray(heap)).cuda()
    t = torch.from_numpy(np.ascontiguousarray(toast)).cuda()
    status, part = segment.decode_pages(h, t, **kw)
    return status, part


def part_to_dict(part, shapes=None):
    ind_h, dep_h = part.ind.cpu().numpy(), part.dep.cpu().numpy()
    out = {}
    for bid, ind, dep in part.value_views(ind_h, dep_h):
        if shapes is not None:
            ind, dep = ind.reshape(shapes[bid][0]), dep.reshape(shapes[bid][1])
        out[bid] = {"independent_var": ind, "dependent_var": dep}
    return out


@pytest.mark.parametrize("name", SEGMENT_CASES)
def test_golden_segments_bit_exact(name, torch_cuda):
    from conftest import assert_dict_equal
    heap, toast, order, expected = load_golden_segment(name)
    status, part = gpu_decode(torch_cuda, heap, toast)
    assert status == 0
    shapes = {k: (v["independent_var"].shape, v["dependent_var"].shape) for k, v in expected.items()}
    assert_dict_equal(part_to_dict(part, shapes), expected, order)


def test_pglz_kat_on_gpu(torch_cuda):
    from cerebro_ds_kpgi_b200 import segment
    z = np.load(os.path.join(GOLDEN, "pglz_kat.npz"))
    so, ro = z["src_off"], z["raw_off"]
    datums = [z["src"][so[i]:so[i + 1]].tobytes() for i in range(len(so) - 1)]
    raws = [z["raw"][ro[i]:ro[i + 1]].tobytes() for i in range(len(ro) - 1)]
    out, codes = segment.pglz_decompress_batch(datums)
    assert codes == [0] * len(datums)
    for got, want in zip(out, raws):
        assert got == want


def test_pglz_corrupt_streams_flagged(torch_cuda, coracle):
    """Random damage: GPU and oracle must agree on corrupt / not corrupt and on the bytes when intact."""
    from cerebro_ds_kpgi_b200 import segment, synth
    import da_oracle
    rng = np.random.default_rng(5)
    base = [synth.pglz_varlena(rng.integers(0, 6, size=n, dtype=np.uint8), "lz") for n in (50, 400, 3000, 9000)]
    datums = []
    for v in base:
        for _ in range(25):
            b = bytearray(v)
            k = int(rng.integers(8, len(b)))
            b[k] ^= int(rng.integers(1, 256))
            datums.append(bytes(b))
        datums.append(v[:-1][:4] + v[4:-1])                 # truncated (header still claims full length)
        b = bytearray(v); b[3] = (b[3] - 1) & 0xFF           # header one byte short
        datums.append(bytes(b))
    out, codes = segment.pglz_decompress_batch(datums)
    n_bad = 0
    for i, (d, got, code) in enumerate(zip(datums, out, codes)):
        try:
            want = bytes(coracle.pglz_decompress(d))
        except da_oracle.OracleError:
            assert code == da_oracle.ERR_PGLZ_CORRUPT, "datum %d: oracle says corrupt, GPU decoded it" % i
            n_bad += 1
            continue
        assert code == 0 and got == want, "datum %d" % i
    assert n_bad > 10


def test_error_cases_codes(torch_cuda, coracle):
    import da_oracle
    z = np.load(os.path.join(GOLDEN, "error_cases.npz"))
    checked = 0
    for i, name in enumerate(z["names"]):
        name = str(name)
        heap, toast = z["heap_%d" % i], z["toast_%d" % i]
        if name in ("partial_page", "shape_mismatch"):
            continue                                         # file layer / catalog layer (tests/test_gpu_dropin.py)
        with pytest.raises(da_oracle.OracleError) as ei:
            coracle.decode(heap, toast)
        status, _ = gpu_decode(torch_cuda, heap, toast)
        assert status == ei.value.code, name
        checked += 1
    assert checked >= 7


@pytest.mark.parametrize("dep_storage", ["ext_raw", "inline_pglz"])
def test_config2_uncompressed_vs_oracle(dep_storage, torch_cuda, coracle):
    """BASELINE config 2 shape (uncompressed external 112x112x3 f32 rows), small row count."""
    from cerebro_ds_kpgi_b200 import synth
    ind, dep = synth.imagenet_rows(40, seed=2018, kind="normal")
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage=dep_storage)
    want = coracle.decode(seg.heap, seg.toast, ind_sample_bytes=150528, dep_sample_bytes=2000)
    status, part = gpu_decode(torch_cuda, seg.heap, seg.toast, ind_sample_bytes=150528, dep_sample_bytes=2000)
    assert status == 0
    assert np.array_equal(part.layout["order"], want["order"])
    assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()
    assert np.array_equal(part.rows, want["rows"])
    for k in ("ind_off", "dep_off", "ind_len", "dep_len"):
        assert np.array_equal(part.layout[k], want[k]), k


def test_config3_mixed_pglz_vs_oracle(torch_cuda, coracle):
    from cerebro_ds_kpgi_b200 import synth
    a_ind, a_dep = synth.imagenet_rows(12, seed=2018, kind="norm_u8")
    b_ind, b_dep = synth.imagenet_rows(12, seed=2019, kind="normal")
    rng = np.random.default_rng(1)
    ind, dep, st = [], [], []
    for i in range(12):
        if rng.random() < 0.5:
            ind.append(a_ind[i]); dep.append(a_dep[i]); st.append("ext_pglz")
        else:
            ind.append(b_ind[i]); dep.append(b_dep[i]); st.append(["ext_literal", "ext_raw"][i % 2])
    seg = synth.build_segment(ind, dep, ind_storage=st, dep_storage="inline_pglz")
    want = coracle.decode(seg.heap, seg.toast)
    status, part = gpu_decode(torch_cuda, seg.heap, seg.toast)
    assert status == 0
    assert np.array_equal(part.layout["order"], want["order"])
    assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()


def test_config4_criteo_vs_oracle(torch_cuda, coracle):
    from cerebro_ds_kpgi_b200 import synth
    ind, dep = synth.criteo_buffers(6, seed=2018, rows_per_buffer=256)
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="inline_pglz")
    want = coracle.decode(seg.heap, seg.toast, ind_sample_bytes=7306 * 4, dep_sample_bytes=4)
    status, part = gpu_decode(torch_cuda, seg.heap, seg.toast, ind_sample_bytes=7306 * 4, dep_sample_bytes=4)
    assert status == 0
    assert part.ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()
    assert part.ind.numel() == 6 * 256 * 7306 * 4


def test_consumer_casts_vs_oracle(torch_cuda, coracle):
    from cerebro_ds_kpgi_b200 import segment
    torch = torch_cuda
    rng = np.random.default_rng(3)
    oh = np.zeros((777, 1000), dtype=np.int16)
    oh[np.arange(777), rng.integers(0, 1000, 777)] = 1
    oh[5] = 0                                               # all-equal row -> index 0
    oh[6, [10, 900]] = 1                                    # tie -> first
    got = segment.onehot_argmax(torch.from_numpy(oh).cuda()).cpu().numpy()
    assert np.array_equal(got, coracle.argmax_i16(oh))
    x = rng.standard_normal((5, 112, 112, 3)).astype(np.float32)
    got = segment.nhwc_to_nchw(torch.from_numpy(x).cuda()).cpu().numpy()
    assert got.tobytes() == coracle.nhwc_to_nchw(x).tobytes()
    b = segment.to_bf16(torch.from_numpy(x).cuda())
    assert torch.equal(b.cpu(), torch.from_numpy(x).to(torch.bfloat16))   # RNE, the documented non-reference dtype


def test_one_shot_decode_device_and_host(torch_cuda, coracle):
    """cds_decode_device / cds_decode_host with caller-sized output tensors (no intermediate sync)."""
    from cerebro_ds_kpgi_b200 import segment, synth
    torch = torch_cuda
    ind, dep = synth.imagenet_rows(30, seed=7, kind="normal")
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="ext_raw")
    want = coracle.decode(seg.heap, seg.toast)
    ds = segment.DeviceSegment(seg.heap.size // 32768, seg.toast.size // 32768, stage=True)
    out_ind = torch.zeros(30 * 150528, dtype=torch.uint8, device="cuda")
    out_dep = torch.zeros(30 * 2000, dtype=torch.uint8, device="cuda")
    h, t = torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda()
    s = ds.decode_device(h, t, out_ind, out_dep, 150528, 2000)
    assert s.status == 0 and s.n_rows == 30 and s.ind_bytes == 30 * 150528
    assert out_ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    out_ind.zero_(); out_dep.zero_()
    ph, pt = segment.PinnedBuffer(seg.heap.size), segment.PinnedBuffer(seg.toast.size)
    ph.array[:] = seg.heap; pt.array[:] = seg.toast
    s = ds.decode_host(ph.array, pt.array, out_ind, out_dep, 150528, 2000)
    assert s.status == 0
    assert out_ind.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert out_dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()
    # too-small output: refused, nothing written out of bounds
    small = torch.zeros(1000, dtype=torch.uint8, device="cuda")
    from cerebro_ds_kpgi_b200._lib import CdsError
    with pytest.raises(CdsError) as ei:
        ds.decode_device(h, t, small, out_dep, 150528, 2000)
    assert ei.value.code == 102
    assert int(small.sum()) == 0
    ds.close()


def test_fused_bf16_output_matches_rne_of_oracle(torch_cuda, coracle):
    """Optional output dtype (not in the reference): the gather and the pglz decoder emit bfloat16 directly.
    Must equal round-to-nearest-even of the oracle's float32 output, for every storage form and for
    odd result offsets; dependent_var is untouched."""
    from cerebro_ds_kpgi_b200 import synth
    torch = torch_cuda
    rng = np.random.default_rng(8)
    ind, dep = [], []
    for i in range(24):
        b = 1 + i % 3
        ind.append((rng.standard_normal((b, 37, 29, 3)) * 10.0 ** int(rng.integers(-3, 4))).astype(np.float32))
        dep.append(np.eye(7, dtype=np.int16)[rng.integers(0, 7, b)])
    st = [["ext_raw", "ext_pglz", "ext_literal"][i % 3] for i in range(24)]
    seg = synth.build_segment(ind, dep, ind_storage=st, dep_storage=["inline_pglz", "ext_raw"] * 12)
    want = coracle.decode(seg.heap, seg.toast)
    status, part = gpu_decode(torch, seg.heap, seg.toast, ind_bf16=True)
    assert status == 0
    ref = torch.from_numpy(want["out_ind"].view(np.float32).copy()).to(torch.bfloat16)
    got = part.ind.view(torch.bfloat16).cpu()
    assert got.shape == ref.shape and torch.equal(got.view(torch.int16), ref.view(torch.int16))
    assert part.dep.cpu().numpy().tobytes() == want["out_dep"].tobytes()
    # real shape, uncompressed: 112x112x3 rows
    ind, dep = synth.imagenet_rows(20, seed=3, kind="normal")
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="ext_raw")
    want = coracle.decode(seg.heap, seg.toast)
    status, part = gpu_decode(torch, seg.heap, seg.toast, ind_bf16=True, ind_sample_bytes=150528, dep_sample_bytes=2000)
    assert status == 0
    ref = torch.from_numpy(want["out_ind"].view(np.float32).copy()).to(torch.bfloat16)
    assert torch.equal(part.ind.view(torch.bfloat16).cpu().view(torch.int16), ref.view(torch.int16))


def test_fused_uint8_normalise_is_the_reference_formula_bit_for_bit(torch_cuda, coracle):
    """north_star's "fused cast/normalise": uint8 pixel tables leave the decode as float32
    ((x / 255.0) - mean) / std  (preprocessing/imagenet/generate_h5_file.py:77-81; evaluated in float64 and
    stored to float32, as numpy does there) — 0 ULP, for uncompressed, pglz and literal-only values."""
    from cerebro_ds_kpgi_b200 import synth
    torch = torch_cuda
    rng = np.random.default_rng(21)
    mean, std = np.array([0.485, 0.456, 0.406]), np.array([0.229, 0.224, 0.225])
    lut = synth.norm_lut()
    want_lut = (((np.arange(256)[None, :] / 255.0) - mean[:, None]) / std[:, None]).astype(np.float32)
    assert lut.tobytes() == want_lut.tobytes()
    ind, dep = [], []
    for i in range(18):
        b = 1 + i % 2
        px = rng.integers(0, 256, size=(b, 112, 112, 3), dtype=np.uint8)
        if i % 3 == 1:
            px[:, 20:90] = px[:, 20:21]                        # repeated rows: real matches for pglz
        ind.append(px)
        dep.append(np.eye(9, dtype=np.int16)[rng.integers(0, 9, b)])
    st = [["ext_raw", "ext_pglz", "ext_literal"][i % 3] for i in range(18)]
    seg = synth.build_segment(ind, dep, ind_storage=st, dep_storage="inline_pglz")
    ref = coracle.decode(seg.heap, seg.toast)                   # the stored uint8 bytes, in result order
    px_all = ref["out_ind"].reshape(-1, 3)
    img = px_all.astype(np.float64) / 255.0                     # the reference's three statements
    img = img - mean
    img = img / std
    want = img.astype(np.float32).reshape(-1)
    status, part = gpu_decode(torch, seg.heap, seg.toast, u8_normalize=lut, ind_sample_bytes=112 * 112 * 3, dep_sample_bytes=18)
    assert status == 0
    got = part.ind.view(torch.float32).cpu().numpy()
    assert got.shape == want.shape and got.tobytes() == want.tobytes()
    assert part.dep.cpu().numpy().tobytes() == ref["out_dep"].tobytes()


def test_deliberate_deviation_on_offsets_outside_the_output(torch_cuda, coracle):
    """A match with offset 0, or one that reaches before the start of the output, is reported corrupt here, while
    the reference's Python decoder indexes its zero-filled bytearray with dest[dp - off] and, depending on the sizes,
    returns bytes (offset 0; a negative index that wraps inside the buffer) or dies with an IndexError.  The golden
    file records what it did (tests/golden/make_golden_deviation.py); DESIGN.md §5 lists the deviation."""
    import da_oracle
    from cerebro_ds_kpgi_b200 import segment
    z = np.load(os.path.join(GOLDEN, "pglz_deviation.npz"))
    names = [str(n) for n in z["names"]]
    datums = [z["datum_%d" % i].tobytes() for i in range(len(names))]
    ref = {n: (str(z["raised_%d" % i][0]), z["out_%d" % i].tobytes()) for i, n in enumerate(names)}
    assert ref["offset_zero"] == ("", b"A\x00\x00\x00B")                                   # decodes: copies the zero it is about to write
    assert ref["offset_before_start"][0].startswith("IndexError")
    assert ref["offset_wraps_to_unwritten"][0] == "" and ref["offset_wraps_to_unwritten"][1][:5] == b"A\x00\x00\x00\x00"
    out, codes = segment.pglz_decompress_batch(datums)
    assert codes == [da_oracle.ERR_PGLZ_CORRUPT] * 3 and out == [None] * 3
    for d in datums:                                                                           # the oracle agrees with the library
        with pytest.raises(da_oracle.OracleError):
            coracle.pglz_decompress(d)
