"""Pin the oracle (oracle/da_oracle.py + oracle/da_oracle.c) against outputs of the UNMODIFIED
reference (tests/golden/, see tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, SEGMENT_CASES, assert_dict_equal, load_golden_segment

import da_oracle


def shapes_of(expected):
    return {k: (list(v["independent_var"].shape), list(v["dependent_var"].shape)) for k, v in expected.items()}


@pytest.mark.parametrize("name", SEGMENT_CASES)
def test_python_oracle_matches_reference(name, coracle):
    heap, toast, order, expected = load_golden_segment(name)
    # big cases: route pglz through the C restatement to keep the CPU suite short
    pglz = da_oracle.pglz_decompress if heap.size + toast.size < 200_000 else coracle.pglz_decompress
    got = da_oracle.input_fn_bytes(heap, toast, shapes_of(expected), pglz=pglz)
    assert_dict_equal(got, expected, order)


@pytest.mark.parametrize("name", SEGMENT_CASES)
def test_c_oracle_matches_reference(name, coracle):
    heap, toast, order, expected = load_golden_segment(name)
    got = coracle.decode_dict(heap, toast, shapes_of(expected), nthreads=2)
    assert_dict_equal(got, expected, order)


def kat_vectors():
    z = np.load(os.path.join(GOLDEN, "pglz_kat.npz"))
    so, ro = z["src_off"], z["raw_off"]
    for i in range(len(so) - 1):
        yield z["src"][so[i]:so[i + 1]].tobytes(), z["raw"][ro[i]:ro[i + 1]].tobytes()


def test_pglz_kat_python_and_c(coracle):
    n = 0
    for src, raw in kat_vectors():
        assert bytes(coracle.pglz_decompress(src)) == raw
        if len(raw) <= 1200:
            assert bytes(da_oracle.pglz_decompress(src)) == raw
        n += 1
    assert n >= 200


def test_pglz_kat_reference_c_if_built():
    """When oracle/_ref/pg_lzcompress.so (the reference's own C file, compiled where it lies) is present,
    it must agree too — it is what `cpu_baseline.kind == "reference"` would time."""
    try:
        ref = da_oracle.RefPglz()
    except RuntimeError:
        pytest.skip("oracle/_ref not built on this machine")
    for src, raw in kat_vectors():
        assert bytes(ref.pglz_decompress(src)) == raw


ERR_EXPECT = {
    "pglz_corrupt_external": da_oracle.ERR_PGLZ_CORRUPT,
    "pglz_corrupt_inline": da_oracle.ERR_PGLZ_CORRUPT,
    "partial_page": da_oracle.ERR_PARTIAL_PAGE,
    "toast_special": da_oracle.ERR_TOAST_SPECIAL,
    "itemid_misaligned": da_oracle.ERR_ITEMID_ALIGN,
    "chunk_count": da_oracle.ERR_CHUNK_COUNT,
    "chunk_size_last": da_oracle.ERR_CHUNK_SIZE,
    "chunk_size_middle": da_oracle.ERR_CHUNK_SIZE,
}
# what the reference raised for each code (type, message prefix) — from tests/golden/error_cases.npz
REF_EXC = {
    da_oracle.ERR_PGLZ_CORRUPT: ("Exception", "compressed data is corrupt"),
    da_oracle.ERR_PARTIAL_PAGE: ("Exception", "FATAL"),
    da_oracle.ERR_TOAST_SPECIAL: ("AssertionError", "THERE SHALL NOT BE INDICES"),
    da_oracle.ERR_ITEMID_ALIGN: ("Exception", "Item identifier length do not match."),
    da_oracle.ERR_CHUNK_COUNT: ("AssertionError", ""),
    da_oracle.ERR_CHUNK_SIZE: ("Exception", "unexpected chunk size"),
    da_oracle.ERR_SHAPE: ("ValueError", "cannot reshape array"),
}


def error_cases():
    z = np.load(os.path.join(GOLDEN, "error_cases.npz"))
    for i, name in enumerate(z["names"]):
        yield str(name), z["heap_%d" % i], z["toast_%d" % i], str(z["types"][i]), str(z["msgs"][i])


def test_error_cases_match_reference_exceptions(coracle):
    seen = 0
    for name, heap, toast, etype, emsg in error_cases():
        if name == "shape_mismatch":
            assert etype == "ValueError"
            continue
        code = ERR_EXPECT[name]
        assert (etype, emsg[:len(REF_EXC[code][1])]) == REF_EXC[code], name
        if name == "partial_page":
            assert toast.size % 32768 != 0        # detected by the file layer (read_relation)
            continue
        with pytest.raises(da_oracle.OracleError) as ei:
            da_oracle.input_fn_bytes(heap, toast, None, pglz=coracle.pglz_decompress)
        assert ei.value.code == code, name
        with pytest.raises(da_oracle.OracleError) as ei:
            coracle.decode(heap, toast)
        assert ei.value.code == code, name
        seen += 1
    assert seen >= 7


def test_uncompressed_external_extension(coracle, built):
    """external + uncompressed is outside the reference; both restatements must agree
    with the generator's source arrays."""
    from cerebro_ds_kpgi_b200 import synth
    ind, dep = synth.imagenet_rows(4, seed=5, shape=(30, 30, 3), classes=12, images_per_row=2, kind="normal")
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage=["ext_raw", "inline_pglz"] * 2)
    shapes = {k: (list(v["independent_var"].shape), list(v["dependent_var"].shape)) for k, v in seg.expected.items()}
    got_c = coracle.decode_dict(seg.heap, seg.toast, shapes)
    got_py = da_oracle.input_fn_bytes(seg.heap, seg.toast, shapes)
    assert_dict_equal(got_c, seg.expected, [1, 3, 0, 2])
    assert_dict_equal(got_py, seg.expected, [1, 3, 0, 2])


def test_consumer_casts(coracle):
    rng = np.random.default_rng(0)
    oh = np.zeros((50, 1000), dtype=np.int16)
    lab = rng.integers(0, 1000, 50)
    oh[np.arange(50), lab] = 1
    assert np.array_equal(coracle.argmax_i16(oh), np.argmax(oh, axis=1))
    x = rng.standard_normal((3, 5, 7, 3)).astype(np.float32)
    assert np.array_equal(coracle.nhwc_to_nchw(x), np.stack([np.rollaxis(i, 2, 0) for i in x]))
