"""Host-side logic: the synthetic segment writer and the pglz encoder (inputs of every parity test)."""
import struct

import numpy as np

import da_oracle


def test_encoder_round_trips(built, coracle):
    from cerebro_ds_kpgi_b200 import synth
    rng = np.random.default_rng(0)
    for n in [0, 1, 2, 3, 7, 8, 9, 17, 18, 19, 272, 273, 274, 5000, 70000]:
        for raw in (rng.integers(0, 256, n, dtype=np.uint8), np.zeros(n, np.uint8), rng.integers(0, 3, n, dtype=np.uint8)):
            for mode in ("lz", "literal"):
                v = synth.pglz_varlena(raw, mode)
                assert struct.unpack(">I", v[:4])[0] == (0x40000000 | len(v))
                assert bytes(coracle.pglz_decompress(v)) == raw.tobytes()
    v = synth.pglz_varlena(np.zeros(100000, np.uint8))
    assert len(v) < 2000                      # long overlapping matches (len 273, offset 1)


def test_page_layout_invariants(built):
    from cerebro_ds_kpgi_b200 import synth
    ind, dep = synth.imagenet_rows(9, seed=1, shape=(50, 50, 3), classes=40, kind="normal")
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage=["ext_raw", "inline_pglz", "ext_pglz"] * 3)
    assert seg.heap.size % 32768 == 0 and seg.toast.size % 32768 == 0
    for rel in (seg.heap, seg.toast):
        for p in range(rel.size // 32768):
            page = rel[p * 32768:(p + 1) * 32768]
            lower, upper, special = struct.unpack_from("<HHH", page, 12)
            assert special == 32768 and 24 <= lower <= upper <= 32768 and (lower - 24) % 4 == 0
            prev = 32768
            for i in range((lower - 24) // 4):
                w = struct.unpack_from("<I", page, 24 + 4 * i)[0]
                off, flags, ln = w & 0x7FFF, (w >> 15) & 3, w >> 17
                assert flags == 1 and off % 8 == 0 and off + ln <= prev and page[off + 22] == 24
                prev = off
            assert prev == upper or lower == 24
    # 4 full chunks per TOAST page: 8140 + 4 + 24 + 8 = 8176  (pg_page_reader.py:627-628)
    w = struct.unpack_from("<I", seg.toast, 24)[0]
    assert w >> 17 == 8176


def test_python_and_c_oracle_agree_on_generated(built, coracle):
    from cerebro_ds_kpgi_b200 import synth
    ind, dep = synth.criteo_buffers(3, seed=9, rows_per_buffer=16)
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="inline_pglz", buffer_ids=[5, 1, 3])
    a = da_oracle.input_fn_bytes(seg.heap, seg.toast, None, pglz=coracle.pglz_decompress)
    b = coracle.decode_dict(seg.heap, seg.toast)
    assert list(a.keys()) == list(b.keys()) == [5, 1, 3]
    for k in a:
        assert a[k]["independent_var"].tobytes() == b[k]["independent_var"].tobytes() == seg.expected[k]["independent_var"].tobytes()
        assert a[k]["dependent_var"].tobytes() == b[k]["dependent_var"].tobytes() == seg.expected[k]["dependent_var"].tobytes()


def test_rows_for_toast_bytes(built):
    from cerebro_ds_kpgi_b200 import synth
    n = synth.rows_for_toast_bytes(64 << 20)
    ind, dep = synth.imagenet_rows(8, seed=1, kind="normal")
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="ext_raw", keep_expected=False)
    per_row = seg.toast.size / 8
    assert abs(n * per_row - (64 << 20)) / (64 << 20) < 0.08


def test_catalog_round_trip(built, tmp_path):
    import dill
    from cerebro_ds_kpgi_b200 import synth
    ind, dep = synth.imagenet_rows(2, seed=1, shape=(4, 4, 3), classes=3)
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", table_name="t")
    p = tmp_path / "sys_cat.dill"
    synth.write_catalog(str(p), {"train": [dict(gp_segment_id=3, relname="t", relfilenode=7, toast_relfilenode=8)]}, seg.shape_rows)
    cats = dill.load(open(p, "rb"))
    assert list(cats) == ["train", "valid", "shape"]
    assert int(cats["train"].iloc[0]["toast_relfilenode"]) == 8
    assert cats["shape"].iloc[1]["independent_var_shape"] == [1, 4, 4, 3]
