"""GPU tests of the drop-in reader interface (cerebro_ds_kpgi_b200.pg_page_reader / .da) against what the
UNMODIFIED reference returned for the same files (tests/golden/)."""
import os
import struct

import numpy as np
import pytest

from conftest import GOLDEN, SEGMENT_CASES, assert_dict_equal, load_golden_segment

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    return torch


def write_case(tmp_path, heap, toast, expected, table="t_train", seg_id=0, split_pages=None):
    from cerebro_ds_kpgi_b200 import synth
    d = tmp_path / ("gpseg%d" % seg_id) / "base" / "16384"
    d.mkdir(parents=True, exist_ok=True)
    for name, buf in (("50000", heap), ("50001", toast)):
        if split_pages and buf.size > split_pages * 32768:
            step = split_pages * 32768
            for k, o in enumerate(range(0, buf.size, step)):
                (d / (name if k == 0 else "%s.%d" % (name, k))).write_bytes(buf[o:o + step].tobytes())
        else:
            (d / name).write_bytes(buf.tobytes())
    shape_rows = [{"__dist_key__": 0, "buffer_id": k, "independent_var_shape": list(v["independent_var"].shape),
                   "dependent_var_shape": list(v["dependent_var"].shape), "table_name": table} for k, v in expected.items()]
    cat = tmp_path / "sys_cat.dill"
    synth.write_catalog(str(cat), {"train": [dict(gp_segment_id=seg_id, relname=table, relfilenode=50000,
                                                  toast_relfilenode=50001)], "valid": []}, shape_rows)
    return str(d), str(cat)


@pytest.mark.parametrize("name", SEGMENT_CASES)
def test_input_fn_matches_reference(name, tmp_path, torch_cuda, capsys):
    from cerebro_ds_kpgi_b200 import da
    heap, toast, order, expected = load_golden_segment(name)
    d, cat = write_case(tmp_path, heap, toast, expected, split_pages=3 if name == "seg_imagenet_rows" else None)
    da.SYS_CAT_PATH = cat
    out = da.input_fn(os.path.join(d, "train"))
    assert_dict_equal(out, expected, order)
    logged = capsys.readouterr().out
    for marker in ("CURRENT MODE: train", "START READING TABLE", "END READING TOAST", "START DETOASTING", "END DETOASTING"):
        assert marker in logged
    # consumers index like this (run_da_cerebro_standalone.py:38-56, run_pytorchddp_da.py:51-62)
    k0 = next(iter(out))
    assert out[k0]["dependent_var"].shape[0] == out[k0]["independent_var"].shape[0]
    assert out[k0]["independent_var"][0].dtype == np.float32 and out[k0]["dependent_var"][0].dtype == np.int16


def test_two_call_path_matches_reference_frames(tmp_path, torch_cuda):
    """table_page_read -> toast_page_read, the way da.input_fn chains them (da.py:55-57)."""
    from cerebro_ds_kpgi_b200 import pg_page_reader as ppr
    heap, toast, order, expected = load_golden_segment("seg_small_mixed")
    d, cat = write_case(tmp_path, heap, toast, expected)
    import dill
    df_shape = dill.load(open(cat, "rb"))["shape"]
    df_data, df_toast = ppr.table_page_read(os.path.join(d, "50000"))
    z = np.load(os.path.join(GOLDEN, "frames_small_mixed.npz"))
    assert list(df_data.columns) == list(z["data_columns"]) and list(df_toast.columns) == list(z["toast_columns"])
    for c in ("va_rawsize", "va_extsize", "va_valueid", "va_toastrelid", "dist_key", "buffer_id"):
        assert np.array_equal(df_toast[c].values.astype(np.int64), z[c]), c
    assert np.array_equal(df_toast["external"].values.astype(bool), z["external"])
    assert [str(x) for x in df_toast["attname"].values] == [str(x) for x in z["attname"]]
    off = 0
    for b, n in zip(df_toast["bytea"].values, z["bytea_len"]):
        want = z["bytea"][off:off + n].tobytes()
        off += int(n)
        assert (b is None and n == 0) or b == want
    assert np.array_equal(np.array([list(t[:4]) for t in df_data["independent_var_tuple"].values]), z["ind_tuple_headers"])
    assert np.array_equal(np.array([list(t[:4]) for t in df_data["dependent_var_tuple"].values]), z["dep_tuple_headers"])
    out = ppr.toast_page_read(os.path.join(d, "50001"), df_toast, df_shape, "t_train")
    assert_dict_equal(out, expected, order)
    # generator_page(toast=True) yields the reference's chunk stream
    got = list(ppr.generator_page(os.path.join(d, "50001"), toast=True))
    assert [g[0] for g in got] == list(z["chunk_id"]) and [g[1] for g in got] == list(z["chunk_seq"])
    assert b"".join(g[2] for g in got) == z["chunk_data"].tobytes()
    # read_toast_all_bytes + detoast (host chunk bytes, GPU pglz batch)
    grand = ppr.read_toast_all_bytes(os.path.join(d, "50001"), df_toast)
    out2 = ppr.detoast(grand, df_toast, df_shape, "t_train")
    assert_dict_equal(out2, expected, order)


def test_reference_exceptions(tmp_path, torch_cuda):
    """Same exception type and message as the reference raised on the same corrupted files."""
    from cerebro_ds_kpgi_b200 import da
    z = np.load(os.path.join(GOLDEN, "error_cases.npz"))
    types = {"Exception": Exception, "AssertionError": AssertionError, "ValueError": ValueError}
    for i, name in enumerate(z["names"]):
        name, etype, emsg = str(name), str(z["types"][i]), str(z["msgs"][i])
        heap, toast = z["heap_%d" % i], z["toast_%d" % i]
        shape = [1, 40, 40, 4] if name == "shape_mismatch" else [1, 40, 40, 3]
        expected = {k: {"independent_var": np.zeros(shape, np.float32), "dependent_var": np.zeros([1, 10], np.int16)} for k in range(3)}
        sub = tmp_path / name
        sub.mkdir()
        d, cat = write_case(sub, heap[:heap.size - heap.size % 32768], toast[:toast.size - toast.size % 32768], expected)
        if toast.size % 32768:
            with open(os.path.join(d, "50001"), "ab") as f:
                f.write(toast[toast.size - toast.size % 32768:].tobytes())
        da.SYS_CAT_PATH = cat
        with pytest.raises(types[etype]) as ei:
            da.input_fn(os.path.join(d, "train"))
        assert type(ei.value) is types[etype], name
        if etype != "ValueError":
            assert str(ei.value).split("\n")[0] == emsg.split("\n")[0], name


def test_input_fn_device_and_consumers(tmp_path, torch_cuda, coracle):
    from cerebro_ds_kpgi_b200 import da, synth, consumers
    torch = torch_cuda
    ind, dep = synth.imagenet_rows(10, seed=4, kind="normal", images_per_row=2)
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage=["inline_pglz", "ext_raw"] * 5, table_name="t_train")
    d, cat = write_case(tmp_path, seg.heap, seg.toast, seg.expected)
    da.SYS_CAT_PATH = cat
    part = da.input_fn_device(os.path.join(d, "train"))
    want = coracle.decode(seg.heap, seg.toast)
    assert part.images.shape == (20, 112, 112, 3) and part.onehot.shape == (20, 1000) and part.labels.shape == (20,)
    assert part.images.cpu().numpy().tobytes() == want["out_ind"].tobytes()
    assert part.buffer_ids == [int(want["rows"][r, 1]) for r in want["order"]]
    assert np.array_equal(part.labels.cpu().numpy(), np.argmax(want["out_dep"].view(np.int16).reshape(-1, 1000), 1))
    # torch consumer: NCHW float32 + int64 labels, epoch padded by i % actual_length (run_pytorchddp_da.py:51-62)
    ds = consumers.DeviceDADataset(part, length=50, train=True)
    idx = torch.arange(37, 45, device="cuda")
    x, y = ds.batch(idx)
    ref_x = np.stack([np.rollaxis(want["out_ind"].view(np.float32).reshape(20, 112, 112, 3)[i % 20], 2, 0) for i in range(37, 45)])
    assert x.shape == (8, 3, 112, 112) and x.cpu().numpy().tobytes() == ref_x.tobytes()
    assert np.array_equal(y.cpu().numpy(), part.labels.cpu().numpy()[np.arange(37, 45) % 20])
    xb, yb = ds.batch(idx, dtype=torch.bfloat16)          # optional dtype: RNE of the same values
    assert torch.equal(xb.cpu().view(torch.int16), torch.from_numpy(ref_x).to(torch.bfloat16).view(torch.int16))
    assert torch.equal(yb, y)
    # a shape the staged kernel does not take (odd pixel count): generic path, same answer
    from cerebro_ds_kpgi_b200 import segment
    odd = torch.randn(7, 5, 3, 3, device="cuda")
    assert torch.equal(segment.nhwc_to_nchw(odd), odd.permute(0, 3, 1, 2).contiguous())
    big = torch.randn(33, 112, 112, 3, device="cuda")
    assert torch.equal(segment.nhwc_to_nchw(big), big.permute(0, 3, 1, 2).contiguous())
    # Keras-style consumer (run_da_cerebro_standalone.py:33-56)
    out = da.input_fn(os.path.join(d, "train"))
    g = consumers.generator_df(out)
    assert g.length == 20
    first = next(iter(g()))
    k0 = next(iter(out))
    assert first[0].tobytes() == out[k0]["independent_var"][0].tobytes()


def test_prefetch_device_yields_the_same_partitions_in_order(tmp_path, torch_cuda, coracle):
    from cerebro_ds_kpgi_b200 import da, synth
    segs, dirs, shape_rows, train = [], [], [], []
    for seg_id in range(3):
        ind, dep = synth.imagenet_rows(6, seed=40 + seg_id, kind="normal", images_per_row=2)
        seg = synth.build_segment(ind, dep, ind_storage=["ext_pglz", "ext_raw"] * 3, dep_storage="inline_pglz",
                                  buffer_ids=list(range(100 * seg_id, 100 * seg_id + 6)), table_name="t_train")
        d, _ = write_case(tmp_path, seg.heap, seg.toast, seg.expected, seg_id=seg_id)
        segs.append(seg); dirs.append(d)
        shape_rows += [{"__dist_key__": seg_id, "buffer_id": k, "independent_var_shape": list(v["independent_var"].shape),
                        "dependent_var_shape": list(v["dependent_var"].shape), "table_name": "t_train"} for k, v in seg.expected.items()]
        train.append(dict(gp_segment_id=seg_id, relname="t_train", relfilenode=50000, toast_relfilenode=50001))
    cat = str(tmp_path / "sys_cat_all.dill")
    synth.write_catalog(cat, {"train": train, "valid": []}, shape_rows)
    da.SYS_CAT_PATH = cat
    paths = [os.path.join(d, "train") for d in dirs]
    got = list(da.prefetch_device(paths))
    assert [p for p, _ in got] == paths
    for (p, part), seg in zip(got, segs):
        want = coracle.decode(seg.heap, seg.toast)
        assert part.images.cpu().numpy().tobytes() == want["out_ind"].tobytes()
        assert part.buffer_ids == [int(want["rows"][r, 1]) for r in want["order"]]
    # an error in a later partition surfaces when that partition is reached
    with pytest.raises(Exception):
        list(da.prefetch_device(paths[:1] + [os.path.join(str(tmp_path), "gpseg9", "base", "16384", "train")]))


def test_generator_page_debug_and_value_markers_match_the_reference(tmp_path, torch_cuda, capsys):
    """generator_page(debug=True) (ppr:380-381, 415-420, 433-445) and the per-value log markers (ppr:554, 600-603)
    against what the unmodified reference printed and yielded (tests/golden/frames_debug.npz)."""
    import re
    from cerebro_ds_kpgi_b200 import da
    from cerebro_ds_kpgi_b200 import pg_page_reader as ppr
    heap, toast, order, expected = load_golden_segment("seg_small_mixed")
    z = np.load(os.path.join(GOLDEN, "frames_debug.npz"))
    d, cat = write_case(tmp_path, heap, toast, expected)
    # same two-file TOAST relation the golden run used
    cut = int(z["toast_cut_pages"][0]) * 32768
    tp = os.path.join(d, "50001")
    open(tp, "wb").write(toast.tobytes()[:cut])
    open(tp + ".1", "wb").write(toast.tobytes()[cut:])
    capsys.readouterr()
    for name, path, is_toast in (("heap", os.path.join(d, "50000"), False), ("toast", tp, True)):
        recs = list(ppr.generator_page(path, toast=is_toast, debug=True))
        printed = capsys.readouterr().out.splitlines()
        assert printed == [str(x) for x in z[name + "_printed"]], name
        assert len(recs) == len(z[name + "_lp_off"])
        off = 0
        for i, rec in enumerate(recs):
            page_bytes, header, ids, lp_off, item_num, ident, item_header, tupdata = rec[-1]
            assert len(page_bytes) == 32768
            assert list(header) == list(z[name + "_page_header"][i]), (name, i)
            assert len(ids) == z[name + "_ids_len"][i] and item_num == z[name + "_item_num"][i]
            assert (-1 if lp_off is None else lp_off) == z[name + "_lp_off"][i], (name, i)
            assert list(ident) == list(z[name + "_ident"][i]), (name, i)
            assert list(item_header) == list(z[name + "_item_header"][i]), (name, i)
            n = int(z[name + "_tupdata_len"][i])
            assert bytes(tupdata) == z[name + "_tupdata"][off:off + n].tobytes(), (name, i)
            off += n
    # the reader's log, line for line (timestamps stripped), with the per-value markers switched on
    os.remove(tp + ".1")
    open(tp, "wb").write(toast.tobytes())
    da.SYS_CAT_PATH = cat
    ppr.LOG_VALUE_IDS = True
    try:
        out = da.input_fn(os.path.join(d, "train"))
    finally:
        ppr.LOG_VALUE_IDS = False
    assert_dict_equal(out, expected, order)
    stamp = re.compile(r": \d{4}-\d\d-\d\d \d\d:\d\d:\d\d$")
    got = [stamp.sub("", l).replace(str(tmp_path), "<root>") for l in capsys.readouterr().out.splitlines()]
    want = [str(x) for x in z["input_fn_markers"]]
    assert got[:-1] == want[:-1]
    assert got[-1].startswith("ITEM NUM:6, BUFFER_IDs:")
