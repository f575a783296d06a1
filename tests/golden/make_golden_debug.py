"""Golden data for the diagnostic outputs of the reference reader (build container only: needs /root/reference):

    python tests/golden/make_golden_debug.py

* generator_page(filename, toast=..., debug=True) (/root/reference/cerebro_gpdb/pg_page_reader.py:364-448) on the
  small mixed segment: the printed "PAGE: <pos>" lines and, per yielded item, the fields of the debug tuple
  (lp_off, item_num, ItemIdentifier, HeapTupleHeader, tupdata);
* the per-value log markers of toast_page_read (ppr:554, 600-603) with the timestamps stripped.
Writes tests/golden/frames_debug.npz.  The other golden files are not touched.
"""
import contextlib
import io
import os
import re
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import import_reference          # noqa: E402


def main():
    ref_reader, ref_da = import_reference()
    z = np.load(os.path.join(HERE, "seg_small_mixed.npz"))
    root = tempfile.mkdtemp(prefix="golden_dbg_")
    hp, tp = os.path.join(root, "heap"), os.path.join(root, "toast")
    # the TOAST relation split into two files, so that "PAGE:" positions restart (ppr:369-381)
    open(hp, "wb").write(z["heap"].tobytes())
    tb = z["toast"].tobytes()
    cut = (len(tb) // 32768 // 2) * 32768
    open(tp, "wb").write(tb[:cut])
    open(tp + ".1", "wb").write(tb[cut:])
    arrays = {"toast_cut_pages": np.array([cut // 32768])}
    for name, path, toast in (("heap", hp, False), ("toast", tp, True)):
        buf = io.StringIO()
        recs = []
        with contextlib.redirect_stdout(buf):
            for item in ref_reader.generator_page(path, toast=toast, debug=True):
                dbg = item[-1]
                page_bytes, page_header, ids, lp_off, item_num, ident, item_header, tupdata = dbg
                recs.append((list(page_header), len(ids), -1 if lp_off is None else lp_off, item_num, list(ident), list(item_header),
                             bytes(tupdata), hash(bytes(page_bytes)) & 0xFFFFFFFF))
        arrays[name + "_printed"] = np.array(buf.getvalue().splitlines())
        arrays[name + "_page_header"] = np.array([r[0] for r in recs], dtype=np.int64)
        arrays[name + "_ids_len"] = np.array([r[1] for r in recs], dtype=np.int64)
        arrays[name + "_lp_off"] = np.array([r[2] for r in recs], dtype=np.int64)
        arrays[name + "_item_num"] = np.array([r[3] for r in recs], dtype=np.int64)
        arrays[name + "_ident"] = np.array([r[4] for r in recs], dtype=np.int64)
        arrays[name + "_item_header"] = np.array([r[5] for r in recs], dtype=np.int64)
        arrays[name + "_tupdata"] = np.frombuffer(b"".join(r[6] for r in recs), dtype=np.uint8)
        arrays[name + "_tupdata_len"] = np.array([len(r[6]) for r in recs], dtype=np.int64)
    # log markers of the whole reader, timestamps removed
    import dill  # noqa: F401
    from cerebro_ds_kpgi_b200 import synth
    order = [int(k) for k in z["order"]]
    shape_rows = [{"__dist_key__": 0, "buffer_id": k, "independent_var_shape": list(z["ind_%d" % k].shape),
                   "dependent_var_shape": list(z["dep_%d" % k].shape), "table_name": "t_train"} for k in order]
    d = os.path.join(root, "gpseg0", "base", "16384")
    os.makedirs(d)
    open(os.path.join(d, "50000"), "wb").write(z["heap"].tobytes())
    open(os.path.join(d, "50001"), "wb").write(tb)
    synth.write_catalog(os.path.join(root, "sys_cat.dill"),
                        {"train": [dict(gp_segment_id=0, relname="t_train", relfilenode=50000, toast_relfilenode=50001)], "valid": []},
                        shape_rows)
    ref_da.SYS_CAT_PATH = os.path.join(root, "sys_cat.dill")
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        ref_da.input_fn(os.path.join(root, "gpseg0/base/16384/train"))
    stamp = re.compile(r": \d{4}-\d\d-\d\d \d\d:\d\d:\d\d$")
    lines = [stamp.sub("", l).replace(root, "<root>") for l in buf.getvalue().splitlines()]
    arrays["input_fn_markers"] = np.array(lines)
    path = os.path.join(HERE, "frames_debug.npz")
    np.savez_compressed(path, **arrays)
    print("frames_debug: heap items %d, toast items %d, %d log lines -> %d KB" % (
        len(arrays["heap_lp_off"]), len(arrays["toast_lp_off"]), len(lines), os.path.getsize(path) // 1024))


if __name__ == "__main__":
    main()
