"""Time the UNMODIFIED reference reader on BASELINE.json configs[0] (BASELINE.md §3 "reference-py").

Build-container only: it imports /root/reference (absent on the GPU box).

    python tools/reference_py_config0.py [--toast-mib 1024] [--out profiles/r2_reference_py_config0.json]

Workload: one synthetic segment with a 1 GiB TOAST file of ImageNet-shaped rows (112x112x3 float32, external,
pglz-compressed; int16[1000] one-hot, inline pglz), SURVEY.md §8(d) "Config 1".  Runs
cerebro_gpdb.da.input_fn (/root/reference/cerebro_gpdb/da.py:29-58) in ONE process with
PGLZ_IMPLEMENTATION = 'py' (pg_page_reader.py:46), as the reference ships it.  Phase times come from the
reference's own START/END markers (pg_page_reader.py:452,493,600-603,616-623): the module-level name `logs` is
wrapped so that every marker is also stamped with perf_counter (the reference prints whole seconds only); no
reference source is changed.  The result is checked against the generator's arrays.
"""
import argparse
import contextlib
import io
import json
import os
import shutil
import sys
import tempfile
import time
import types
import warnings

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def import_reference():
    ps = types.ModuleType("psycopg2")
    ps.extensions = types.SimpleNamespace(POLL_OK=0, POLL_READ=1, POLL_WRITE=2)
    ps.OperationalError = Exception
    sys.modules["psycopg2"] = ps
    sys.modules["psycopg2.extensions"] = ps.extensions
    sys.path.insert(0, "/root/reference")
    warnings.simplefilter("ignore")
    import cerebro_gpdb.da as ref_da
    import cerebro_gpdb.pg_page_reader as ref_reader
    return ref_reader, ref_da


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--toast-mib", type=int, default=1024)
    ap.add_argument("--out", default=os.path.join(REPO, "profiles", "r2_reference_py_config0.json"))
    args = ap.parse_args()
    from cerebro_ds_kpgi_b200 import synth

    # rows whose compressed payloads fill the TOAST file (ratio ~0.67 for the normalised-uint8 floats)
    n = int(synth.rows_for_toast_bytes(args.toast_mib << 20, dep_external=False) / 0.675)
    t0 = time.time()
    ind, dep = synth.imagenet_rows(n, seed=2018, kind="norm_u8")
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="inline_pglz", dist_key=0, effort=4,
                              table_name="t_train")
    print("segment: %d rows, heap %d pages, TOAST %d pages (%.1f MiB), stored/raw = %.3f, built in %.0f s" % (
        n, seg.heap.size // 32768, seg.toast.size // 32768, seg.toast.size / 2 ** 20, seg.stored_bytes / seg.raw_bytes,
        time.time() - t0), flush=True)

    root = tempfile.mkdtemp(prefix="refpy_")
    try:
        seg.write(root, 0, 16384, 50000, 50001)
        synth.write_catalog(os.path.join(root, "sys_cat.dill"),
                            {"train": [dict(gp_segment_id=0, relname="t_train", relfilenode=50000, toast_relfilenode=50001)],
                             "valid": []}, seg.shape_rows)
        ref_reader, ref_da = import_reference()
        assert ref_reader.PGLZ_IMPLEMENTATION == "py"
        ref_da.SYS_CAT_PATH = os.path.join(root, "sys_cat.dill")
        marks = []
        orig_logs = ref_reader.logs

        def stamped(message):
            if not message.startswith("ID:"):
                marks.append((message.split(":")[0].strip(), time.perf_counter()))
            return orig_logs(message)

        ref_reader.logs = stamped
        with contextlib.redirect_stdout(io.StringIO()):
            t_start = time.perf_counter()
            out = ref_da.input_fn(os.path.join(root, "gpseg0/base/16384/train"))
            t_end = time.perf_counter()
    finally:
        shutil.rmtree(root, ignore_errors=True)

    assert len(out) == n
    for k, v in out.items():
        e = seg.expected[int(k)]
        assert v["independent_var"].tobytes() == e["independent_var"].tobytes()
        assert v["dependent_var"].tobytes() == e["dependent_var"].tobytes()

    def phase(a, b):
        ta = [t for m, t in marks if m == a]
        tb = [t for m, t in marks if m == b]
        return (tb[-1] - ta[0]) if ta and tb else None

    dec_starts = [t for m, t in marks if m == "START DECOMPRESSING"]
    dec_ends = [t for m, t in marks if m == "END DECOMPRESSING"]
    decompress_s = float(sum(e - s for s, e in zip(dec_starts, dec_ends)))
    total = t_end - t_start
    res = {
        "what": "unmodified cerebro_gpdb.da.input_fn, one process, PGLZ_IMPLEMENTATION='py' (BASELINE.md §3 reference-py), "
                "BASELINE.json configs[0]; measured in the build container (CPU only), not on the GPU box",
        "rows": n, "toast_mib": seg.toast.size >> 20, "heap_pages": seg.heap.size // 32768,
        "decoded_bytes": int(seg.raw_bytes), "stored_over_raw": seg.stored_bytes / seg.raw_bytes,
        "total_s": total, "rows_per_s": n / total, "decoded_mb_per_s": seg.raw_bytes / total / 1e6,
        "phases_s": {"read_table": phase("START READING TABLE", "END READING TABLE"),
                     "read_toast": phase("START READING TOAST", "END READING TOAST"),
                     "detoast": phase("START DETOASTING", "END DETOASTING"),
                     "pglz_inside_detoast (external values: sum of START/END DECOMPRESSING pairs)": decompress_s},
        "external_values_decompressed": len(dec_ends),
        "python": sys.version.split()[0], "numpy": np.__version__, "cpu_count": os.cpu_count(),
        "bit_exact_vs_generator": True,
    }
    with open(args.out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
