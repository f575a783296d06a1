"""Generate tests/golden/*.npz by running the UNMODIFIED reference reader on synthetic segments.

Run in the build container only (it needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

The reference has no golden vectors of its own, so parity is pinned on outputs of
the reference itself: /root/reference/cerebro_gpdb/da.py:29-58 (`input_fn`) for whole segments,
/root/reference/cerebro_gpdb/pg_page_reader.py:191-231 (`pglz_decompress_py`) and the compiled
/root/reference/cerebro_gpdb/pg_lzcompress.c for pglz known-answer vectors, and the reference's own
exception type/message on corrupted inputs.  Nothing under tests/ reads /root/reference at test time;
tests load the .npz files written here.
"""
import contextlib
import io
import os
import struct
import sys
import tempfile
import types
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "oracle"))


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


def run_reference(ref_da, seg, table_name="t_train"):
    from cerebro_ds_kpgi_b200 import synth
    root = tempfile.mkdtemp(prefix="golden_")
    seg.write(root, 0, 16384, 50000, 50001)
    synth.write_catalog(os.path.join(root, "sys_cat.dill"),
                        {"train": [dict(gp_segment_id=0, relname=table_name, relfilenode=50000,
                                        toast_relfilenode=50001)], "valid": []}, seg.shape_rows)
    ref_da.SYS_CAT_PATH = os.path.join(root, "sys_cat.dill")
    with contextlib.redirect_stdout(io.StringIO()):
        return ref_da.input_fn(os.path.join(root, "gpseg0/base/16384/train"))


def save_case(name, seg, out):
    keys = [int(k) for k in out.keys()]
    arrays = {"heap": seg.heap, "toast": seg.toast, "order": np.array(keys, dtype=np.int64)}
    for k, v in out.items():
        arrays["ind_%d" % int(k)] = np.ascontiguousarray(v["independent_var"])
        arrays["dep_%d" % int(k)] = np.ascontiguousarray(v["dependent_var"])
        assert v["independent_var"].dtype == np.float32 and v["dependent_var"].dtype == np.int16
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **arrays)
    print("%-28s rows=%d heap_pages=%d toast_pages=%d -> %d KB" % (
        name, len(keys), seg.heap.size // 32768, seg.toast.size // 32768, os.path.getsize(path) // 1024))


def segment_cases(ref_da):
    from cerebro_ds_kpgi_b200 import synth
    # 1. small mixed: every storage form the reference can read, B=2
    ind, dep = synth.imagenet_rows(6, seed=11, shape=(16, 16, 3), classes=10, images_per_row=2)
    seg = synth.build_segment(ind, dep,
                              ind_storage=["ext_pglz", "ext_literal", "ext_pglz", "ext_pglz", "ext_literal", "ext_pglz"],
                              dep_storage=["inline_pglz", "ext_pglz", "inline_literal", "ext_literal", "inline_pglz", "ext_pglz"],
                              table_name="t_train", buffer_ids=[7, 3, 11, 0, 5, 9])
    save_case("seg_small_mixed", seg, run_reference(ref_da, seg))

    # 2. one real-shape ImageNet row pair (112x112x3 f32 + int16[1,1000]), compressible + incompressible
    ind_a, dep_a = synth.imagenet_rows(1, seed=2018, kind="norm_u8")
    ind_b, dep_b = synth.imagenet_rows(1, seed=2019, kind="normal")
    seg = synth.build_segment(ind_a + ind_b, dep_a + dep_b, ind_storage=["ext_pglz", "ext_literal"],
                              dep_storage="inline_pglz", table_name="t_train")
    save_case("seg_imagenet_rows", seg, run_reference(ref_da, seg))

    # 3. Criteo-shaped buffer (B=8 to stay small)
    ind, dep = synth.criteo_buffers(2, seed=5, rows_per_buffer=8)
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="inline_pglz", table_name="t_train")
    save_case("seg_criteo", seg, run_reference(ref_da, seg))

    # 4. many small rows: multi-page heap, many chunks per TOAST page, TOAST order != table order
    ind, dep = synth.imagenet_rows(300, seed=3, shape=(4, 4, 3), classes=5, images_per_row=3)
    st_dep = ["inline_pglz" if i % 3 else "ext_pglz" for i in range(300)]
    n_ext = 300 + sum(1 for s in st_dep if s.startswith("ext"))
    perm = list(np.random.default_rng(4).permutation(n_ext))
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage=st_dep, table_name="t_train",
                              toast_perm=perm, buffer_ids=list(range(1000, 1300)))
    save_case("seg_many_shuffled", seg, run_reference(ref_da, seg))

    # 5. multi-chunk values with small chunk size written in scrambled value order
    ind, dep = synth.imagenet_rows(5, seed=8, shape=(40, 40, 3), classes=30, images_per_row=1, kind="normal")
    seg = synth.build_segment(ind, dep, ind_storage="ext_literal", dep_storage="ext_literal", table_name="t_train",
                              toast_perm=[9, 0, 3, 2, 7, 6, 5, 4, 1, 8])
    save_case("seg_multichunk_shuffled", seg, run_reference(ref_da, seg))


def frame_case(ref_reader):
    """df_data / df_toast of table_page_read (ppr:451-494) and the chunk stream of
    generator_page(toast=True) (ppr:385-422) for the small mixed segment."""
    z = np.load(os.path.join(HERE, "seg_small_mixed.npz"))
    root = tempfile.mkdtemp(prefix="golden_")
    hp, tp = os.path.join(root, "heap"), os.path.join(root, "toast")
    open(hp, "wb").write(z["heap"].tobytes())
    open(tp, "wb").write(z["toast"].tobytes())
    with contextlib.redirect_stdout(io.StringIO()):
        df_data, df_toast = ref_reader.table_page_read(hp)
        chunks = [(a, b, bytes(c)) for a, b, c, _ in ref_reader.generator_page(tp, toast=True)]
    bytea = [b if b is not None else b"" for b in df_toast["bytea"].values]
    arrays = {c: df_toast[c].values.astype(np.int64) for c in
              ("va_rawsize", "va_extsize", "va_valueid", "va_toastrelid", "dist_key", "buffer_id")}
    arrays["external"] = df_toast["external"].values.astype(np.bool_)
    arrays["attname"] = np.array([str(x) for x in df_toast["attname"].values])
    arrays["bytea"] = np.frombuffer(b"".join(bytea), dtype=np.uint8)
    arrays["bytea_len"] = np.array([len(b) for b in bytea], dtype=np.int64)
    arrays["data_columns"] = np.array(list(df_data.columns))
    arrays["toast_columns"] = np.array(list(df_toast.columns))
    arrays["ind_tuple_headers"] = np.array([list(t[:4]) for t in df_data["independent_var_tuple"].values], dtype=np.int64)
    arrays["dep_tuple_headers"] = np.array([list(t[:4]) for t in df_data["dependent_var_tuple"].values], dtype=np.int64)
    arrays["chunk_id"] = np.array([c[0] for c in chunks], dtype=np.int64)
    arrays["chunk_seq"] = np.array([c[1] for c in chunks], dtype=np.int64)
    arrays["chunk_len"] = np.array([len(c[2]) for c in chunks], dtype=np.int64)
    arrays["chunk_data"] = np.frombuffer(b"".join(c[2] for c in chunks), dtype=np.uint8)
    path = os.path.join(HERE, "frames_small_mixed.npz")
    np.savez_compressed(path, **arrays)
    print("frames_small_mixed: %d df_toast rows, %d chunks -> %d KB" % (len(df_toast), len(chunks), os.path.getsize(path) // 1024))


def text_etl_cases():
    """Outputs of the reference's own text parsers (etl_imagenet.py:30-37, etl_criteo.py:28-37) on synthetic
    gpfdist TEXT unload lines.  h5py / tqdm are stubbed: the parsers do not use them."""
    for m in ("h5py", "tqdm"):
        sys.modules.setdefault(m, types.ModuleType(m))
    sys.path.insert(0, "/root/reference/cerebro_gpdb")
    import etl_criteo
    import etl_imagenet
    rng = np.random.default_rng(99)
    fmts = ["%.6g", "%.9g", "%.3f", "%d", "%.6e", "%.15g"]
    lines = []
    for r in range(40):                                       # ImageNet style: nested 4x3x3, int label
        vals = rng.standard_normal(36) * 10.0 ** rng.integers(-5, 9, 36)
        vals[rng.integers(0, 36, 4)] = 0.0
        toks = []
        for v in vals:
            f = fmts[int(rng.integers(0, len(fmts)))]
            toks.append(f % (int(v) if f == "%d" else v))
        rows = ["{" + ",".join("{" + ",".join(toks[(i * 3 + j) * 3:(i * 3 + j) * 3 + 3]) + "}" for j in range(3)) + "}" for i in range(4)]
        lines.append("%d|{%s}|%d\n" % (1000 + r, ",".join(rows), int(rng.integers(0, 1000))))
    lines.append("7|{{{-0,+1.5,1e22},{123456789012345,0.000001,9007199254740991}," + "{1e-22,2.5e+10,-3E5}},{{1,2,3},{4,5,6},{7,8,9}},"
                 "{{1,2,3},{4,5,6},{7,8,9}},{{1,2,3},{4,5,6},{7,8,9}}}|0")      # edge literals, no trailing newline
    text_a = "".join(lines)
    parsed = [etl_imagenet.parser(l) for l in text_a.splitlines(True)]
    feats_a = np.stack([p[1] for p in parsed]).astype(np.float32)
    assert feats_a.shape == (41, 4, 3, 3)
    lines = []
    for r in range(30):                                       # Criteo style: flat 0/1 floats, digit-string label
        v = (rng.random(64) < 0.1).astype(int)
        lines.append("%d|{%s}|%s\n" % (r, ",".join(str(x) for x in v), "01" if rng.random() < 0.3 else "10"))
    text_b = "".join(lines)
    parsed_b = [etl_criteo.parser(l) for l in text_b.splitlines(True)]
    path = os.path.join(HERE, "text_etl.npz")
    np.savez_compressed(path,
                        text_a=np.frombuffer(text_a.encode(), dtype=np.uint8), ids_a=np.array([int(p[0]) for p in parsed]),
                        feats_a=feats_a, labels_a=np.array([p[2] for p in parsed], dtype=np.int64),
                        text_b=np.frombuffer(text_b.encode(), dtype=np.uint8), ids_b=np.array([int(p[0]) for p in parsed_b]),
                        feats_b=np.array([p[1] for p in parsed_b], dtype=np.float32),
                        labels_b=np.array([p[2] for p in parsed_b], dtype=np.int64))
    print("text_etl: %d + %d lines -> %d KB" % (len(parsed), len(parsed_b), os.path.getsize(path) // 1024))


def pglz_cases(ref_reader):
    """Known-answer vectors: (varlena incl. 8-byte header) -> bytes, decoded by BOTH reference decoders."""
    import ctypes
    from cerebro_ds_kpgi_b200 import build, synth
    build.build_oracle()
    ref_c = ctypes.CDLL(build.LIB_REF_PGLZ)
    c_fn = ref_reader.wrap_function(ref_c, "pglz_decompress", None, [ctypes.c_char_p, ctypes.c_char_p])
    rng = np.random.default_rng(77)
    raws = []
    for n in list(range(0, 20)) + [31, 32, 33, 255, 256, 272, 273, 274, 300, 1000, 4095, 4096, 4097, 9000]:
        raws.append(rng.integers(0, 256, size=n, dtype=np.uint8).tobytes())          # incompressible
        raws.append(bytes(n))                                                        # zeros: overlapping matches, max length
        raws.append((b"abcabcabd" * (n // 9 + 1))[:n])                               # short period
        raws.append(rng.integers(0, 4, size=n, dtype=np.uint8).tobytes())            # low entropy
    text = (b"the quick brown fox jumps over the lazy dog. " * 300)
    raws.append(text)
    lut_img, _ = synth.imagenet_rows(1, seed=9, shape=(24, 24, 3), classes=3)
    raws.append(lut_img[0].tobytes())
    raws.append(np.repeat(rng.integers(0, 256, size=700, dtype=np.uint8), 13).tobytes())   # runs of 13
    far = rng.integers(0, 256, size=4095, dtype=np.uint8).tobytes()
    raws.append(far + far[:600])                                                     # offset 4095 exactly
    blobs, offs, raw_blobs, raw_offs = [], [0], [], [0]
    for raw in raws:
        for mode in ("lz", "literal"):
            v = synth.pglz_varlena(np.frombuffer(raw, dtype=np.uint8), mode, effort=16)
            d_py = ref_reader.pglz_decompress_py(v, bytearray(len(raw)))
            d_c = ctypes.create_string_buffer(max(len(raw), 1))
            c_fn(v, d_c)
            assert bytes(d_py) == raw, "reference python decoder disagrees with encoder input"
            assert d_c.raw[:len(raw)] == raw, "reference C decoder disagrees with encoder input"
            blobs.append(v)
            offs.append(offs[-1] + len(v))
            raw_blobs.append(raw)
            raw_offs.append(raw_offs[-1] + len(raw))
    path = os.path.join(HERE, "pglz_kat.npz")
    np.savez_compressed(path, src=np.frombuffer(b"".join(blobs), dtype=np.uint8), src_off=np.array(offs, dtype=np.int64),
                        raw=np.frombuffer(b"".join(raw_blobs), dtype=np.uint8), raw_off=np.array(raw_offs, dtype=np.int64))
    print("pglz_kat: %d vectors -> %d KB" % (len(blobs), os.path.getsize(path) // 1024))


def error_cases(ref_reader, ref_da):
    """Corrupted inputs: record the reference's exception type and message."""
    from cerebro_ds_kpgi_b200 import synth
    ind, dep = synth.imagenet_rows(3, seed=21, shape=(40, 40, 3), classes=10, images_per_row=1, kind="normal")

    def fresh(**kw):
        return synth.build_segment(ind, dep, ind_storage="ext_literal", dep_storage=kw.pop("dep", "inline_pglz"),
                                   table_name="t_train", **kw)
    out_names, heaps, toasts, types_, msgs = [], [], [], [], []

    def record(name, seg):
        try:
            run_reference(ref_da, seg)
            et, em = "", ""
        except BaseException as e:  # noqa
            et, em = type(e).__name__, str(e)
        out_names.append(name); heaps.append(seg.heap.copy()); toasts.append(seg.toast.copy())
        types_.append(et); msgs.append(em)
        print("error case %-24s -> %s(%r)" % (name, et, em[:60]))

    # a) corrupt pglz stream of an external value: overwrite a literal group's control byte with 0xFF
    seg = fresh()
    # first TOAST tuple physically on page 0 is the LAST inserted; use ItemId 0 -> first inserted chunk (seq 0)
    iid = struct.unpack_from("<I", seg.toast, 24)[0]
    lp_off = iid & 0x7FFF
    payload = lp_off + 24 + 8 + 4
    seg.toast[payload + 4] = 0xFF          # payload[0:4] is rawsize; payload[4] the first control byte
    record("pglz_corrupt_external", seg)

    # b) corrupt inline stream in the heap tuple (dependent_var)
    seg = fresh()
    iid = struct.unpack_from("<I", seg.heap, 24)[0]
    lp_off = iid & 0x7FFF
    td = lp_off + 24
    seg.heap[td + 24 + 8] ^= 0xFF           # first control byte of the inline stream
    record("pglz_corrupt_inline", seg)

    # c) partial trailing page
    seg = fresh()
    seg.toast = np.concatenate([seg.toast, np.zeros(100, dtype=np.uint8)])
    record("partial_page", seg)

    # d) pd_special != BLCKSZ on a TOAST page
    seg = fresh()
    struct.pack_into("<H", seg.toast, 16, 32760)
    record("toast_special", seg)

    # e) pd_lower not a multiple of 4 past the header
    seg = fresh()
    struct.pack_into("<H", seg.heap, 12, struct.unpack_from("<H", seg.heap, 12)[0] + 2)
    record("itemid_misaligned", seg)

    # f) pointer claims one more chunk than stored (va_extsize bumped past a chunk boundary)
    seg = fresh()
    iid = struct.unpack_from("<I", seg.heap, 24)[0]
    td = (iid & 0x7FFF) + 24
    ext = struct.unpack_from("<i", seg.heap, td + 4 + 8)[0]
    struct.pack_into("<i", seg.heap, td + 4 + 8, ext + 8140)
    record("chunk_count", seg)

    # g) last chunk shorter than the pointer says (same chunk count)
    seg = fresh()
    iid = struct.unpack_from("<I", seg.heap, 24)[0]
    td = (iid & 0x7FFF) + 24
    ext = struct.unpack_from("<i", seg.heap, td + 4 + 8)[0]
    struct.pack_into("<i", seg.heap, td + 4 + 8, ext + 1)
    record("chunk_size_last", seg)

    # h) a middle chunk that is not full size
    seg = synth.build_segment(ind, dep, ind_storage="ext_literal", dep_storage="inline_pglz", table_name="t_train")
    iid = struct.unpack_from("<I", seg.toast, 24 + 4)[0]        # second inserted tuple: seq 1 of the first value
    lp_off = iid & 0x7FFF
    struct.pack_into(">I", seg.toast, lp_off + 24 + 8, 4 + 8139)
    record("chunk_size_middle", seg)

    # i) shape catalog disagrees with the decoded size
    seg = fresh()
    seg.shape_rows[0]["independent_var_shape"] = [1, 40, 40, 4]
    record("shape_mismatch", seg)

    path = os.path.join(HERE, "error_cases.npz")
    arrays = {"names": np.array(out_names), "types": np.array(types_), "msgs": np.array(msgs)}
    for i, (h, t) in enumerate(zip(heaps, toasts)):
        arrays["heap_%d" % i] = h
        arrays["toast_%d" % i] = t
    np.savez_compressed(path, **arrays)
    print("error_cases -> %d KB" % (os.path.getsize(path) // 1024))


if __name__ == "__main__":
    ref_reader, ref_da = import_reference()
    segment_cases(ref_da)
    frame_case(ref_reader)
    pglz_cases(ref_reader)
    error_cases(ref_reader, ref_da)
    text_etl_cases()
