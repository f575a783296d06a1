"""Synthetic Greenplum-5 segment files for the Direct-Access reader.

The reference has no test data and no writer: its inputs were produced by a live
Greenplum + MADlib `training_preprocessor_dl` (/root/reference/cerebro_gpdb/load_imagenet.py:118-135).
This module writes files in the byte layout the reference reader parses (SURVEY.md Appendix A;
/root/reference/cerebro_gpdb/pg_page_reader.py:253-269,285-305,328-351,385-422) so that the
unmodified reference, the oracle and the CUDA path can all be run on identical input.

It is input generation only — nothing here is on the decode path.  The heavy lifting (page
packing, pglz encoding) is in csrc/synth_host.c, loaded through ctypes.
"""
import ctypes
import os
import struct

import numpy as np

from . import build as _build

BLCKSZ = 32768
TOAST_MAX_CHUNK_SIZE = 8140
IMAGENET_SHAPE = (112, 112, 3)      # /root/reference/cerebro_gpdb/imagenetcat.py:17
IMAGENET_CLASSES = 1000             # imagenetcat.py:18
CRITEO_FEATURES = 7306              # /root/reference/cerebro_gpdb/criteocat.py:15
CRITEO_CLASSES = 2                  # criteocat.py:16
SEED = 2018                         # imagenetcat.py:16

_lib = None


class _Rel(ctypes.Structure):
    _fields_ = [("buf", ctypes.c_void_p), ("cap_pages", ctypes.c_uint64), ("n_pages", ctypes.c_uint64),
                ("natts", ctypes.c_uint32), ("xmin", ctypes.c_uint32)]


def lib():
    global _lib
    if _lib is None:
        path = _build.build_synth()
        L = ctypes.CDLL(path)
        L.cds_synth_rel_append.restype = ctypes.c_int
        L.cds_synth_rel_append.argtypes = [ctypes.POINTER(_Rel), ctypes.c_void_p, ctypes.c_uint32]
        L.cds_synth_toast_value.restype = ctypes.c_int
        L.cds_synth_toast_value.argtypes = [ctypes.POINTER(_Rel), ctypes.c_uint32, ctypes.c_void_p,
                                            ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32]
        L.cds_synth_pglz_literal.restype = ctypes.c_int64
        L.cds_synth_pglz_literal.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64]
        L.cds_synth_pglz_encode.restype = ctypes.c_int64
        L.cds_synth_pglz_encode.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64,
                                            ctypes.c_int]
        L.cds_synth_norm_lut.restype = None
        L.cds_synth_norm_lut.argtypes = [ctypes.c_void_p]
        L.cds_synth_apply_lut.restype = None
        L.cds_synth_apply_lut.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]
        _lib = L
    return _lib


def _as_u8(data):
    a = np.frombuffer(data, dtype=np.uint8) if isinstance(data, (bytes, bytearray, memoryview)) else data
    a = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
    return a


def pglz_stream(raw, mode="lz", effort=8):
    """Encode `raw` into a pglz item stream (no header).  mode: 'lz' or 'literal'."""
    src = _as_u8(raw)
    n = src.size
    cap = n + n // 8 + 16
    dst = np.empty(cap, dtype=np.uint8)
    if mode == "literal":
        m = lib().cds_synth_pglz_literal(src.ctypes.data, n, dst.ctypes.data, cap)
    elif mode == "lz":
        m = lib().cds_synth_pglz_encode(src.ctypes.data, n, dst.ctypes.data, cap, effort)
    else:
        raise ValueError(mode)
    if m < 0:
        raise RuntimeError("pglz encode overflow")
    return dst[:m].tobytes()


def pglz_varlena(raw, mode="lz", effort=8):
    """Full inline-compressed varlena: BE header (0x40000000 | total size), LE rawsize, stream
    (the form pglz_decompress takes, pg_page_reader.py:191-197, pg_lzcompress.c:34-42)."""
    raw = _as_u8(raw)
    s = pglz_stream(raw, mode, effort)
    return struct.pack(">I", 0x40000000 | (8 + len(s))) + struct.pack("<I", raw.size) + s


class RelWriter:
    """One relation file being filled page by page (heap or TOAST)."""

    def __init__(self, cap_pages, natts, xmin=1000):
        self.buf = np.zeros(max(1, int(cap_pages)) * BLCKSZ, dtype=np.uint8)
        self.rel = _Rel(self.buf.ctypes.data, max(1, int(cap_pages)), 0, natts, xmin)

    def append(self, data):
        d = _as_u8(data)
        if lib().cds_synth_rel_append(ctypes.byref(self.rel), d.ctypes.data, d.size) != 0:
            raise RuntimeError("relation capacity exceeded")

    def toast_value(self, valueid, payload, chunk_size=TOAST_MAX_CHUNK_SIZE, first_seq=0):
        p = _as_u8(payload)
        if lib().cds_synth_toast_value(ctypes.byref(self.rel), int(valueid), p.ctypes.data, p.size,
                                       int(chunk_size), int(first_seq)) != 0:
            raise RuntimeError("relation capacity exceeded")

    @property
    def n_pages(self):
        return int(self.rel.n_pages)

    def pages(self):
        return self.buf[: self.n_pages * BLCKSZ]


def toast_pointer(va_rawsize, va_extsize, va_valueid, va_toastrelid):
    """20-byte external TOAST pointer as the reference unpacks it ('@BBBBiiII',
    pg_page_reader.py:337,342): 0x80, length byte, two pad bytes, then varatt_external."""
    return struct.pack("<BBBBiiII", 0x80, 20, 0, 0, va_rawsize, va_extsize, va_valueid, va_toastrelid)


def _int_array_column(values):
    """An int[] datum (the *_shape columns).  Never read by the reader; realistic filler."""
    body = struct.pack("<iiIii", 1, 0, 23, len(values), 1) + struct.pack("<%di" % len(values), *values)
    return struct.pack(">I", 4 + len(body)) + body


class SyntheticSegment:
    """Result of build_segment(): page buffers plus what a correct reader must return."""

    def __init__(self):
        self.heap = None            # np.uint8, multiple of 32768
        self.toast = None
        self.rows = []              # per heap tuple, table order: dicts
        self.expected = {}          # buffer_id -> {'independent_var': arr, 'dependent_var': arr}
        self.shape_rows = []        # for the catalog 'shape' frame
        self.raw_bytes = 0
        self.stored_bytes = 0

    def write(self, root, seg_id, dboid, relfilenode, toast_relfilenode, split_pages=None):
        """Write <root>/gpseg<seg>/base/<dboid>/<relfilenode> (+ .1, .2 … when split_pages)."""
        d = os.path.join(root, "gpseg%d" % seg_id, "base", str(dboid))
        os.makedirs(d, exist_ok=True)
        for name, buf in ((str(relfilenode), self.heap), (str(toast_relfilenode), self.toast)):
            path = os.path.join(d, name)
            for stale in [p for p in os.listdir(d) if p == name or p.startswith(name + ".")]:
                os.remove(os.path.join(d, stale))
            if split_pages and buf.size > split_pages * BLCKSZ:
                step = split_pages * BLCKSZ
                for k, o in enumerate(range(0, buf.size, step)):
                    with open(path if k == 0 else "%s.%d" % (path, k), "wb") as f:
                        f.write(buf[o:o + step].tobytes())
            else:
                with open(path, "wb") as f:
                    f.write(buf.tobytes())
        return d


def build_segment(ind_arrays, dep_arrays, ind_storage="ext_raw", dep_storage="inline_pglz", dist_key=0,
                  first_valueid=20000, toastrelid=16500, effort=8, buffer_ids=None,
                  chunk_size=TOAST_MAX_CHUNK_SIZE, table_name="synthetic_packed", keep_expected=True,
                  toast_cap_pages=None, toast_perm=None):
    """Pack rows (independent_var f32[B,...], dependent_var int16[B,C]) into heap + TOAST pages.

    Storage per value (a string, or a sequence with one entry per row):
      'ext_pglz'     external, compressed by the hash-chain encoder
      'ext_literal'  external, literal-only pglz stream (readable by the unmodified reference)
      'ext_raw'      external, uncompressed (PostgreSQL semantics; the reference cannot read it,
                     SURVEY.md §8c oracle extension)
      'inline_pglz' / 'inline_literal'   dependent_var only: compressed varlena inside the heap tuple
    toast_perm: optional permutation of the external values' write order into the TOAST file
    (default: the order the heap rows reference them), to make physical order differ from table order.
    """
    n = len(ind_arrays)
    assert len(dep_arrays) == n
    ind_st = [ind_storage] * n if isinstance(ind_storage, str) else list(ind_storage)
    dep_st = [dep_storage] * n if isinstance(dep_storage, str) else list(dep_storage)
    if buffer_ids is None:
        buffer_ids = list(range(n))

    raw_total = sum(int(a.nbytes) for a in ind_arrays) + sum(int(a.nbytes) for a in dep_arrays)
    if toast_cap_pages is None:
        n_chunks = sum(-(-int(a.nbytes * 1.13 + 16) // chunk_size) + 1 for a in ind_arrays) + 2 * n
        toast_cap_pages = n_chunks // 3 + 8 + raw_total // BLCKSZ // 8
    heap_cap_pages = 4 + n * 2400 // BLCKSZ + sum(int(a.nbytes) for a, s in zip(dep_arrays, dep_st)
                                                    if s.startswith("inline")) * 2 // BLCKSZ
    heap = RelWriter(heap_cap_pages, natts=6)
    toast = RelWriter(toast_cap_pages, natts=3)
    seg = SyntheticSegment()
    valueid = first_valueid
    pending = []

    encoded = {}                     # (id(array), storage) -> payload: callers may pass the same array for many rows

    def store_external(arr, how):
        nonlocal valueid
        raw = _as_u8(arr)
        if how == "ext_raw":
            payload, rawsize = raw, raw.size
        else:
            key = (id(arr), how)
            if key not in encoded:
                stream = pglz_stream(raw, "literal" if how == "ext_literal" else "lz", effort)
                encoded[key] = np.frombuffer(struct.pack("<I", raw.size) + stream, dtype=np.uint8)
            payload = encoded[key]
            rawsize = raw.size
        vid = valueid
        valueid += 1
        if toast_perm is None:
            toast.toast_value(vid, payload, chunk_size)
        else:
            pending.append((vid, payload))
        seg.stored_bytes += payload.size
        return toast_pointer(rawsize + 4, payload.size, vid, toastrelid), vid

    for r in range(n):
        ind, dep = ind_arrays[r], dep_arrays[r]
        assert ind.dtype in (np.float32, np.uint8) and dep.dtype == np.int16
        row = {"buffer_id": int(buffer_ids[r]), "dist_key": int(dist_key),
               "ind_storage": ind_st[r], "dep_storage": dep_st[r]}
        if not ind_st[r].startswith("ext_"):
            raise ValueError("independent_var must be external (pg_page_reader.py:348)")
        ind_ptr, row["ind_valueid"] = store_external(ind, ind_st[r])
        if dep_st[r].startswith("inline"):
            v = pglz_varlena(dep, "literal" if dep_st[r] == "inline_literal" else "lz", effort)
            dep_field = v + b"\0" * (-len(v) % 4)
            seg.stored_bytes += len(v)
            row["dep_valueid"] = None
        else:
            dep_field, row["dep_valueid"] = store_external(dep, dep_st[r])
        data = (struct.pack("<I", dist_key) + ind_ptr + dep_field +
                _int_array_column(list(ind.shape)) + _int_array_column(list(dep.shape)) +
                struct.pack("<I", int(buffer_ids[r])))
        heap.append(data)
        seg.rows.append(row)
        seg.shape_rows.append({"__dist_key__": int(dist_key), "buffer_id": int(buffer_ids[r]),
                               "independent_var_shape": list(ind.shape),
                               "dependent_var_shape": list(dep.shape), "table_name": table_name})
        if keep_expected:
            seg.expected[int(buffer_ids[r])] = {"independent_var": ind, "dependent_var": dep}
    if toast_perm is not None:
        assert sorted(toast_perm) == list(range(len(pending)))
        for k in toast_perm:
            toast.toast_value(pending[k][0], pending[k][1], chunk_size)
    seg.raw_bytes = raw_total
    seg.heap = heap.pages().copy() if heap.n_pages * BLCKSZ < heap.buf.size // 2 else heap.pages()
    seg.toast = toast.pages()
    seg.table_name = table_name
    return seg


# --------------------------------------------------------------------------------------------
# Value distributions of SURVEY.md §8(d)
# --------------------------------------------------------------------------------------------

def norm_lut():
    lut = np.empty((3, 256), dtype=np.float32)
    lib().cds_synth_norm_lut(lut.ctypes.data)
    return lut


def imagenet_rows(n_rows, seed=SEED, kind="norm_u8", images_per_row=1, shape=IMAGENET_SHAPE,
                  classes=IMAGENET_CLASSES):
    """ImageNet-shaped rows.  kind 'norm_u8': uint8 ~ U[0,255] through the reference normalisation
    (compressible: 256 distinct bit patterns per channel); 'normal': standard normal f32
    (incompressible).  Labels ~ U[0, classes) as int16 one-hot."""
    rng = np.random.default_rng(seed)
    B = images_per_row
    ind, dep = [], []
    lut = norm_lut()
    for _ in range(n_rows):
        if kind == "norm_u8":
            px = rng.integers(0, 256, size=(B,) + tuple(shape), dtype=np.uint8)
            if shape[-1] == 3:
                img = np.empty(px.shape, dtype=np.float32)
                lib().cds_synth_apply_lut(px.ctypes.data, px.size // 3, lut.ctypes.data, img.ctypes.data)
            else:
                img = lut[0][px]
        elif kind == "normal":
            img = rng.standard_normal(size=(B,) + tuple(shape), dtype=np.float32)
        else:
            raise ValueError(kind)
        lab = rng.integers(0, classes, size=B)
        oh = np.zeros((B, classes), dtype=np.int16)
        oh[np.arange(B), lab] = 1
        ind.append(img)
        dep.append(oh)
    return ind, dep


def criteo_buffers(n_buffers, seed=SEED, rows_per_buffer=256, features=CRITEO_FEATURES):
    """Criteo-shaped buffers: 13 bucket groups of 50 + 26 hash groups of 256, one hot each
    (/root/reference/cerebro_gpdb/preprocessing/criteo/preprocessing_criteo.py:50-58,88-106);
    labels one-hot int16[B,2] with P(1) ~ 0.03."""
    rng = np.random.default_rng(seed)
    groups = [50] * 13 + [256] * 26
    assert sum(groups) == CRITEO_FEATURES
    starts = np.cumsum([0] + groups[:-1])
    B = rows_per_buffer
    ind, dep = [], []
    for _ in range(n_buffers):
        x = np.zeros((B, features), dtype=np.float32)
        for g, s in zip(groups, starts):
            if s + g <= features:
                x[np.arange(B), s + rng.integers(0, g, size=B)] = 1.0
        lab = (rng.random(B) < 0.03).astype(np.int64)
        oh = np.zeros((B, CRITEO_CLASSES), dtype=np.int16)
        oh[np.arange(B), lab] = 1
        ind.append(x)
        dep.append(oh)
    return ind, dep


def rows_for_toast_bytes(target_toast_bytes, ind_bytes=112 * 112 * 3 * 4, dep_bytes=2000, dep_external=True):
    """How many uncompressed ImageNet rows fill `target_toast_bytes` of TOAST file."""
    def tuples(nbytes):
        full, rest = divmod(nbytes, TOAST_MAX_CHUNK_SIZE)
        return full * 8176 + (((36 + rest + 7) & ~7) if rest else 0)
    per_row = tuples(ind_bytes) + (tuples(dep_bytes) if dep_external else 0)
    return max(1, int(target_toast_bytes * 0.995) // per_row)


# --------------------------------------------------------------------------------------------
# Catalog (sys_cat.dill) as da.input_fn loads it (/root/reference/cerebro_gpdb/da.py:38-49,174-182)
# --------------------------------------------------------------------------------------------

def write_catalog(path, entries, shape_rows):
    """entries: {'train': [dict(gp_segment_id, relname, relfilenode, toast_relfilenode)], 'valid': [...]}"""
    import dill
    import pandas as pd
    cols = ["gp_segment_id", "oid", "relname", "relfilenode", "toast_oid", "toast_relname",
            "toast_relfilenode", "reltoastidxid"]                       # da.py:105-108
    cats = {}
    for mode in ("train", "valid"):
        rows = []
        for e in entries.get(mode, []):
            rows.append([e["gp_segment_id"], e.get("oid", e["relfilenode"]), e["relname"], e["relfilenode"],
                         e.get("toast_oid", e["toast_relfilenode"]), "pg_toast_%d" % e["relfilenode"],
                         e["toast_relfilenode"], e["toast_relfilenode"] + 1])
        cats[mode] = pd.DataFrame(rows, columns=cols)
    cats["shape"] = pd.DataFrame(shape_rows, columns=["__dist_key__", "buffer_id", "independent_var_shape",
                                                      "dependent_var_shape", "table_name"])   # da.py:117-125
    with open(path, "wb") as f:
        dill.dump(cats, f)
    return cats
