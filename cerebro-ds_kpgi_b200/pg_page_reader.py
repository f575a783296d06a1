"""Drop-in mirror of the reference reader interface, computed on the GPU.

Same module-level names, signatures, return structures and exceptions as
/root/reference/cerebro_gpdb/pg_page_reader.py ("ppr" below), so that `da.py` and the drivers
(`run_da_cerebro_standalone.py`, `run_pytorchddp_da.py`) can import this module instead:

    table_page_read(page_dir)                                   ppr:451-494
    toast_page_read(page_dir, df_toast, df_shape, table_name)   ppr:615-624
    generator_page(filename, toast=False, debug=False)          ppr:364-448
    read_toast_all_bytes(toast_page_path, df_toast)             ppr:497-524
    pglz_decompress(source, dest, implementation)               ppr:234-236
    pre_alloc_dest(bytea, implementation)                       ppr:239-250

All page walking, chunk matching, reassembly and pglz decoding runs in libcds_da.so
(include/cds_da.h); this file only moves bytes between files, pinned memory and the GPU, and
shapes the results into the pandas / numpy objects the callers expect.  There is no CPU decode
path here: without the CUDA library or a GPU every entry point raises.
"""
import ctypes
import datetime
import glob
import struct
import sys
from collections import namedtuple

import numpy as np

from . import _lib
from . import segment as _segment

# ---- constants (ppr:25-46, 627-628) -----------------------------------------------------------
IS_PYTHON3 = sys.version_info.major >= 3
HEAP_HASNULL = 0x0001
HEAP_NATTS_MASK = 0x07FF
VARHDRSZ_EXTERNAL = 4
SIZE_OF_VARATT_EXTERNAL = 16
BLOCK_SIZE = 32768
BLCKSZ = BLOCK_SIZE
PAGE_HEADER_LEN = 24
ITEM_IDENTIFIER_LEN = 4
VARHDRSZ = 4
SIZE_OF_ITEMIDDATA = ITEM_IDENTIFIER_LEN
ITEM_HEADER_LEN = 23
MAXIMUM_ALIGNOF = 8
SIZE_OF_PGLZ_HEADER = 8
CHUNK_ID_AND_CHUNK_SEQ_LEN = 8
TOAST_MAX_CHUNK_SIZE = 8140
PGLZ_IMPLEMENTATION = 'cuda'          # the reference says 'py' (ppr:46); here the decoder is the sm_100a kernel
LP_LENGTH_GLOBAL = TOAST_MAX_CHUNK_SIZE + VARHDRSZ + 24 + CHUNK_ID_AND_CHUNK_SEQ_LEN

PageHeader = namedtuple('PageHeader', ['pd_lsn', 'pd_tli', 'pd_flags', 'pd_lower', 'pd_upper', 'pd_special',
                                       'pd_pagesize_version', 'pd_prune_xid'])
HeapTupleHeader = namedtuple('HeapTupleHeader', ['t_xmin', 't_xmax', 't_cid', 'bi_hi', 'bi_lo', 'ip_posid',
                                                 't_infomask2', 't_infomask', 't_hoff'])
ItemIdentifier = namedtuple('ItemIdentifier', ['lp_off', 'lp_flags', 'lp_len'])
Tuple = namedtuple('Tuple', ['header', 'length', 'dummy1', 'dummy2', 'va_rawsize', 'va_extsize', 'va_valueid',
                             'va_toastrelid', 'external', 'bytea'])

# The reference prints three markers per external value while it detoasts (ppr:554, 600-603; parsed by
# plots/data_analytics.py): "ID:<valueid>", "START DECOMPRESSING", "END DECOMPRESSING".  Here all values of a
# segment are decoded by one batch of kernels, so there is no per-value moment to stamp; with LOG_VALUE_IDS = True
# the same lines are printed after the decode, in the reference's order (first-seen physical TOAST order), so that
# log parsers keep working (every value then shows the same second).  Off by default: it is one line per value.
LOG_VALUE_IDS = False


def logs(message):
    """Same format as cerebro_gpdb.utils.logs (utils.py:93-98): '<message>: YYYY-mm-dd HH:MM:SS'."""
    printed = "{}: {}".format(message, datetime.datetime.now().strftime("%Y-%m-%d %H:%M:%S"))
    print(printed)
    sys.stdout.flush()
    return printed


# ---- small varlena helpers kept for API compatibility (ppr:52-156) ------------------------------

def wrap_function(lib, funcname, restype, argtypes):
    func = getattr(lib, funcname)
    func.restype = restype
    func.argtypes = argtypes
    return func


def TYPEALIGN(ALIGNVAL, LEN):
    return (LEN + ALIGNVAL - 1) & ~(ALIGNVAL - 1)


def MAX_ALIGN(LEN):
    return TYPEALIGN(MAXIMUM_ALIGNOF, LEN)


def VARSIZE_1B_E():
    return SIZE_OF_VARATT_EXTERNAL + VARHDRSZ_EXTERNAL


def get_1b_header(bytea):
    return bytea if isinstance(bytea, int) else bytea[0]


def VARATT_IS_1B(bytea):
    return (get_1b_header(bytea) & 0x80) == 0x80


def VARATT_IS_4B_U(bytea):
    return (get_1b_header(bytea) & 0xC0) == 0x00


def VARATT_IS_EXTENDED(bytea):
    return not VARATT_IS_4B_U(bytea)


def VARATT_IS_1B_E(bytea):
    return get_1b_header(bytea) == 0x80


def VARATT_IS_4B_C(bytea):
    return (get_1b_header(bytea) & 0xC0) == 0x40


VARATT_IS_COMPRESSED = VARATT_IS_4B_C
VARATT_IS_EXTERNAL = VARATT_IS_1B_E


def VARSIZE(bytea):
    """Big-endian 4-byte length word, low 30 bits (ppr:130-131)."""
    return struct.unpack('>I', bytes(bytea[:4]))[0] & 0x3FFFFFFF


def VARDATA(bytea):
    return bytea[4:]


def GET_VARSIZE_4B_C(total_length):
    return struct.pack('>I', (total_length & 0x3FFFFFFF) | 0x40000000)


def GET_VARSIZE_4B(total_length):
    return struct.pack('>I', total_length & 0x3FFFFFFF)


def GET_RAWSIZE_FROM_COMPRESSED(bytea):
    return struct.unpack('<I', bytes(bytea[4:8]))[0]


def SET_VARSIZE(bytea, header):            # ppr:60-61
    bytea[:4] = header


def att_align_nominal(off):                # ppr:64-69 (int alignment)
    return TYPEALIGN(4, off)


def VARSIZE_1B(bytea):                     # ppr:84-86
    return get_1b_header(bytea) & 0x7F


def acc(bytea, i):                         # ppr:188-189
    return bytea[i:i + 1]


def expand_input_dims(input_data):         # ppr:159-162
    return np.expand_dims(np.array(input_data, dtype=np.float32), axis=0)


def read_page_at(f, pos):                  # ppr:358-361
    f.seek(pos)
    return f.read(BLOCK_SIZE)


# The per-page parse helpers of the reference's CPU loop (deserialize_page_header, deserialize_item,
# deserialize_item_header, extract_tuple, desirialize_tupdata; ppr:253-351) have no counterpart here: that
# parse runs inside k_page_index / k_heap_rows / k_toast_walk, and table_page_read / generator_page return
# its results in the reference's own record layout.


def get_num_chunks(ressize):
    return ((ressize - 1) // TOAST_MAX_CHUNK_SIZE) + 1


def np_array_float32(var, var_shape):
    arr = np.frombuffer(var, dtype=np.float32)
    arr.shape = var_shape
    return arr


def np_array_int16(var, var_shape):
    arr = np.frombuffer(var, dtype=np.int16)
    arr.shape = var_shape
    return arr


def deserialize_bytea(attname, dest_data, var_shape):
    if attname == 'dependent_var':
        return np_array_int16(dest_data, var_shape)
    elif attname == 'independent_var':
        return np_array_float32(dest_data, var_shape)


# ---- error mapping -----------------------------------------------------------------------------

def raise_for_status(code):
    """Re-raise a data error of the CUDA library as the exception the reference raises there."""
    code = int(code)
    if code == _lib.CDS_OK:
        return
    if code == _lib.CDS_ERR_PARTIAL_PAGE:
        raise Exception("FATAL")                                                       # ppr:377
    if code == _lib.CDS_ERR_ITEMID_ALIGN:
        raise Exception("""Item identifier length do not match.
                        Must be a multiply of {}.""".format(ITEM_IDENTIFIER_LEN))     # ppr:263-265
    if code == _lib.CDS_ERR_TOAST_SPECIAL:
        raise AssertionError("THERE SHALL NOT BE INDICES")                             # ppr:388-389
    if code == _lib.CDS_ERR_TUPLE:
        raise struct.error("unpack requires a buffer of 23 bytes")                     # ppr:273-274
    if code == _lib.CDS_ERR_RAWSIZE:
        raise AssertionError('raw size does not match!')                               # ppr:539
    if code in (_lib.CDS_ERR_CHUNK_COUNT, _lib.CDS_ERR_CHUNK_EXTENDED):
        raise AssertionError()                                                         # ppr:572, 580
    if code == _lib.CDS_ERR_MAX_CHUNK:
        raise Exception("max chunk size error")                                        # ppr:586
    if code == _lib.CDS_ERR_CHUNK_SIZE:
        raise Exception("unexpected chunk size")                                       # ppr:588, 591
    if code == _lib.CDS_ERR_FINAL_SIZE:
        raise AssertionError("final size does not match")                              # ppr:595-596
    if code == _lib.CDS_ERR_PGLZ_CORRUPT:
        raise Exception("compressed data is corrupt")                                  # ppr:230
    if code == _lib.CDS_ERR_SHAPE:
        raise ValueError("cannot reshape array: decoded size does not match the catalog shape")   # ppr:167
    if code == _lib.CDS_ERR_CATALOG_MISS:
        raise IndexError("single positional indexer is out-of-bounds")                 # ppr:563-564
    if code == _lib.CDS_ERR_MISSING_VALUE:
        raise ValueError("a TOAST pointer has no chunks in the TOAST relation")
    raise _lib.CdsError(code)


# ---- file layer (ppr:358-379) ------------------------------------------------------------------

def relation_files(filename):
    """Base file, then its `.N` continuation files ordered by (length, name) (ppr:365-368)."""
    seg_files = glob.glob(filename + '.*')
    seg_files.sort()
    seg_files.sort(key=len)
    return [filename] + seg_files


def read_relation(filename, pinned=True):
    """All pages of a relation in one host buffer (page-locked when `pinned`, so the H2D copy is
    asynchronous and full-speed).  Returns (holder, uint8 array, partial) where `partial` says the
    reference would raise Exception("FATAL") at the trailing short page (ppr:375-377)."""
    import os
    files = relation_files(filename)
    sizes = [os.path.getsize(f) for f in files]
    whole = [s - s % BLOCK_SIZE for s in sizes]
    partial = False
    total = 0
    for s, w in zip(sizes, whole):            # pages after the first short page are never reached
        total += w
        if s != w:
            partial = True
            break
    holder = _segment.PinnedBuffer(total) if pinned else None
    arr = holder.array if pinned else np.empty(total, dtype=np.uint8)
    off = 0
    for f, s, w in zip(files, sizes, whole):
        if w:
            with open(f, 'rb') as fh:
                got = fh.readinto(memoryview(arr[off:off + w]))
            assert got == w
        off += w
        if s != w:
            break
    return holder, arr, partial


def _to_device(arr):
    import torch
    t = torch.empty(arr.size, dtype=torch.uint8, device='cuda')
    if arr.size:
        t.copy_(torch.from_numpy(arr), non_blocking=True)
    return t


def _rows_to_frames(rows, heap_host):
    """df_data / df_toast exactly as table_page_read builds them (ppr:465-492)."""
    import pandas as pd
    n = rows.shape[0]
    all_data, all_toast = [], []
    hdr = rows[:, [2, 7]].copy().view(np.uint8).reshape(n, 2, 4)
    signed = rows.view(np.int32)
    inline_off = rows[:, 13].astype(np.uint64) | (rows[:, 14].astype(np.uint64) << np.uint64(32))
    for r in range(n):
        tuples = []
        for a, base in ((0, 2), (1, 7)):
            external = True if a == 0 else bool(rows[r, 12])
            bytea = None
            if not external:
                o = int(inline_off[r])
                bytea = bytes(heap_host[o:o + int(rows[r, 15])])
            tuples.append(Tuple(int(hdr[r, a, 0]), int(hdr[r, a, 1]), int(hdr[r, a, 2]), int(hdr[r, a, 3]),
                                int(signed[r, base + 1]), int(signed[r, base + 2]), int(rows[r, base + 3]),
                                int(rows[r, base + 4]), external, bytea))
        dist_key, buffer_id = int(rows[r, 0]), int(rows[r, 1])
        all_data.append((dist_key, tuples[0], tuples[1], buffer_id))
        all_toast.append(list(tuples[0][4:]) + [dist_key, buffer_id, 'independent_var'])
        all_toast.append(list(tuples[1][4:]) + [dist_key, buffer_id, 'dependent_var'])
    df_data = pd.DataFrame(all_data, columns=['dist_key', 'independent_var_tuple', 'dependent_var_tuple', 'buffer_id'])
    df_toast = pd.DataFrame(all_toast, columns=['va_rawsize', 'va_extsize', 'va_valueid', 'va_toastrelid', 'external',
                                                'bytea', 'dist_key', 'buffer_id', 'attname'])
    return df_data, df_toast


def _scan_heap(page_dir):
    import torch
    _lib.require_device()
    holder, heap_host, partial = read_relation(page_dir, pinned=False)     # small; its bytes are needed on the host too
    n_pages = heap_host.size // BLOCK_SIZE
    seg = _segment.DeviceSegment(max(1, n_pages), 0)
    try:
        d_heap = _to_device(heap_host)
        seg.heap_scan(d_heap)
        s = seg.finish()
        raise_for_status(s.status)
        if partial:
            raise_for_status(_lib.CDS_ERR_PARTIAL_PAGE)
        rows = seg.rows()
        torch.cuda.synchronize()
    finally:
        seg.close()
    return rows, heap_host, holder


def table_page_read(page_dir):
    """ppr:451-494.  Returns (df_data, df_toast)."""
    logs("START READING TABLE: {}".format(page_dir))
    rows, heap_host, holder = _scan_heap(page_dir)
    df_data, df_toast = _rows_to_frames(rows, heap_host)
    logs("END READING TABLE: {}".format(page_dir))
    return df_data, df_toast


def _file_of_page(filename):
    """[(first page, pages)] per file of the relation, for the per-file positions debug=True prints (ppr:380-381)."""
    import os
    out, first = [], 0
    for f in relation_files(filename):
        n = os.path.getsize(f) // BLOCK_SIZE
        out.append((first, n))
        first += n
    return out


def _debug_page(host, page, spans):
    """What the reference holds per page when debug=True: prints "PAGE: <position in its file>" (ppr:380-381) and
    returns (page_bytes, PageHeader, item_identifiers_bytes, item_num) as deserialize_page_header does (ppr:253-269)."""
    pos = next((page - first) * BLOCK_SIZE for first, n in spans if first <= page < first + n)
    print("PAGE: {}".format(pos))
    page_bytes = bytes(host[page * BLOCK_SIZE:(page + 1) * BLOCK_SIZE])
    hdr = list(struct.unpack('@qHHHHHHI', page_bytes[:PAGE_HEADER_LEN]))
    hdr[-2] &= 0x00FF
    header = PageHeader(*hdr)
    ids = page_bytes[PAGE_HEADER_LEN:header.pd_lower]
    return page_bytes, header, ids, len(ids) // ITEM_IDENTIFIER_LEN


def generator_page(filename, toast=False, debug=False):
    """ppr:364-448.  Yields (dist_key, independent_var_tuple, dependent_var_tuple, buffer_id, None) for a
    heap relation or (chunk_id, chunk_seq, chunk_data, None) for a TOAST relation, in the reference's
    order; the walk itself runs on the GPU.  With debug=True the last element is the reference's tuple
    (page_bytes, page_header, item_identifiers_bytes, lp_off | None, item_num, item_identifier, item_header, tupdata)
    (ppr:415-420, 433-445) and "PAGE: <pos>" is printed as each page is entered: positions come from the GPU's row
    table / chunk directory, the headers shown are unpacked from the page bytes for display."""
    if not toast:
        rows, heap_host, holder = _scan_heap(filename)
        df_data, _ = _rows_to_frames(rows, heap_host)
        if not debug:
            for rec in df_data.itertuples(index=False):
                yield rec[0], rec[1], rec[2], rec[3], None
            return
        spans = _file_of_page(filename)
        inline_off = rows[:, 13].astype(np.uint64) | (rows[:, 14].astype(np.uint64) << np.uint64(32))
        r = 0
        recs = list(df_data.itertuples(index=False))
        for page in range(heap_host.size // BLOCK_SIZE):
            page_bytes, header, ids, item_num = _debug_page(heap_host, page, spans)
            for i in range(item_num):
                word = struct.unpack('<I', ids[4 * i:4 * i + 4])[0]
                ident = ItemIdentifier(word & 0x7FFF, (word >> 15) & 3, word >> 17)       # ppr:285-300
                item = page_bytes[ident.lp_off:ident.lp_off + ident.lp_len]
                item_header = HeapTupleHeader(*struct.unpack('@IIIHHHHHB', item[:ITEM_HEADER_LEN]))
                tupdata = item[item_header.t_hoff:]
                rec = recs[r]
                assert int(inline_off[r]) == page * BLOCK_SIZE + ident.lp_off + item_header.t_hoff + 24
                r += 1
                yield rec[0], rec[1], rec[2], rec[3], (page_bytes, header, ids, None, item_num, ident, item_header, tupdata)
        return
    _lib.require_device()
    holder, toast_host, partial = read_relation(filename)
    n_pages = toast_host.size // BLOCK_SIZE
    seg = _segment.DeviceSegment(1, n_pages)
    try:
        seg.set_rows(np.zeros((0, _lib.ROW_WORDS), dtype=np.uint32), None)
        d_toast = _to_device(toast_host)
        seg.toast_scan(d_toast)
        s = seg.finish()
        cid, seq, off, ln = seg.chunks(max(1, int(s.n_chunks)))
        # the reference yields every chunk before the failing page, then raises
        err = int(s.status)
        stop_at = int(s.err_where) if err and s.err_stage == 1 else None
    finally:
        seg.close()
    if not debug:
        for i in range(cid.size):
            o, n = int(off[i]), int(ln[i])
            if stop_at is not None and o // BLOCK_SIZE >= stop_at:
                break
            yield int(cid[i]), int(seq[i]), bytes(toast_host[o - 4:o + n]), None
    else:
        spans = _file_of_page(filename)
        i = 0
        last_page = n_pages if stop_at is None else min(n_pages, stop_at + 1)
        for page in range(last_page):
            page_bytes, header, ids, item_num = _debug_page(toast_host, page, spans)
            if stop_at is not None and page >= stop_at:
                break                                                 # the reference fails inside this page
            lp_off = header.pd_upper
            while i < cid.size and int(off[i]) // BLOCK_SIZE == page:
                o, n = int(off[i]) - page * BLOCK_SIZE, int(ln[i])
                lp_off = MAX_ALIGN(lp_off)                                # ppr:394
                t_hoff = o - 12 - lp_off                                  # payload = item + t_hoff + 8 (ids) + 4 (length word)
                lp_len = t_hoff + 8 + 4 + n
                ident = ItemIdentifier(lp_off, 1, lp_len)                 # ppr:391,407-408
                item = page_bytes[lp_off:lp_off + lp_len]
                item_header = HeapTupleHeader(*struct.unpack('@IIIHHHHHB', item[:ITEM_HEADER_LEN]))
                assert item_header.t_hoff == t_hoff
                lp_off += lp_len
                yield int(cid[i]), int(seq[i]), page_bytes[o - 4:o + n], (page_bytes, header, ids, lp_off, item_num, ident,
                                                                         item_header, item[t_hoff:])
                i += 1
    raise_for_status(err)
    if partial:
        raise_for_status(_lib.CDS_ERR_PARTIAL_PAGE)


def read_toast_all_bytes(toast_page_path, df_toast, debug=False):
    """ppr:497-524: {chunk_id: [[chunk_seq, chunk_data], ...]} for chunk ids present in df_toast."""
    wanted = set(int(x) for x in df_toast['va_valueid'].values)
    grand_bytes = {}
    for chunk_id, chunk_seq, chunk_data, _ in generator_page(toast_page_path, toast=True, debug=debug):
        if chunk_id in wanted:
            grand_bytes.setdefault(chunk_id, []).append([chunk_seq, chunk_data])
    return grand_bytes


def _rows_from_df_toast(df_toast):
    """Inverse of _rows_to_frames: exported-row array + blob of the inline datums."""
    att = df_toast['attname'].values
    n = len(df_toast)
    if n % 2 or not (np.all(att[0::2] == 'independent_var') and np.all(att[1::2] == 'dependent_var')):
        raise ValueError("df_toast must hold one independent_var row followed by one dependent_var row per tuple "
                         "(the layout table_page_read produces, ppr:466-474)")
    rows = np.zeros((n // 2, _lib.ROW_WORDS), dtype=np.uint32)
    col = {c: df_toast[c].values for c in ('va_rawsize', 'va_extsize', 'va_valueid', 'va_toastrelid', 'external',
                                           'dist_key', 'buffer_id')}
    bytea = df_toast['bytea'].values
    rows[:, 0] = col['dist_key'][0::2].astype(np.int64) & 0xFFFFFFFF
    rows[:, 1] = col['buffer_id'][0::2].astype(np.int64) & 0xFFFFFFFF
    for a, base in ((0, 2), (1, 7)):
        rows[:, base] = 0x80 if a == 0 else 0
        rows[:, base + 1] = col['va_rawsize'][a::2].astype(np.int64) & 0xFFFFFFFF
        rows[:, base + 2] = col['va_extsize'][a::2].astype(np.int64) & 0xFFFFFFFF
        rows[:, base + 3] = col['va_valueid'][a::2].astype(np.int64) & 0xFFFFFFFF
        rows[:, base + 4] = col['va_toastrelid'][a::2].astype(np.int64) & 0xFFFFFFFF
    rows[:, 12] = col['external'][1::2].astype(bool).astype(np.uint32)
    parts, off = [], 0
    for r in range(n // 2):
        if not rows[r, 12]:
            b = bytea[2 * r + 1]
            rows[r, 13], rows[r, 14], rows[r, 15] = off & 0xFFFFFFFF, off >> 32, len(b)
            pad = (-len(b)) % 16
            parts.append(bytes(b) + b"\0" * pad)
            off += len(b) + pad
    blob = np.frombuffer(b"".join(parts) if parts else b"\0" * 16, dtype=np.uint8)
    return rows, blob


def _shapes_for(df_shape, table_name, buffer_ids):
    """{buffer_id: row} lookups of ppr:544-547 / 565-569, done once instead of once per value."""
    sub = df_shape.loc[df_shape['table_name'] == table_name]
    first = {}
    for bid, ishape, dshape in zip(sub['buffer_id'].values, sub['independent_var_shape'].values,
                                   sub['dependent_var_shape'].values):
        first.setdefault(int(bid), (ishape, dshape))
    for bid in buffer_ids:
        if bid not in first:
            raise_for_status(_lib.CDS_ERR_CATALOG_MISS)
    return first


def _assemble(part, ind_host, dep_host, df_shape, table_name):
    out = {}
    shapes = _shapes_for(df_shape, table_name, part.buffer_ids) if df_shape is not None else None
    for bid, ind, dep in part.value_views(ind_host, dep_host):
        if shapes is not None:
            ishape, dshape = shapes[bid]
            dep.shape = tuple(dshape)       # same statement the reference uses: ValueError on a size mismatch
            ind.shape = tuple(ishape)
        entry = out.setdefault(bid, {"dependent_var": None, "independent_var": None})
        entry["dependent_var"] = dep
        entry["independent_var"] = ind
    return out


def _log_value_markers(part, rows):
    wanted = set(int(x) for x in rows[:, 5]) | set(int(rows[r, 10]) for r in range(rows.shape[0]) if rows[r, 12])
    seen = set()
    for cid in part.chunk_ids:
        cid = int(cid)
        if cid in wanted and cid not in seen:
            seen.add(cid)
            logs("ID:{}".format(cid))
            logs("START DECOMPRESSING")
            logs("END DECOMPRESSING")


def _detoast_on_device(d_heap, d_toast, rows, inline_blob, df_shape, table_name, partial):
    import torch
    status, part = _segment.decode_pages(d_heap, d_toast, rows=rows, inline_blob=inline_blob, want_chunk_ids=LOG_VALUE_IDS)
    if LOG_VALUE_IDS and status == 0:
        _log_value_markers(part, part.rows)
    # errors of the TOAST walk surface in read_toast_all_bytes, before detoast; a trailing short page
    # is met by the walk only after every whole page was read (ppr:375-377)
    if status and part.summary.err_stage <= 1:
        raise_for_status(status)
    if partial:
        raise_for_status(_lib.CDS_ERR_PARTIAL_PAGE)
    raise_for_status(status)
    from . import hostio
    return _assemble(part, hostio.device_to_numpy(part.ind), hostio.device_to_numpy(part.dep), df_shape, table_name)


def toast_page_read(page_dir, df_toast, df_shape, table_name):
    """ppr:615-624.  Returns {buffer_id: {'dependent_var': int16 array, 'independent_var': float32 array}}
    in the reference's insertion order."""
    import torch
    _lib.require_device()
    logs("START READING TOAST: {}".format(page_dir))
    from . import hostio
    rows, blob = _rows_from_df_toast(df_toast)
    d_toast, partial = hostio.relation_to_device(page_dir, relation_files)
    d_blob = torch.from_numpy(blob.copy()).cuda()
    logs("END READING TOAST: {}".format(page_dir))
    logs("START DETOASTING")
    out = _detoast_on_device(None, d_toast, rows, d_blob, df_shape, table_name, partial)
    logs("END DETOASTING")
    logs("ITEM NUM:{}, BUFFER_IDs:{}".format(len(out.keys()), out.keys()))
    return out


def detoast(grand_bytes, df_toast, df_shape, table_name):
    """ppr:527-612 for callers that hold chunk bytes on the host: the joins happen here, every pglz
    stream is decoded by the GPU in one batch."""
    external = df_toast.loc[df_toast['external'] == True]      # noqa: E712
    internal = df_toast.loc[df_toast['external'] == False]     # noqa: E712
    datums, owners = [], []
    for _, row in internal.iterrows():
        bytea = row['bytea']
        rawsize = struct.unpack('<i', bytes(bytea[4:8]))[0]
        assert rawsize == row['va_rawsize'], 'raw size does not match!'
        datums.append(bytes(bytea[:VARSIZE(bytea)]) if VARSIZE(bytea) <= len(bytea) else bytes(bytea))
        owners.append((row['buffer_id'], row['attname']))
    for k, v in grand_bytes.items():
        if LOG_VALUE_IDS:
            logs("ID:{}".format(k))
            logs("START DECOMPRESSING")
            logs("END DECOMPRESSING")
        hit = external.loc[external['va_valueid'] == k].iloc[0]
        va_extsize, buffer_id, attname = hit['va_extsize'], hit['buffer_id'], hit['attname']
        numchunks = get_num_chunks(va_extsize)
        assert numchunks == len(v)
        v = sorted(v, key=lambda x: x[0])
        parts = [GET_VARSIZE_4B_C(VARHDRSZ + va_extsize)]
        for idx, chunk in v:
            assert not VARATT_IS_EXTENDED(chunk)
            chunksize = VARSIZE(chunk) - VARHDRSZ
            parts.append(bytes(chunk[4:4 + chunksize]))
            if idx == 0 and chunksize < va_extsize and chunksize != TOAST_MAX_CHUNK_SIZE:
                raise Exception("max chunk size error")
            if idx < numchunks - 1 and chunksize != TOAST_MAX_CHUNK_SIZE:
                raise Exception("unexpected chunk size")
            if idx == numchunks - 1 and idx * TOAST_MAX_CHUNK_SIZE + chunksize != va_extsize:
                raise Exception("unexpected chunk size")
        bytea = b''.join(parts)
        assert len(bytea) == VARHDRSZ + va_extsize, "final size does not match"
        datums.append(bytea)
        owners.append((buffer_id, attname))
    decoded, codes = _segment.pglz_decompress_batch(datums)
    out = {}
    for (buffer_id, attname), raw, code in zip(owners, decoded, codes):
        raise_for_status(code)
        shape_info = df_shape.loc[(df_shape['table_name'] == table_name) & (df_shape['buffer_id'] == buffer_id)].iloc[0]
        arr = deserialize_bytea(attname, bytearray(raw), shape_info[attname + '_shape'])
        out.setdefault(buffer_id, {"dependent_var": None, "independent_var": None})[attname] = arr
    return out


# ---- pglz entry points (ppr:234-250) -------------------------------------------------------------

def pre_alloc_dest(bytea, implementation=PGLZ_IMPLEMENTATION):
    """ppr:239-250: destination buffer sized from the rawsize field."""
    dest_data_size = struct.unpack('<I', bytes(bytea[4:8]))[0]
    if implementation == 'c':
        dest_data = ctypes.create_string_buffer(dest_data_size)
    else:
        dest_data = bytearray(dest_data_size)
    return dest_data, dest_data_size


def pglz_decompress(source, dest, implementation=PGLZ_IMPLEMENTATION):
    """ppr:234-236 / ppr:191-231.  Decodes `source` (8-byte header + stream) into `dest` on the GPU,
    whatever `implementation` says, and returns `dest`; raises Exception("compressed data is corrupt")."""
    decoded, codes = _segment.pglz_decompress_batch([bytes(source)])
    raise_for_status(codes[0])
    n = len(decoded[0])
    if isinstance(dest, (bytearray, memoryview)):
        dest[:n] = decoded[0]
    else:
        ctypes.memmove(dest, decoded[0], n)
    return dest


pglz_decompress_py = pglz_decompress      # name kept for callers of the reference's Python decoder
