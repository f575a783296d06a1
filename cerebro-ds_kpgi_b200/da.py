"""Drop-in mirror of /root/reference/cerebro_gpdb/da.py: the Direct-Access entry point.

`input_fn(file_path, overwrite_table_page_path=None, overwrite_toast_page_path=None)` keeps the
reference's signature, path convention and result (da.py:29-58): the last path component is the
mode marker ('train' / 'valid'), the segment id comes from 'gpseg<N>' in the path, the relation
files are looked up in the catalog at SYS_CAT_PATH (a module global, as in the reference, so
callers may re-point it), and the result is {buffer_id: {'dependent_var': int16 array,
'independent_var': float32 array}} in the reference's insertion order.

`input_fn_device` is the B200-native variant of the same call: the decoded partition stays in HBM
as torch tensors (NHWC float32 + int16 one-hot, plus int64 labels), ready for training.

The catalog *client* of the reference (DirectAccessClient, da.py:61-183) runs SQL against a live
Greenplum and is outside this path; `build_catalog` writes the same sys_cat.dill from frames.
"""
import os
import re

import numpy as np

from . import _lib
from . import hostio
from . import pg_page_reader
from . import segment as _segment
from .pg_page_reader import logs, table_page_read, toast_page_read  # noqa: F401  (same names as da.py:20-22)

SYS_CAT_PATH = '/mnt/nfs/sys_cat.dill'          # da.py:26


def _load_catalog(file_path):
    import dill
    head, mode = os.path.split(file_path)                                   # da.py:34-37
    logs("CURRENT MODE: {}".format(mode))
    with open(SYS_CAT_PATH, "rb") as f:
        sys_cats = dill.load(f)                                             # da.py:38-39
    df_shape = sys_cats['shape']
    frame = sys_cats[mode]
    gp_segment_id = int(re.search(r"gpseg(\d+)", file_path).group(1))       # da.py:43
    sys_cat = frame.loc[frame['gp_segment_id'] == gp_segment_id].iloc[0]    # da.py:44 (IndexError when absent)
    table_page_path = os.path.join(head, str(int(sys_cat['relfilenode'])))
    toast_page_path = os.path.join(head, str(int(sys_cat['toast_relfilenode'])))
    return df_shape, sys_cat['relname'], table_page_path, toast_page_path


def _decode_files(table_page_path, toast_page_path, ind_sample_bytes=0, dep_sample_bytes=0, ind_bf16=False,
                  u8_normalize=None):
    """Both relations -> pinned host memory -> HBM -> one fused decode.  Returns (part, partial flags)."""
    import torch
    _lib.require_device()
    logs("START READING TABLE: {}".format(table_page_path))
    d_heap, heap_partial = hostio.relation_to_device(table_page_path, pg_page_reader.relation_files)
    logs("END READING TABLE: {}".format(table_page_path))
    logs("START READING TOAST: {}".format(toast_page_path))
    d_toast, toast_partial = hostio.relation_to_device(toast_page_path, pg_page_reader.relation_files)
    logs("END READING TOAST: {}".format(toast_page_path))
    logs("START DETOASTING")
    status, part = _segment.decode_pages(d_heap, d_toast, ind_sample_bytes=ind_sample_bytes,
                                         dep_sample_bytes=dep_sample_bytes, ind_bf16=ind_bf16, u8_normalize=u8_normalize,
                                         want_chunk_ids=pg_page_reader.LOG_VALUE_IDS)
    torch.cuda.synchronize()
    if pg_page_reader.LOG_VALUE_IDS and status == 0:
        pg_page_reader._log_value_markers(part, part.rows)          # ppr:554, 600-603
    # reference order of failures: heap scan, short heap page, TOAST walk, short TOAST page, detoast
    stage = part.summary.err_stage if status else 99
    if status and stage == 0:
        pg_page_reader.raise_for_status(status)
    if heap_partial:
        pg_page_reader.raise_for_status(_lib.CDS_ERR_PARTIAL_PAGE)
    if status and stage == 1:
        pg_page_reader.raise_for_status(status)
    if toast_partial:
        pg_page_reader.raise_for_status(_lib.CDS_ERR_PARTIAL_PAGE)
    pg_page_reader.raise_for_status(status)
    return part


def input_fn(file_path, overwrite_table_page_path=None, overwrite_toast_page_path=None):
    """da.py:29-58.  Same arguments, same result structure, decoded on the GPU and copied back to
    host numpy arrays (what the reference's callers index into)."""
    import torch
    df_shape, table_name, table_page_path, toast_page_path = _load_catalog(file_path)
    if overwrite_table_page_path:
        table_page_path = overwrite_table_page_path                         # da.py:50-53
    if overwrite_toast_page_path:
        toast_page_path = overwrite_toast_page_path
    part = _decode_files(table_page_path, toast_page_path)
    ind_host = hostio.device_to_numpy(part.ind)
    dep_host = hostio.device_to_numpy(part.dep)
    out = pg_page_reader._assemble(part, ind_host, dep_host, df_shape, table_name)
    logs("END DETOASTING")
    logs("ITEM NUM:{}, BUFFER_IDs:{}".format(len(out.keys()), out.keys()))
    return out


class DevicePartition:
    """A decoded partition resident on the GPU: what `input_fn_device` returns.

    images   float32 [N, *sample_shape]   (NHWC for ImageNet; [N, 7306] for Criteo)
    onehot   int16   [N, classes]
    labels   int64   [N]                  np.argmax(onehot, 1)  (run_pytorchddp_da.py:55)
    buffer_ids, rows_per_buffer           result order of the reference's dict
    """

    def __init__(self, images, onehot, labels, buffer_ids, rows_per_buffer):
        self.images, self.onehot, self.labels = images, onehot, labels
        self.buffer_ids, self.rows_per_buffer = buffer_ids, rows_per_buffer

    def __len__(self):
        return int(self.images.shape[0])

    def as_dict(self):
        """The reference's {buffer_id: {...}} view, as CUDA tensor slices (no copy)."""
        out, o = {}, 0
        for bid, n in zip(self.buffer_ids, self.rows_per_buffer):
            out[bid] = {"dependent_var": self.onehot[o:o + n], "independent_var": self.images[o:o + n]}
            o += n
        return out


def input_fn_device(file_path, overwrite_table_page_path=None, overwrite_toast_page_path=None, dtype=None,
                    u8_normalize=None):
    """Same call as input_fn, result left on the GPU.  Sample shapes come from the catalog's shape
    frame (every buffer of a table shares its trailing dimensions).  dtype=torch.bfloat16 makes the decode
    kernels emit bfloat16 images directly (round to nearest even); default: the stored float32 bits.
    u8_normalize = float32 [channels, 256] table: the table stores uint8 pixels and the kernels emit
    float32 table[channel][x] (e.g. synth.norm_lut(): the reference's (x/255 - mean)/std, exact)."""
    import torch
    df_shape, table_name, table_page_path, toast_page_path = _load_catalog(file_path)
    if overwrite_table_page_path:
        table_page_path = overwrite_table_page_path
    if overwrite_toast_page_path:
        toast_page_path = overwrite_toast_page_path
    sub = df_shape.loc[df_shape['table_name'] == table_name]
    if len(sub) == 0:
        pg_page_reader.raise_for_status(_lib.CDS_ERR_CATALOG_MISS)
    ishape = tuple(int(x) for x in sub['independent_var_shape'].iloc[0][1:])
    dshape = tuple(int(x) for x in sub['dependent_var_shape'].iloc[0][1:])
    isb, dsb = (1 if u8_normalize is not None else 4) * int(np.prod(ishape)), 2 * int(np.prod(dshape))
    bf16 = dtype is not None and dtype == torch.bfloat16 and u8_normalize is None
    part = _decode_files(table_page_path, toast_page_path, isb, dsb, ind_bf16=bf16, u8_normalize=u8_normalize)
    images = part.ind.view(torch.bfloat16 if bf16 else torch.float32).view((-1,) + ishape)
    onehot = part.dep.view(torch.int16).view((-1,) + dshape)
    labels = _segment.onehot_argmax(onehot.reshape(onehot.shape[0], -1))
    lay = part.layout
    rows_per = [int(lay["ind_len"][r]) // isb for r in lay["order"]]
    logs("END DETOASTING")
    return DevicePartition(images, onehot, labels, part.buffer_ids, rows_per)


def prefetch_device(file_paths, dtype=None, u8_normalize=None):
    """input_fn_device over several partitions (e.g. 'train' then 'valid', or the partitions a model hopper visits),
    one ahead: while the caller works on partition i, the files of partition i+1 are read, copied to HBM and decoded
    on a worker thread with its own CUDA stream.  Yields (file_path, DevicePartition) in order.  The optional
    async variant of the per-partition call; the reference reads partitions one after another."""
    import threading
    import torch
    paths = list(file_paths)
    if not paths:
        return
    dev = torch.cuda.current_device()
    results = {}

    def work(i):
        try:
            with torch.cuda.device(dev), torch.cuda.stream(torch.cuda.Stream(device=dev)):
                part = input_fn_device(paths[i], dtype=dtype, u8_normalize=u8_normalize)
                torch.cuda.current_stream().synchronize()
            results[i] = (part, None)
        except BaseException as e:  # noqa  (re-raised in the caller's thread)
            results[i] = (None, e)

    th = threading.Thread(target=work, args=(0,))
    th.start()
    for i in range(len(paths)):
        th.join()
        part, err = results.pop(i)
        if i + 1 < len(paths):
            th = threading.Thread(target=work, args=(i + 1,))
            th.start()
        if err is not None:
            if i + 1 < len(paths):
                th.join()
            raise err
        yield paths[i], part


def build_catalog(path, df_train, df_valid, df_shape):
    """Write sys_cat.dill as DirectAccessClient.generate_cats does (da.py:174-182), from frames the
    caller already has (columns of da.py:105-108 and da.py:117-125)."""
    import dill
    with open(path, "wb") as f:
        dill.dump({'train': df_train, 'valid': df_valid, 'shape': df_shape}, f)


def data_catalog(segment_ids, dboid, data_root='/mnt/gpdata'):
    """The Cerebro data catalog of DirectAccessClient.cat_factory / generate_cats (da.py:149-172):
    one partition per segment, identity availability, mode-marker paths."""
    n = len(segment_ids)
    eye = np.eye(n, dtype=int).tolist()
    return {
        'data_root': data_root,
        'train': ['gpseg{}/base/{}/train'.format(i, dboid) for i in segment_ids],
        'train_availability': eye,
        'valid': ['gpseg{}/base/{}/valid'.format(i, dboid) for i in segment_ids],
        'valid_availability': eye,
    }
