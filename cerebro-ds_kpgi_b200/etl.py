"""GPU version of the reference's filesystem text ETL.

Mirrors /root/reference/cerebro_gpdb/etl_imagenet.py (`parser`, `main`, lines 30-74) and
/root/reference/cerebro_gpdb/etl_criteo.py (`parser`, lines 28-37): gpfdist TEXT unloads
(`id|{v,v,...}|label` per line, /root/reference/cerebro_gpdb/unload_imagenet.sql:1-14) are parsed into
float32 feature arrays and integer labels — on the GPU, every number by one thread, with the
reference's double rounding (Python float, then numpy float32).  The reference writes .h5 through h5py
(not available here) or .npy; `main` writes .npy in both cases.
"""
import ctypes

import numpy as np

from . import _lib
from . import segment as _segment
from .pg_page_reader import logs


def parse_text(text, values_per_row, label_digits=0, feature_shape=None, device=None, keep_on_device=False):
    """text: bytes / uint8 array of whole lines.  Returns (ids int64[rows], features float32[rows, *feature_shape or
    (values_per_row,)], labels int64[rows] or [rows, label_digits]) as numpy arrays (or CUDA tensors)."""
    import torch
    L = _lib.require_device()
    if device is None:
        device = torch.cuda.current_device()
    arr = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray, memoryview)) else np.ascontiguousarray(text, dtype=np.uint8)
    n = int(arr.size)
    cap_rows = int(np.count_nonzero(arr == 10)) + 1
    with torch.cuda.device(device):
        d_text = torch.from_numpy(arr.copy() if n else np.zeros(1, np.uint8)).cuda()
        feats = torch.empty((cap_rows, int(values_per_row)), dtype=torch.float32, device="cuda")
        labels = torch.empty((cap_rows, max(1, int(label_digits))), dtype=torch.int64, device="cuda")
        ids = torch.empty(cap_rows, dtype=torch.int64, device="cuda")
        n_rows, err_pos = ctypes.c_uint64(), ctypes.c_uint64()
        rc = L.cds_text_etl(int(device), _segment._ptr(d_text), n, int(values_per_row), int(label_digits), _segment._ptr(feats),
                            _segment._ptr(labels), _segment._ptr(ids), cap_rows, ctypes.byref(n_rows), ctypes.byref(err_pos),
                            _segment._stream_ptr())
    _lib.check(rc)
    if rc == _lib.CDS_ERR_TEXT_NUMBER:
        at = int(err_pos.value)
        raise ValueError("malformed node or string: %r at byte %d" % (bytes(arr[at:at + 24]), at))     # what literal_eval reports
    if rc == _lib.CDS_ERR_SHAPE:
        raise ValueError("a line does not hold 1 id + %d values + 1 label (near byte %d)" % (values_per_row, int(err_pos.value)))
    if rc:
        raise _lib.CdsError(rc)
    r = int(n_rows.value)
    feats, labels, ids = feats[:r], labels[:r], ids[:r]
    if feature_shape is not None:
        feats = feats.view((r,) + tuple(feature_shape))
    if not label_digits:
        labels = labels[:, 0]
    if keep_on_device:
        return ids, feats, labels
    return ids.cpu().numpy(), feats.cpu().numpy(), labels.cpu().numpy()


def parser(line, values_per_row=None, label_digits=0):
    """Single-line form of etl_imagenet.parser (etl_imagenet.py:30-37): (record_id, feature, label)."""
    if isinstance(line, str):
        line = line.encode()
    if values_per_row is None:
        values_per_row = line.split(b'|')[1].count(b',') + 1
    ids, feats, labels = parse_text(line, values_per_row, label_digits)
    return str(int(ids[0])), feats[0], (int(labels[0]) if not label_digits else [int(x) for x in labels[0]])


def main(file_list, output_list, values_per_row, label_digits=0, outformat='npy'):
    """etl_imagenet.main (etl_imagenet.py:39-74) with the parsing on the GPU; writes
    np.hstack([features, labels]) as .npy (etl_imagenet.py:70-72)."""
    logs("Start")
    for purpose in ['valid', 'train']:
        if purpose not in file_list:
            continue
        logs("Starting {}".format(purpose))
        logs("Start loading into memory {}".format(purpose))
        with open(file_list[purpose], 'rb') as f:
            text = f.read()
        logs("End loading into memory {}".format(purpose))
        logs("Start parsing {}".format(purpose))
        _, feats, labels = parse_text(text, values_per_row, label_digits)
        logs("End parsing {}".format(purpose))
        lab = labels[:, None] if labels.ndim == 1 else labels
        np.save(output_list[purpose], np.hstack([feats.reshape(feats.shape[0], -1), lab]))     # float64, as numpy promotes
        logs("Ending {}".format(purpose))
    logs("Ending")
