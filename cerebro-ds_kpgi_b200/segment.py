"""Torch-facing wrapper of the C-ABI segment handle (include/cds_da.h).

`DeviceSegment` is the B200-native form of one `da.input_fn` call
(/root/reference/cerebro_gpdb/da.py:29-58): page buffers in, decoded values written straight into
caller-owned torch CUDA tensors.  PyTorch is used for device memory and streams only; all compute
happens in libcds_da.so.
"""
import ctypes

import numpy as np

from . import _lib
from ._lib import BLCKSZ, ROW_WORDS, Summary


def _torch():
    import torch
    return torch


def _ptr(t):
    """Device/host address of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if isinstance(t, np.ndarray):
        return ctypes.c_void_p(t.ctypes.data)
    return ctypes.c_void_p(t.data_ptr())


def _stream_ptr(stream=None):
    torch = _torch()
    s = stream if stream is not None else torch.cuda.current_stream()
    return ctypes.c_void_p(s.cuda_stream)


class PinnedBuffer:
    """Page-locked host bytes from cds_host_alloc, exposed as a numpy uint8 array."""

    def __init__(self, nbytes):
        L = _lib.require_device()
        self._p = ctypes.c_void_p()
        _lib.check(L.cds_host_alloc(ctypes.byref(self._p), int(nbytes)))
        self.nbytes = int(nbytes)
        self.array = np.ctypeslib.as_array(ctypes.cast(self._p, ctypes.POINTER(ctypes.c_uint8)), shape=(max(self.nbytes, 1),))[:self.nbytes]

    def close(self):
        if self._p:
            _lib.load().cds_host_free(self._p)
            self._p = ctypes.c_void_p()
            self.array = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa
            pass


class DeviceSegment:
    """One decode context on one GPU.  Capacities are in 32 KB pages; all scratch is allocated here."""

    def __init__(self, max_heap_pages, max_toast_pages, device=None, max_chunks=0, scratch_bytes=None, stage=False):
        L = _lib.require_device()
        torch = _torch()
        if device is None:
            device = torch.cuda.current_device()
        self.device = int(torch.device("cuda", device).index if not isinstance(device, int) else device)
        self._h = ctypes.c_void_p()
        scratch = (2 ** 64 - 1) if scratch_bytes is None else int(scratch_bytes)
        _lib.check(L.cds_segment_create(ctypes.byref(self._h), self.device, int(max(1, max_heap_pages)),
                                        int(max_toast_pages), int(max_chunks), scratch, 1 if stage else 0))
        self.L = L
        self.max_heap_pages = int(max(1, max_heap_pages))
        self.max_toast_pages = int(max_toast_pages)
        self.cap_rows = self.max_heap_pages * 430

    def close(self):
        if getattr(self, "_h", None) and self._h:
            self.L.cds_segment_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa
            pass

    def set_output_format(self, ind_bf16=False, u8_normalize=None):
        """independent_var as stored (default), float32 -> bfloat16 on the fly, or — with `u8_normalize` =
        a float32 [channels, 256] table — uint8 pixels -> normalised float32 through that table."""
        if u8_normalize is not None:
            lut = np.ascontiguousarray(u8_normalize, dtype=np.float32)
            assert lut.ndim == 2 and lut.shape[1] == 256
            _lib.check(self.L.cds_set_normalize_lut(self._h, _ptr(lut), lut.shape[0]))
            _lib.check(self.L.cds_set_output_format(self._h, 2))
        else:
            _lib.check(self.L.cds_set_output_format(self._h, 1 if ind_bf16 else 0))

    # ---- step API -------------------------------------------------------------------------
    def heap_scan(self, heap, stream=None):
        _lib.check(self.L.cds_heap_scan(self._h, _ptr(heap), heap.numel() // BLCKSZ, _stream_ptr(stream)))

    def set_rows(self, rows, inline_blob, stream=None):
        rows = np.ascontiguousarray(rows, dtype=np.uint32).reshape(-1, ROW_WORDS)
        _lib.check(self.L.cds_set_rows(self._h, _ptr(rows), rows.shape[0], _ptr(inline_blob), _stream_ptr(stream)))

    def toast_scan(self, toast, stream=None):
        _lib.check(self.L.cds_toast_scan(self._h, _ptr(toast), toast.numel() // BLCKSZ, _stream_ptr(stream)))

    def plan(self, ind_sample_bytes=0, dep_sample_bytes=0, stream=None):
        _lib.check(self.L.cds_plan(self._h, int(ind_sample_bytes), int(dep_sample_bytes), _stream_ptr(stream)))

    def materialize(self, out_ind, out_dep, stream=None):
        _lib.check(self.L.cds_materialize(self._h, _ptr(out_ind), out_ind.numel() * out_ind.element_size(),
                                          _ptr(out_dep), out_dep.numel() * out_dep.element_size(), _stream_ptr(stream)))

    def finish(self, stream=None):
        """Synchronise and return the Summary (status = first data error in reference order)."""
        s = Summary()
        _lib.check(self.L.cds_finish(self._h, _stream_ptr(stream), ctypes.byref(s)))
        return s

    # ---- whole path -----------------------------------------------------------------------
    def decode_device(self, heap, toast, out_ind, out_dep, ind_sample_bytes=0, dep_sample_bytes=0, stream=None,
                      sync=True):
        """Pages resident in HBM (uint8 CUDA tensors) -> values in `out_ind` / `out_dep`."""
        s = Summary()
        rc = self.L.cds_decode_device(self._h, _ptr(heap), heap.numel() // BLCKSZ, _ptr(toast), toast.numel() // BLCKSZ,
                                      int(ind_sample_bytes), int(dep_sample_bytes),
                                      _ptr(out_ind), out_ind.numel() * out_ind.element_size(),
                                      _ptr(out_dep), out_dep.numel() * out_dep.element_size(),
                                      _stream_ptr(stream), ctypes.byref(s) if sync else None)
        _lib.check(rc)
        return s if sync else None

    def decode_host(self, heap_host, toast_host, out_ind, out_dep, ind_sample_bytes=0, dep_sample_bytes=0,
                    stream=None, sync=True):
        """Pinned HOST page buffers (numpy uint8 / pinned tensors) -> values in device tensors.
        Needs `stage=True` at construction."""
        s = Summary()
        hb = heap_host.size if isinstance(heap_host, np.ndarray) else heap_host.numel()
        tb = toast_host.size if isinstance(toast_host, np.ndarray) else toast_host.numel()
        rc = self.L.cds_decode_host(self._h, _ptr(heap_host), hb // BLCKSZ, _ptr(toast_host), tb // BLCKSZ,
                                    int(ind_sample_bytes), int(dep_sample_bytes),
                                    _ptr(out_ind), out_ind.numel() * out_ind.element_size(),
                                    _ptr(out_dep), out_dep.numel() * out_dep.element_size(),
                                    _stream_ptr(stream), ctypes.byref(s) if sync else None)
        _lib.check(rc)
        return s if sync else None

    # ---- measurement ----------------------------------------------------------------------
    def timing_enable(self, max_samples):
        _lib.check(self.L.cds_timing_enable(self._h, int(max_samples)))

    def timing_read(self, cap=4096):
        ms = np.zeros(cap, dtype=np.float32)
        n = ctypes.c_uint32()
        _lib.check(self.L.cds_timing_read(self._h, _ptr(ms), cap, ctypes.byref(n)))
        return ms[:n.value].copy()

    def timing_read_stage(self, stage, cap=4096):
        """Durations (ms) of one timed stage (0 gather, 1 pglz stage, 2 / 3 the wide block passes); call before timing_read."""
        ms = np.zeros(cap, dtype=np.float32)
        n = ctypes.c_uint32()
        _lib.check(self.L.cds_timing_read_stage(self._h, int(stage), _ptr(ms), cap, ctypes.byref(n)))
        return ms[:n.value].copy()

    # ---- exports --------------------------------------------------------------------------
    def rows(self, stream=None):
        out = np.zeros((self.cap_rows, ROW_WORDS), dtype=np.uint32)
        n = ctypes.c_uint64()
        _lib.check(self.L.cds_export_rows(self._h, _ptr(out), self.cap_rows, ctypes.byref(n), _stream_ptr(stream)))
        return out[:n.value].copy()

    def layout(self, n_rows, n_positions, stream=None):
        cap = max(1, int(n_rows))
        order = np.zeros(cap, dtype=np.uint64)
        arrs = [np.zeros(cap, dtype=np.uint64) for _ in range(4)]
        _lib.check(self.L.cds_export_layout(self._h, _ptr(order), _ptr(arrs[0]), _ptr(arrs[1]), _ptr(arrs[2]),
                                            _ptr(arrs[3]), cap, _stream_ptr(stream)))
        return {"order": order[:int(n_positions)].astype(np.int64), "ind_off": arrs[0][:n_rows], "dep_off": arrs[1][:n_rows],
                "ind_len": arrs[2][:n_rows], "dep_len": arrs[3][:n_rows]}

    def chunks(self, cap, stream=None):
        cap = max(1, int(cap))
        cid, seq = np.zeros(cap, dtype=np.uint32), np.zeros(cap, dtype=np.uint32)
        off, ln = np.zeros(cap, dtype=np.uint64), np.zeros(cap, dtype=np.uint32)
        n = ctypes.c_uint64()
        _lib.check(self.L.cds_export_chunks(self._h, _ptr(cid), _ptr(seq), _ptr(off), _ptr(ln), cap, ctypes.byref(n),
                                            _stream_ptr(stream)))
        k = int(n.value)
        return cid[:k], seq[:k], off[:k], ln[:k]


# ---- casts (a14) ----------------------------------------------------------------------------

def onehot_argmax(onehot, stream=None):
    """int16[N, C] CUDA tensor -> int64[N] labels (np.argmax semantics: first maximum)."""
    torch = _torch()
    L = _lib.require_device()
    assert onehot.dtype == torch.int16 and onehot.is_cuda and onehot.is_contiguous() and onehot.dim() == 2
    out = torch.empty(onehot.shape[0], dtype=torch.int64, device=onehot.device)
    with torch.cuda.device(onehot.device):
        _lib.check(L.cds_onehot_argmax_i16(_ptr(onehot), onehot.shape[0], onehot.shape[1], _ptr(out), _stream_ptr(stream)))
    return out


def nhwc_to_nchw(x, stream=None):
    """f32[N, H, W, C] CUDA tensor -> f32[N, C, H, W] (np.rollaxis(img, 2, 0) per sample)."""
    torch = _torch()
    L = _lib.require_device()
    assert x.dtype == torch.float32 and x.is_cuda and x.is_contiguous() and x.dim() == 4
    n, h, w, c = x.shape
    out = torch.empty((n, c, h, w), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(L.cds_nhwc_to_nchw_f32(_ptr(x), n, h * w, c, _ptr(out), _stream_ptr(stream)))
    return out


def to_bf16(x, stream=None):
    torch = _torch()
    L = _lib.require_device()
    assert x.dtype == torch.float32 and x.is_cuda and x.is_contiguous()
    out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(L.cds_f32_to_bf16(_ptr(x), x.numel(), _ptr(out), _stream_ptr(stream)))
    return out


def pglz_decompress_batch(datums, device=None):
    """Decode a list of pglz varlenas (bytes, each starting with the 8-byte header) on the GPU.
    Returns (list of bytes or None, list of status codes)."""
    import struct
    torch = _torch()
    L = _lib.require_device()
    if device is None:
        device = torch.cuda.current_device()
    n = len(datums)
    if n == 0:
        return [], []
    src_off = np.zeros(n + 1, dtype=np.uint64)
    dst_off = np.zeros(n + 1, dtype=np.uint64)
    for i, d in enumerate(datums):
        src_off[i + 1] = src_off[i] + len(d)            # exact: `avail` bounds the header's length word
        raw = struct.unpack_from("<I", d, 4)[0] if len(d) >= 8 else 0
        dst_off[i + 1] = dst_off[i] + ((raw + 15) & ~15)
    blob = np.zeros(int(src_off[-1]) + 16, dtype=np.uint8)
    for i, d in enumerate(datums):
        blob[int(src_off[i]):int(src_off[i]) + len(d)] = np.frombuffer(d, dtype=np.uint8)
    status = np.zeros(n, dtype=np.int32)
    with torch.cuda.device(device):
        d_src = torch.from_numpy(blob).cuda()
        d_dst = torch.zeros(int(dst_off[-1]) + 16, dtype=torch.uint8, device="cuda")
        # output slots are padded to 16 bytes; the pad only loosens the room bound
        _lib.check(L.cds_pglz_decompress_batch(int(device), _ptr(d_src), _ptr(src_off), _ptr(d_dst), _ptr(dst_off), n,
                                               _ptr(status), _stream_ptr()))
        host = d_dst.cpu().numpy()
    out, codes = [], []
    for i, d in enumerate(datums):
        raw = struct.unpack_from("<I", d, 4)[0] if len(d) >= 8 else 0
        codes.append(int(status[i]))
        out.append(host[int(dst_off[i]):int(dst_off[i]) + raw].tobytes() if status[i] == 0 else None)
    return out, codes


class DecodedPartition:
    """Result of one segment decode, resident on the GPU.

    ind / dep are flat uint8 CUDA tensors holding the values back to back in result order
    (the reference's dict insertion order); `order`, `rows`, offsets and lengths are small host arrays.
    """

    def __init__(self, summary, rows, layout, ind, dep):
        self.summary, self.rows, self.layout, self.ind, self.dep = summary, rows, layout, ind, dep

    @property
    def buffer_ids(self):
        return [int(self.rows[r, 1]) for r in self.layout["order"]]

    def value_views(self, ind_host, dep_host):
        """Per-buffer numpy views (float32 / int16, flat) over host copies of ind / dep."""
        out = []
        lay = self.layout
        for r in lay["order"]:
            io, il = int(lay["ind_off"][r]), int(lay["ind_len"][r])
            do, dl = int(lay["dep_off"][r]), int(lay["dep_len"][r])
            out.append((int(self.rows[r, 1]), ind_host[io:io + il].view(np.float32), dep_host[do:do + dl].view(np.int16)))
        return out


def decode_pages(heap, toast, device=None, ind_sample_bytes=0, dep_sample_bytes=0, rows=None, inline_blob=None,
                 segment=None, ind_bf16=False, u8_normalize=None, want_chunk_ids=False):
    """Two-phase decode of page buffers already on the GPU (uint8 CUDA tensors): scan + plan, one
    host sync to size the outputs, then materialize.  With `rows` (exported-row layout) the heap scan
    is replaced by an upload of those rows (`toast_page_read` path).  Returns (status, DecodedPartition)."""
    torch = _torch()
    dev = toast.device if toast is not None else heap.device
    with torch.cuda.device(dev):
        seg = segment or DeviceSegment(max(1, (heap.numel() // BLCKSZ) if heap is not None else (len(rows) // 400 + 1)),
                                       toast.numel() // BLCKSZ, device=dev.index)
        try:
            seg.set_output_format(ind_bf16, u8_normalize)
            if rows is None:
                seg.heap_scan(heap)
            else:
                seg.set_rows(rows, inline_blob)
            seg.toast_scan(toast)
            seg.plan(ind_sample_bytes, dep_sample_bytes)
            s1 = seg.finish()
            if s1.status != 0:
                return int(s1.status), DecodedPartition(s1, None, None, None, None)
            ind_out_bytes = int(s1.ind_bytes) * 4 if u8_normalize is not None else (int(s1.ind_bytes) // 2 if ind_bf16 else int(s1.ind_bytes))
            ind = torch.empty(ind_out_bytes + 16, dtype=torch.uint8, device=dev)
            dep = torch.empty(int(s1.dep_bytes) + 16, dtype=torch.uint8, device=dev)
            seg.materialize(ind, dep)
            s2 = seg.finish()
            out_rows = seg.rows()
            lay = seg.layout(int(s2.n_rows), int(s2.n_positions))
            part = DecodedPartition(s2, out_rows, lay, ind[:ind_out_bytes], dep[:int(s1.dep_bytes)])
            if want_chunk_ids:                                   # chunk ids in the reference's yield order (ppr:412-422)
                part.chunk_ids = seg.chunks(max(1, int(s2.n_chunks)))[0]
            return int(s2.status), part
        finally:
            if segment is None:
                seg.close()
