"""ctypes binding of libcds_da.so (include/cds_da.h).

There is no CPU fallback: if the library is missing, or no CUDA device is present when a compute
entry point is called, this raises.  `load()` only dlopens and declares prototypes, so symbol
checks work on a machine without a GPU.
"""
import ctypes
import os

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CDS_DA_LIB") or os.path.join(PKG_DIR, "libcds_da.so")     # (override: A/B runs of two builds)

(CDS_OK, CDS_ERR_PARTIAL_PAGE, CDS_ERR_ITEMID_ALIGN, CDS_ERR_TOAST_SPECIAL, CDS_ERR_TUPLE, CDS_ERR_RAWSIZE,
 CDS_ERR_CHUNK_COUNT, CDS_ERR_CHUNK_EXTENDED, CDS_ERR_MAX_CHUNK, CDS_ERR_CHUNK_SIZE, CDS_ERR_FINAL_SIZE,
 CDS_ERR_PGLZ_CORRUPT, CDS_ERR_SHAPE, CDS_ERR_CATALOG_MISS, CDS_ERR_MISSING_VALUE) = range(15)
CDS_ERR_TEXT_NUMBER = 15
CDS_ERR_CUDA, CDS_ERR_ARG, CDS_ERR_CAPACITY, CDS_ERR_NO_DEVICE = 100, 101, 102, 103

ROW_WORDS = 16
BLCKSZ = 32768


class Summary(ctypes.Structure):
    _fields_ = [("n_rows", ctypes.c_uint64), ("n_positions", ctypes.c_uint64), ("ind_bytes", ctypes.c_uint64),
                ("dep_bytes", ctypes.c_uint64), ("n_chunks", ctypes.c_uint64), ("n_inline", ctypes.c_uint64),
                ("scratch_bytes", ctypes.c_uint64), ("err_where", ctypes.c_uint64), ("status", ctypes.c_uint32),
                ("err_stage", ctypes.c_uint32), ("n_wide_values", ctypes.c_uint64)]


class CdsError(RuntimeError):
    """Infrastructure failure of the CUDA library (not a data error)."""

    def __init__(self, code, detail=""):
        super().__init__("libcds_da: %s (code %d) %s" % (_status_str(code), code, detail))
        self.code = code


# Exported symbols and their prototypes: (restype, argtypes).  tests/test_abi.py checks that the
# built library exports exactly the names include/cds_da.h declares.
_vp, _u64, _u32, _int = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int
_pu64, _pu32 = ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint32)
PROTOTYPES = {
    "cds_version": (_int, []),
    "cds_summary_size": (_u32, []),
    "cds_status_str": (ctypes.c_char_p, [_int]),
    "cds_last_error": (ctypes.c_char_p, []),
    "cds_device_count": (_int, []),
    "cds_host_alloc": (_int, [ctypes.POINTER(_vp), _u64]),
    "cds_host_free": (_int, [_vp]),
    "cds_segment_create": (_int, [ctypes.POINTER(_vp), _int, _u64, _u64, _u64, _u64, _int]),
    "cds_segment_destroy": (_int, [_vp]),
    "cds_set_output_format": (_int, [_vp, _int]),
    "cds_set_normalize_lut": (_int, [_vp, _vp, _u32]),
    "cds_heap_scan": (_int, [_vp, _vp, _u64, _vp]),
    "cds_set_rows": (_int, [_vp, _vp, _u64, _vp, _vp]),
    "cds_toast_scan": (_int, [_vp, _vp, _u64, _vp]),
    "cds_plan": (_int, [_vp, _u32, _u32, _vp]),
    "cds_materialize": (_int, [_vp, _vp, _u64, _vp, _u64, _vp]),
    "cds_finish": (_int, [_vp, _vp, ctypes.POINTER(Summary)]),
    "cds_decode_device": (_int, [_vp, _vp, _u64, _vp, _u64, _u32, _u32, _vp, _u64, _vp, _u64, _vp,
                                 ctypes.POINTER(Summary)]),
    "cds_decode_host": (_int, [_vp, _vp, _u64, _vp, _u64, _u32, _u32, _vp, _u64, _vp, _u64, _vp,
                               ctypes.POINTER(Summary)]),
    "cds_set_host_streaming": (_int, [_vp, _u32]),
    "cds_stream_counts": (_int, [ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(ctypes.c_ulonglong)]),
    "cds_launch_count": (ctypes.c_ulonglong, []),
    "cds_timing_enable": (_int, [_vp, _u32]),
    "cds_timing_read": (_int, [_vp, _vp, _u32, _pu32]),
    "cds_timing_read_stage": (_int, [_vp, _u32, _vp, _u32, _pu32]),
    "cds_export_rows": (_int, [_vp, _vp, _u64, _pu64, _vp]),
    "cds_export_layout": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp]),
    "cds_export_chunks": (_int, [_vp, _vp, _vp, _vp, _vp, _u64, _pu64, _vp]),
    "cds_pglz_decompress_batch": (_int, [_int, _vp, _vp, _vp, _vp, _u64, _vp, _vp]),
    "cds_pglz_decompress_batch_static": (_int, [_int, _vp, _vp, _vp, _vp, _u64, _vp, _vp, _vp]),
    "cds_set_wide_pglz_threshold": (_int, [_u32, _int]),
    "cds_set_wide_store_limit": (_int, [_u64]),
    "cds_set_wide_pglz": (_int, [_vp, _int]),
    "cds_trim_work_memory": (_int, [_int]),
    "cds_wide_decoded_count": (ctypes.c_ulonglong, []),
    "pglz_decompress": (None, [ctypes.c_char_p, ctypes.c_char_p]),
    "cds_onehot_argmax_i16": (_int, [_vp, _u64, _u32, _vp, _vp]),
    "cds_nhwc_to_nchw_f32": (_int, [_vp, _u64, _u32, _u32, _vp, _vp]),
    "cds_f32_to_bf16": (_int, [_vp, _u64, _vp, _vp]),
    "cds_checksum64": (_int, [_vp, _u64, _vp, _vp]),
    "cds_peer_copy": (_int, [_vp, _vp, _u64, _vp]),
    "cds_text_etl": (_int, [_int, _vp, _u64, _u32, _u32, _vp, _vp, _vp, _u64, _pu64, _pu64, _vp]),
    "cds_batch_assemble": (_int, [_vp, _u64, _u32, _u32, _vp, _vp, _u32, _vp, _int, _vp, _vp]),
}

_lib = None


def load():
    """dlopen libcds_da.so and declare prototypes.  Raises if the library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libcds_da.so is missing (%s). Build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "This package has no CPU fallback." % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _status_str(code):
    try:
        return load().cds_status_str(int(code)).decode()
    except Exception:  # noqa
        return "status"


def require_device():
    L = load()
    if L.cds_device_count() <= 0:
        raise CdsError(CDS_ERR_NO_DEVICE, "- a CUDA device is required; there is no CPU path")
    return L


def check(code):
    """Raise CdsError for infrastructure errors (>= 100); data errors (1..14) are returned."""
    if code >= 100:
        raise CdsError(code, load().cds_last_error().decode())
    return code
