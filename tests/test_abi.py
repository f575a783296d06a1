"""The C-ABI library loads on a machine without a GPU and exports exactly what include/cds_da.h declares.
No compute entry point is called here."""
import os
import re
import subprocess

import pytest

from conftest import REPO


@pytest.fixture(scope="module")
def lib_path():
    from cerebro_ds_kpgi_b200 import build
    return build.build_cuda()


def header_functions():
    src = open(os.path.join(REPO, "include", "cds_da.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b([a-z_][a-z0-9_]*)\s*\([^;{}]*\)\s*;", src)
    return sorted(set(n for n in names if n.startswith("cds_") or n == "pglz_decompress"))


def test_header_and_library_agree(lib_path):
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True, check=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l and (" cds_" in l or l.endswith(" pglz_decompress")))
    declared = header_functions()
    assert declared, "no declarations parsed"
    assert exported == declared


def test_ctypes_prototypes_cover_header(lib_path):
    from cerebro_ds_kpgi_b200 import _lib
    L = _lib.load()
    assert sorted(_lib.PROTOTYPES) == header_functions()
    assert L.cds_version() >= 100
    assert L.cds_status_str(11) == b"compressed data is corrupt"
    assert L.cds_status_str(1) == b"FATAL"


def test_documented_summary_binding_matches_the_library(lib_path):
    """The Summary struct printed in INTEGRATION.md §2 is what a maintainer pastes: it must have the size the
    library writes (cds_finish fills the whole cds_summary)."""
    import ctypes
    from cerebro_ds_kpgi_b200 import _lib
    L = _lib.load()
    text = open(os.path.join(REPO, "INTEGRATION.md")).read()
    m = re.search(r"class Summary\(ctypes\.Structure\):\n(?:    .*\n|\s*\[.*\n)+", text)
    assert m, "INTEGRATION.md no longer shows the Summary binding"
    ns = {"ctypes": ctypes}
    exec(m.group(0), ns)
    assert ctypes.sizeof(ns["Summary"]) == L.cds_summary_size() == ctypes.sizeof(_lib.Summary) == 80
    assert [f[0] for f in ns["Summary"]._fields_] == [f[0] for f in _lib.Summary._fields_]


def test_no_cpu_fallback(lib_path):
    """Without a CUDA device every compute entry point must fail loudly."""
    import ctypes
    from cerebro_ds_kpgi_b200 import _lib
    L = _lib.load()
    if L.cds_device_count() > 0:
        pytest.skip("a GPU is present")
    h = ctypes.c_void_p()
    assert L.cds_segment_create(ctypes.byref(h), 0, 1, 1, 0, 0, 0) == _lib.CDS_ERR_NO_DEVICE
    with pytest.raises(_lib.CdsError):
        _lib.require_device()
    from cerebro_ds_kpgi_b200 import pg_page_reader
    with pytest.raises(_lib.CdsError):
        pg_page_reader.pglz_decompress(b"\x40\x00\x00\x09\x01\x00\x00\x00\x00a", bytearray(1), "py")


def test_product_does_not_import_oracle():
    """oracle/ is test infrastructure: nothing in the package may import, load or execute it."""
    pkg = os.path.join(REPO, "cerebro-ds_kpgi_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".c", ".h")):
                text = open(os.path.join(d, f)).read()
                for bad in ("import da_oracle", "from da_oracle", "libda_oracle", "oracle/_ref", "da_oracle_"):
                    if f == "build.py" and bad in ("libda_oracle", "oracle/_ref"):
                        continue        # build.py only compiles the checker
                    assert bad not in text, (f, bad)


def test_sass_shows_tma_bulk_copy(lib_path):
    """The gather kernel stages pages with TMA bulk copies: SASS must show UBLKCP + mbarrier ops."""
    out = subprocess.run(["cuobjdump", "-sass", lib_path], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    fn = out.stdout.split("Function : ")
    gather = [f for f in fn if "k_toast_gather" in f.split("\n")[0]]
    assert gather and "UBLKCP" in gather[0] and "SYNCS" in gather[0]
