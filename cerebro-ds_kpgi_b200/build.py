"""In-tree build recipes: explicit nvcc / gcc command lines, no JIT cache.

The built `.so` files land next to the sources so that they travel to the GPU box with the
repo snapshot (they are git-ignored).  Run as `python cerebro-ds_kpgi_b200/build.py [target...]`
or through `__graft_entry__.build()`.
"""
import hashlib
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
ORACLE_DIR = os.path.join(REPO_ROOT, "oracle")
REFERENCE_DIR = "/root/reference"

LIB_DA = os.path.join(PKG_DIR, "libcds_da.so")
LIB_SYNTH = os.path.join(PKG_DIR, "libcds_synth.so")
LIB_ORACLE = os.path.join(ORACLE_DIR, "libda_oracle.so")
LIB_REF_PGLZ = os.path.join(ORACLE_DIR, "_ref", "pg_lzcompress.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "--shared", "-Xcompiler", "-fPIC,-O3",
    "-Xptxas", "-v",
]


def _digest(paths, extra=""):
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def _sources(root, exts):
    out = []
    for d, _, files in os.walk(root):
        for f in files:
            if f.endswith(exts):
                out.append(os.path.join(d, f))
    return out


def _up_to_date(target, stamp_value):
    stamp = target + ".stamp"
    if not (os.path.exists(target) and os.path.exists(stamp)):
        return False
    with open(stamp) as f:
        return f.read().strip() == stamp_value


def _write_stamp(target, stamp_value):
    with open(target + ".stamp", "w") as f:
        f.write(stamp_value)


def _run(cmd, log_name=None):
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log_name:
        with open(log_name, "w") as f:
            f.write(" ".join(cmd) + "\n" + res.stdout)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd))
    return res.stdout


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found; the CUDA library cannot be built on this machine")
    return p


def build_cuda(force=False):
    """libcds_da.so: the C-ABI library (include/cds_da.h) with every sm_100a kernel."""
    srcs = _sources(CSRC, (".cu", ".cuh", ".h", ".inc")) + _sources(os.path.join(REPO_ROOT, "include"), (".h",))
    stamp = _digest(srcs, " ".join(NVCC_FLAGS))
    if not force and _up_to_date(LIB_DA, stamp):
        return LIB_DA
    cmd = [nvcc_path()] + NVCC_FLAGS + [
        "-I", os.path.join(REPO_ROOT, "include"), "-I", CSRC,
        os.path.join(CSRC, "cds_da.cu"), "-o", LIB_DA, "-lcudart",
    ]
    _run(cmd, log_name=os.path.join(PKG_DIR, "build_cuda.log"))
    _write_stamp(LIB_DA, stamp)
    return LIB_DA


def build_synth(force=False):
    """libcds_synth.so: host-only synthetic segment writer (test / bench input generator)."""
    src = os.path.join(CSRC, "synth_host.c")
    stamp = _digest([src])
    if not force and _up_to_date(LIB_SYNTH, stamp):
        return LIB_SYNTH
    _run(["gcc", "-O3", "-shared", "-fPIC", "-o", LIB_SYNTH, src])
    _write_stamp(LIB_SYNTH, stamp)
    return LIB_SYNTH


def build_oracle(force=False):
    """oracle/libda_oracle.so (our C restatement) and, when /root/reference is mounted,
    oracle/_ref/pg_lzcompress.so compiled from the reference's own source where it lies."""
    src = os.path.join(ORACLE_DIR, "da_oracle.c")
    out = []
    if os.path.exists(src):
        stamp = _digest([src])
        if force or not _up_to_date(LIB_ORACLE, stamp):
            _run(["gcc", "-O3", "-fopenmp", "-shared", "-fPIC", "-o", LIB_ORACLE, src])
            _write_stamp(LIB_ORACLE, stamp)
        out.append(LIB_ORACLE)
    ref_src = os.path.join(REFERENCE_DIR, "cerebro_gpdb", "pg_lzcompress.c")
    if os.path.exists(ref_src):
        os.makedirs(os.path.dirname(LIB_REF_PGLZ), exist_ok=True)
        stamp = _digest([ref_src])
        if force or not _up_to_date(LIB_REF_PGLZ, stamp):
            _run(["gcc", "-O2", "-shared", "-fPIC", "-o", LIB_REF_PGLZ, ref_src])
            _write_stamp(LIB_REF_PGLZ, stamp)
        out.append(LIB_REF_PGLZ)
    return out


def build_all(force=False):
    return {
        "synth": build_synth(force),
        "cuda": build_cuda(force),
        "oracle": build_oracle(force),
    }


if __name__ == "__main__":
    targets = sys.argv[1:] or ["all"]
    force = "--force" in targets
    targets = [t for t in targets if not t.startswith("--")] or ["all"]
    for t in targets:
        fn = {"all": build_all, "cuda": build_cuda, "synth": build_synth, "oracle": build_oracle}[t]
        print(t, "->", fn(force))
