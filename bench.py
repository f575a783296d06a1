#!/usr/bin/env python
"""bench.py — Direct-Access decode throughput on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--only]

A "step" is one pass of the hot path over one synthetic segment per GPU (heap scan, TOAST walk,
plan, gather, pglz, int64 labels).  Workloads (BASELINE.json `configs`):

    imagenet_raw   configs[1]: ImageNet-shaped rows, uncompressed external bytea  (default; the
                   configuration the metric is quoted on)
    imagenet_pglz  configs[2]: mixed pglz-compressed / incompressible rows
    criteo         configs[3]: Criteo-shaped buffers (256 x 7306 f32, pglz) + inline labels
    imagenet_packed  SURVEY.md 8 f-3: the paper's packed buffers (--images-per-buffer B, default 3210 = 483 MB per
                   value, pglz): decoded by the intra-value parallel decoder

One JSON line on stdout (rank 0).  The top-level keys describe the headline workload (`--workload`, default
imagenet_raw): `value` = rows/s over all ranks with pages resident in HBM; `e2e` = the same through
cds_decode_host with pinned HOST page buffers (H2D inside the timed region, labels + status read back);
`roofline` = the workload's dominant kernel against measured HBM copy bandwidth; `cpu_baseline` = the CPU
restatement (oracle/) on this host's cores.  Unless `--only` is given the same invocation then runs the other
configurations for a few steps each and attaches them under `configs` — each with its own value, e2e, roofline
block for ITS dominant kernel and CPU rate — and, on more than one GPU, a model-hopper leg under `mop`.
Every workload is gated by a byte-exact check: a 64-bit checksum of the decoded buffers on the device against the
same checksum of the arrays the generator packed, before the warm-up and after the timed region.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

IMAGENET_ROW_BYTES = 112 * 112 * 3 * 4
IMAGES_PER_BUFFER = int(os.environ.get("CDS_BENCH_IMAGES_PER_BUFFER", "3210"))   # imagenet_packed (--images-per-buffer); workers of the reference arm inherit it through the environment
LABEL_ROW_BYTES = 1000 * 2
CRITEO_BUFFERS = int(os.environ.get("CDS_BENCH_CRITEO_BUFFERS", "0"))          # --criteo-buffers (0: sized from --toast-mib, at most 1024)
ALL_WORKLOADS = ["imagenet_raw", "imagenet_pglz", "criteo", "imagenet_packed"]
NVSMI_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
               "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
               "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
SUM_MUL = 0x9E3779B97F4A7C15
MASK64 = (1 << 64) - 1


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------------
# workloads
# ------------------------------------------------------------------------------------------------

class ExpectedSum:
    """The checksum cds_checksum64 computes on the device (sum of w_i * (2 i + 1) * SUM_MUL mod 2^64 over the 64-bit
    words of a buffer), accumulated over arrays as they are appended to the buffer.  Linear in the position, so an
    array that is packed many times is summed once: S(base) = A + 2 * base * SUM_MUL * W."""

    def __init__(self):
        self.total, self.words, self.cache = 0, 0, {}

    def add(self, arr):
        key = id(arr)
        if key not in self.cache:
            b = np.ascontiguousarray(arr).view(np.uint8).reshape(-1)
            assert b.size % 8 == 0, "value sizes of the bench workloads are multiples of 8"
            w = b.view(np.uint64)
            idx = np.arange(w.size, dtype=np.uint64)
            with np.errstate(over="ignore"):
                a = int(np.sum(w * ((idx * np.uint64(2) + np.uint64(1)) * np.uint64(SUM_MUL)), dtype=np.uint64))
                s = int(np.sum(w, dtype=np.uint64))
            self.cache[key] = (a, s, w.size, arr)                          # (keeps arr alive: ids stay unique)
        a, s, n, _ = self.cache[key]
        self.total = (self.total + a + 2 * self.words * SUM_MUL * s) & MASK64
        self.words += n


def workload_units(name, toast_mib):
    """Heap rows (= values per column) of the full-size workload."""
    from cerebro_ds_kpgi_b200 import synth
    target = toast_mib << 20
    if name == "imagenet_raw":
        return synth.rows_for_toast_bytes(target)
    if name == "imagenet_pglz":
        return int(synth.rows_for_toast_bytes(target) * 1.15)
    if name == "criteo":
        if CRITEO_BUFFERS:
            return CRITEO_BUFFERS      # --criteo-buffers: a partition-sized segment; distinct buffers repeat after 256
        return min(max(1, (target * 20) // (256 * 7306 * 4)), 1024)   # ~20x compression: size the TOAST file, not the output
    if name == "imagenet_packed":
        return max(1, int(target // (IMAGES_PER_BUFFER * IMAGENET_ROW_BYTES * 0.67)))
    raise SystemExit("unknown workload " + name)


def make_workload(name, seed, toast_mib, units=None, expected=True):
    """Synthetic segment per SURVEY.md §8(d).  Returns (segment, meta).  `units` overrides the row count that
    `toast_mib` implies (the reference arm splits a workload's rows over its worker processes)."""
    from cerebro_ds_kpgi_b200 import synth
    n = workload_units(name, toast_mib) if units is None else int(units)
    if name == "imagenet_raw":
        ind, dep = synth.imagenet_rows(n, seed=seed, kind="norm_u8")
        ind_st, dep_st = ["ext_raw"] * n, ["ext_raw"] * n
        meta = dict(isb=IMAGENET_ROW_BYTES, dsb=LABEL_ROW_BYTES, samples=n, classes=1000,
                    desc="ImageNet-shaped rows (112x112x3 f32 + int16[1000] one-hot), uncompressed external bytea")
        kw = {}
    elif name == "imagenet_pglz":
        a_ind, dep = synth.imagenet_rows(n, seed=seed, kind="norm_u8")
        rng = np.random.default_rng(seed + 1000)
        noise = np.random.default_rng(seed + 2000)
        ind, ind_st = [], []
        for i in range(n):
            if rng.random() < 0.5:
                ind.append(a_ind[i]); ind_st.append("ext_pglz")
            else:
                ind.append(noise.standard_normal(a_ind[i].shape, dtype=np.float32)); ind_st.append("ext_literal")
        dep_st = ["inline_pglz"] * n
        meta = dict(isb=IMAGENET_ROW_BYTES, dsb=LABEL_ROW_BYTES, samples=n, classes=1000,
                    desc="ImageNet-shaped rows, 50% pglz-compressed / 50% literal-only pglz, inline-compressed labels")
        kw = dict(effort=4)
    elif name == "criteo":
        B = 256            # a real partition holds ~6 300 such buffers (1 624 157 rows / 256)
        ind, dep = synth.criteo_buffers(min(n, 256), seed=seed, rows_per_buffer=B)
        ind, dep = [ind[i % len(ind)] for i in range(n)], [dep[i % len(dep)] for i in range(n)]
        ind_st, dep_st = ["ext_pglz"] * n, ["inline_pglz"] * n
        meta = dict(isb=7306 * 4, dsb=4, samples=n * B, classes=2,
                    desc="Criteo-shaped buffers (256 x 7306 f32 pglz, int16[256,2] inline pglz)")
        kw = dict(effort=4, toast_cap_pages=n * 16 + 64)
    elif name == "imagenet_packed":
        # the paper's table layout (load_imagenet.py:27-31): B images per buffer, one huge pglz value each
        B = IMAGES_PER_BUFFER
        ind, dep = synth.imagenet_rows(min(n, 4), seed=seed, kind="norm_u8", images_per_row=B)     # distinct buffers repeat after 4
        ind, dep = [ind[i % len(ind)] for i in range(n)], [dep[i % len(dep)] for i in range(n)]
        ind_st, dep_st = ["ext_pglz"] * n, ["ext_pglz"] * n
        meta = dict(isb=IMAGENET_ROW_BYTES, dsb=LABEL_ROW_BYTES, samples=n * B, classes=1000,
                    desc="ImageNet-shaped packed buffers: %d images per buffer (%d MB per value), pglz-compressed" % (B, B * IMAGENET_ROW_BYTES // 10 ** 6))
        kw = dict(effort=4)
    else:
        raise SystemExit("unknown workload " + name)
    seg = synth.build_segment(ind, dep, ind_storage=ind_st, dep_storage=dep_st, dist_key=seed - 2018, keep_expected=False, **kw)
    meta["units"] = n
    meta["ind_bytes"] = meta["samples"] * meta["isb"]
    meta["dep_bytes"] = meta["samples"] * meta["dsb"]
    meta["rows"] = meta["samples"]
    # which bytes go through which kernel: uncompressed and literal-only external values are unpacked by the gather,
    # everything else by the pglz stage (its input = the compressed streams, its output = their decoded bytes)
    gather_out = pglz_out = lit_stored = raw_stored = 0
    for arrs, sts in ((ind, ind_st), (dep, dep_st)):
        for a, st in zip(arrs, sts):
            nb = int(a.nbytes)
            if st == "ext_raw":
                gather_out += nb; raw_stored += nb
            elif st == "ext_literal":
                gather_out += nb; lit_stored += 4 + nb + (nb + 7) // 8
            else:
                pglz_out += nb
    meta["gather_out_bytes"] = gather_out
    meta["pglz_out_bytes"] = pglz_out
    meta["pglz_in_bytes"] = int(seg.stored_bytes) - lit_stored - raw_stored
    if expected:
        # the reader's result (dict insertion order = table order for these segments) as two checksums + the labels
        es_i, es_d = ExpectedSum(), ExpectedSum()
        labels = []
        lab_cache = {}
        for a, d in zip(ind, dep):
            es_i.add(a); es_d.add(d)
            if id(d) not in lab_cache:
                lab_cache[id(d)] = np.argmax(d.reshape(-1, meta["classes"]), axis=1).astype(np.int64)
            labels.append(lab_cache[id(d)])
        meta["sum_ind"], meta["sum_dep"] = es_i.total, es_d.total
        meta["labels"] = np.concatenate(labels) if labels else np.zeros(0, dtype=np.int64)
    return seg, meta


# ------------------------------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md recipe)
# ------------------------------------------------------------------------------------------------

class ClockSampler:
    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + NVSMI_QUERY, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except Exception as e:  # noqa
            log("clock sampler unavailable:", e)
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:  # noqa
                self.proc.kill()

    def summary(self, t0, t1):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.lines:
            if t < t0 - 0.15 or t > t1 + 0.15:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle on the host cores
# ------------------------------------------------------------------------------------------------

def python_pglz_rate():
    """MB/s of decoded output of the reference's as-shipped decoder, pglz_decompress_py (a Python loop per output byte,
    ppr:191-231; restated in oracle/da_oracle.py), on one ImageNet-shaped value of 150 KB, one process."""
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import da_oracle
    from cerebro_ds_kpgi_b200 import synth
    ind, _ = synth.imagenet_rows(1, seed=2018, kind="norm_u8")
    datum = synth.pglz_varlena(ind[0], "lz")
    t0 = time.perf_counter()
    out = da_oracle.pglz_decompress(datum)
    dt = time.perf_counter() - t0
    assert bytes(out) == ind[0].tobytes()
    return ind[0].nbytes / dt / 1e6


ORIGINAL_AFFINITY = None          # the cores this process was given, before any NUMA binding of the GPU arm


def host_threads():
    if ORIGINAL_AFFINITY is not None:
        return len(ORIGINAL_AFFINITY)
    try:
        return len(os.sched_getaffinity(0))
    except Exception:  # noqa
        return os.cpu_count() or 1


def cpu_decode_rate(seg, meta, reps, threads=0):
    """rows/s of oracle/da_oracle.c (OpenMP over values) on the whole segment; best of `reps`.
    threads=0: every core this process may run on (torchrun exports OMP_NUM_THREADS=1, so the count is
    passed explicitly)."""
    threads = threads or host_threads()
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import da_oracle
    from cerebro_ds_kpgi_b200 import build
    build.build_oracle()
    orc = da_oracle.COracle()
    out_ind = np.empty(meta["ind_bytes"] + 64, dtype=np.uint8)
    out_dep = np.empty(meta["dep_bytes"] + 64, dtype=np.uint8)
    out_ind[::4096] = 0
    out_dep[::4096] = 0          # fault the pages in before timing
    best = None
    for _ in range(reps):
        t0 = time.perf_counter()
        orc.decode(seg.heap, seg.toast, ind_sample_bytes=meta["isb"], dep_sample_bytes=meta["dsb"], nthreads=threads,
                   out_ind=out_ind, out_dep=out_dep)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return meta["rows"] / best, best, threads


# The reference arm ("reference-C x N"): the reference's reader is Python; its structure is
# restated in oracle/da_oracle.py, pglz goes through the reference's own pg_lzcompress.c (oracle/_ref) when that
# was compiled, and the work is spread over all host cores as disjoint row ranges, one sub-segment per process.
_REF_PART = None          # (heap, toast, buffers, pglz) of THIS worker process
_REF_BARRIER = None       # all workers of the pool: makes every worker take exactly one task of a map


def _ref_pglz():
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import da_oracle
    try:
        return da_oracle.RefPglz().pglz_decompress, "oracle/_ref/pg_lzcompress.so (the reference's own C decoder)"
    except RuntimeError:
        return da_oracle.COracle().pglz_decompress, "oracle/da_oracle.c pglz (oracle/_ref not built on this machine)"


def split_units(units, procs):
    """`units` rows over `procs` workers, the remainder spread over the first ones: the parts sum to `units` exactly."""
    return [units // procs + (1 if i < units % procs else 0) for i in range(procs)]


def _ref_init(counter, barrier, workload, toast_mib, parts):
    """Worker start-up (spawned, so nothing CUDA-related is inherited): build this worker's sub-segment."""
    global _REF_PART, _REF_BARRIER
    _REF_BARRIER = barrier
    with counter.get_lock():
        idx = counter.value
        counter.value += 1
    if parts[idx] == 0:
        _REF_PART = (None, None, 0, None)
        return
    seg, meta = make_workload(workload, 2018 + 100 * idx, toast_mib, units=parts[idx], expected=False)
    _REF_PART = (seg.heap, seg.toast, len(seg.rows), _ref_pglz()[0])


def _ref_worker(_):
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import da_oracle
    heap, toast, buffers, pglz = _REF_PART
    _REF_BARRIER.wait()                                               # one task per worker: the sub-segments differ in size
    if buffers == 0:
        return 0
    out = da_oracle.input_fn_bytes(heap, toast, None, pglz=pglz)
    assert len(out) == buffers
    return buffers


class PythonPortPool:
    """oracle/da_oracle.py (pure-Python page walk / detoast, the reference's loop structure) over all cores:
    `procs` spawned workers, each holding one sub-segment; together exactly the rows of the full workload."""

    def __init__(self, workload, toast_mib, procs):
        import multiprocessing as mp
        ctx = mp.get_context("spawn")
        units = workload_units(workload, toast_mib)
        self.procs = max(1, min(procs, units))
        self.parts = split_units(units, self.procs)
        self.buffers = units
        self.rows = units * {"criteo": 256, "imagenet_packed": IMAGES_PER_BUFFER}.get(workload, 1)
        self.pool = ctx.Pool(self.procs, initializer=_ref_init, initargs=(ctx.Value("i", 0), ctx.Barrier(self.procs), workload, toast_mib, self.parts))
        self.pglz_note = _ref_pglz()[1]
        self.step(1)                                              # every worker up and its sub-segment built

    def step(self, repeats=1):
        t0 = time.perf_counter()
        for _ in range(repeats):
            got = sum(self.pool.map(_ref_worker, range(self.procs), chunksize=1))
            assert got == self.buffers, (got, self.buffers)
        return time.perf_counter() - t0

    def close(self):
        self.pool.close()
        self.pool.join()

    def describe(self, repeats):
        return ("%d x the %d-row workload per step as %d sub-segments, one process each: oracle/da_oracle.py (Python restatement "
                "of pg_page_reader.py's page walk / detoast) with pglz through %s" % (repeats, self.rows, self.procs, self.pglz_note))


def workload_config(name, seg, meta, segments):
    """The `config` block: the same for both arms of a workload."""
    step_bytes = seg.heap.size + seg.toast.size + meta["ind_bytes"] + meta["dep_bytes"] + 8 * meta["rows"]
    return {"workload": name, "description": meta["desc"], "rows_per_segment": meta["rows"],
            "toast_mib": seg.toast.size >> 20, "heap_pages": seg.heap.size // 32768, "segments": segments,
            "partitioning": "one segment per GPU, no collective on the decode path",
            "l2": "inputs (%d MiB per step) larger than the 126 MB L2; no flush needed" % ((seg.toast.size + meta["ind_bytes"]) >> 20),
            "algorithmic_bytes_per_row": step_bytes / meta["rows"]}


def reference_leg(workload, toast_mib, n_seg, steps, warmup, threads, c_port=True):
    pool = PythonPortPool(workload, toast_mib, threads)
    times = []
    for i in range(warmup + steps):
        dt = pool.step(n_seg)
        if i >= warmup:
            times.append(dt)
    pool.close()
    ms = 1000.0 * float(np.mean(times))
    value = n_seg * pool.rows / (ms / 1000.0)
    seg, meta = make_workload(workload, 2018, toast_mib, expected=False)
    assert meta["rows"] == pool.rows
    alg_bytes = seg.heap.size + seg.toast.size + meta["ind_bytes"] + meta["dep_bytes"] + 8 * meta["rows"]
    out = {"value": value, "unit": "rows/s", "ms_per_step": ms, "steps": steps, "gb_per_s": value * alg_bytes / meta["rows"] / 1e9,
           "config": workload_config(workload, seg, meta, n_seg),
           "cpu_baseline": {"value": value, "unit": "rows/s", "cores": pool.procs, "kind": "port", "sample": pool.describe(n_seg)}}
    if c_port:
        # the much faster all-C restatement, for transparency (not what the reference ships)
        c_rate, c_dt, c_threads = cpu_decode_rate(seg, meta, 2 if workload == "imagenet_packed" else 3)
        out["c_port_openmp"] = {"value": c_rate, "unit": "rows/s", "cores": c_threads, "ms_per_segment": 1000.0 * c_dt,
                                "note": "oracle/da_oracle.c, the whole path in C with OpenMP over values: far faster than anything the "
                                        "reference ships (its reader is a Python loop); reported so that nothing is hidden"}
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    n_seg = max(1, args.gpus)          # the whole job: one segment per GPU of our arm, decoded on this host's cores
    head = reference_leg(args.workload, args.toast_mib, n_seg, args.steps, args.warmup, threads)
    line = {
        "impl": "reference", "metric": "decoded_rows_per_s", "value": head["value"], "unit": "rows/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic", "gb_per_s": head["gb_per_s"],
        "config": head["config"], "cpu_baseline": head["cpu_baseline"],
        "e2e": {"value": head["value"], "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "host_cores": threads, "c_port_openmp": head.get("c_port_openmp"),
    }
    if not args.only:
        line["configs"] = {}
        for w in extra_workloads(args):
            sub = reference_leg(w, args.toast_mib, n_seg, max(1, min(args.steps, 2)), 1, threads)
            sub["e2e"] = {"value": sub["value"], "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
            line["configs"][w] = sub
    emit(line)


def extra_workloads(args):
    """The configurations that ride along with the headline one: configs[2], configs[3] and the packed layout."""
    return [w for w in ("imagenet_raw", "imagenet_pglz", "criteo", "imagenet_packed") if w != args.workload]


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------

def load_profile_traffic(workload):
    """dram bytes per launch of the workload's dominant kernel from the committed ncu summary (profiles/), or None."""
    p = os.path.join(REPO, "profiles", "kernel_traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(workload)
    except Exception:  # noqa
        return None


class GpuContext:
    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from cerebro_ds_kpgi_b200 import _lib
        self.torch, self.dist, self.args = torch, dist, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            log("note: WORLD_SIZE=%d, --gpus=%d" % (self.world, args.gpus))
        self.L = _lib.require_device()
        torch.cuda.set_device(self.local)
        global ORIGINAL_AFFINITY
        try:
            ORIGINAL_AFFINITY = os.sched_getaffinity(0)
        except Exception:  # noqa
            ORIGINAL_AFFINITY = None
        from cerebro_ds_kpgi_b200 import hostnuma
        self.numa = hostnuma.bind_to_gpu_node(self.local)          # before any staging buffer is allocated or touched
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        peaks = {}
        try:
            with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:  # noqa
            pass
        self.peak = float(peaks.get("hbm_gbs", 6650.0))
        self.peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s (B200_PROFILING.md)"

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def restore_affinity(self):
        if ORIGINAL_AFFINITY is not None:
            try:
                os.sched_setaffinity(0, ORIGINAL_AFFINITY)
            except Exception:  # noqa
                pass


def device_checksum(ctx, t, nbytes):
    from cerebro_ds_kpgi_b200 import _lib, segment
    torch = ctx.torch
    out = torch.zeros(1, dtype=torch.int64, device="cuda")
    _lib.check(ctx.L.cds_checksum64(segment._ptr(t), int(nbytes), segment._ptr(out), segment._stream_ptr(torch.cuda.current_stream())))
    return int(out.item()) & MASK64


def gpu_leg(ctx, name, steps, warmup, graph_wanted, with_cpu, sampler=None):
    """One workload on this rank (+ the reduction over ranks).  Returns the result block (rank 0) or None."""
    from cerebro_ds_kpgi_b200 import _lib, segment
    torch, dist, args, L = ctx.torch, ctx.dist, ctx.args, ctx.L
    rank, world, local = ctx.rank, ctx.world, ctx.local
    t_gen = time.time()
    seg, meta = make_workload(name, 2018 + rank, args.toast_mib)
    heap_pages, toast_pages = seg.heap.size // 32768, seg.toast.size // 32768
    log("[rank %d] %s: %d rows, heap %d pages, TOAST %d pages (%.1f MiB), generated in %.1fs" % (
        rank, name, meta["rows"], heap_pages, toast_pages, seg.toast.size / 2 ** 20, time.time() - t_gen))

    ds = segment.DeviceSegment(heap_pages, toast_pages, device=local, stage=True)
    bf16 = args.out_dtype == "bf16"
    ds.set_output_format(bf16)
    _lib.check(L.cds_set_wide_pglz(ds._h, {"off": 0, "force": 1, "auto": 2, "host": 3}[args.wide]))
    if args.wide_store_limit >= 0:
        _lib.check(L.cds_set_wide_store_limit(int(args.wide_store_limit)))
    d_heap = torch.from_numpy(seg.heap).cuda()
    d_toast = torch.from_numpy(seg.toast).cuda()
    out_ind = torch.empty(meta["ind_bytes"] // (2 if bf16 else 1), dtype=torch.uint8, device="cuda")
    out_dep = torch.empty(meta["dep_bytes"], dtype=torch.uint8, device="cuda")
    onehot = out_dep.view(torch.int16).view(-1, meta["classes"])
    labels = torch.empty(meta["samples"], dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream()
    sp = segment._stream_ptr(stream)
    lab_ref = torch.from_numpy(meta["labels"]).cuda()

    def step_device():
        ds.decode_device(d_heap, d_toast, out_ind, out_dep, meta["isb"], meta["dsb"], sync=False)
        _lib.check(L.cds_onehot_argmax_i16(segment._ptr(onehot), meta["samples"], meta["classes"], segment._ptr(labels), sp))

    def poison():
        out_ind.fill_(0xA5); out_dep.fill_(0x5A); labels.fill_(-1)

    def gate(where):
        """Byte-exact check of the decoded buffers against the generator's arrays (64-bit checksums) + the labels."""
        if not bf16:
            got = device_checksum(ctx, out_ind, meta["ind_bytes"])
            if got != meta["sum_ind"]:
                raise SystemExit("%s %s: independent_var checksum %016x, expected %016x" % (name, where, got, meta["sum_ind"]))
        got = device_checksum(ctx, out_dep, meta["dep_bytes"])
        if got != meta["sum_dep"]:
            raise SystemExit("%s %s: dependent_var checksum %016x, expected %016x" % (name, where, got, meta["sum_dep"]))
        if not torch.equal(labels, lab_ref):
            raise SystemExit("%s %s: labels differ" % (name, where))

    # ---- correctness gate before timing
    poison()
    ds.decode_device(d_heap, d_toast, out_ind, out_dep, meta["isb"], meta["dsb"], sync=False)
    s = ds.finish()                  # (the decode is complete when cds_finish returns; consumers are enqueued after it)
    _lib.check(L.cds_onehot_argmax_i16(segment._ptr(onehot), meta["samples"], meta["classes"], segment._ptr(labels), sp))
    if s.status != 0 or s.ind_bytes != meta["ind_bytes"] or s.dep_bytes != meta["dep_bytes"]:
        raise SystemExit("%s: decode failed: status %d, %d/%d bytes" % (name, s.status, s.ind_bytes, meta["ind_bytes"]))
    gate("before the warm-up")

    # ---- device-resident timing (value) ---------------------------------------------------------
    for _ in range(warmup):
        step_device()
    ds.finish()
    wide0 = int(L.cds_wide_decoded_count())
    poison()
    ds.timing_enable(steps)
    ctx.barrier()
    launches0 = L.cds_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    graph = None
    if graph_wanted:
        # The K steps of the timed region as ONE CUDA graph (the same kernels and memsets; the timing events become
        # event-record nodes): replay removes the host launch gaps between the short front-end kernels.
        try:
            graph = torch.cuda.CUDAGraph()
            cap_stream = torch.cuda.Stream()
            cap_stream.wait_stream(stream)
            with torch.cuda.graph(graph, stream=cap_stream):
                cur = torch.cuda.current_stream()
                for _ in range(steps):
                    ds.decode_device(d_heap, d_toast, out_ind, out_dep, meta["isb"], meta["dsb"], stream=cur, sync=False)
                    _lib.check(L.cds_onehot_argmax_i16(segment._ptr(onehot), meta["samples"], meta["classes"], segment._ptr(labels),
                                                       segment._stream_ptr(cur)))
            launches = int(L.cds_launch_count() - launches0)          # kernels captured = kernels one replay runs
            graph.replay()                                            # one untimed replay
            torch.cuda.synchronize()
        except Exception as e:  # noqa
            log("[rank %d] CUDA graph capture unavailable (%s): timing eager launches" % (rank, str(e).splitlines()[0]))
            graph = None
            torch.cuda.synchronize()
            ds.timing_enable(steps)
            launches0 = L.cds_launch_count()
    ctx.barrier()
    t_wall0 = time.time()
    ev0.record(stream)
    if graph is not None:
        graph.replay()
    else:
        for _ in range(steps):
            step_device()
    ev1.record(stream)
    torch.cuda.synchronize()
    ctx.barrier()
    if graph is None:
        launches = int(L.cds_launch_count() - launches0)
    s = ds.finish()
    assert s.status == 0
    went_wide = int(L.cds_wide_decoded_count()) > wide0
    blocks_ms = None
    dev_ms = ev0.elapsed_time(ev1) / steps
    stage_ms = ds.timing_read_stage(1, steps)
    b1_ms, b3_ms = ds.timing_read_stage(2, steps), ds.timing_read_stage(3, steps)
    gather_ms = ds.timing_read(steps)
    ds.timing_enable(0)
    gate("after the timed region")

    # ---- end to end: pinned host pages -> H2D -> decode -> labels + status back -------------------
    ph, pt = segment.PinnedBuffer(seg.heap.size), segment.PinnedBuffer(seg.toast.size)
    ph.array[:] = seg.heap
    pt.array[:] = seg.toast
    labels_host = torch.empty(meta["samples"], dtype=torch.int64).pin_memory()

    def step_e2e():
        ds.decode_host(ph.array, pt.array, out_ind, out_dep, meta["isb"], meta["dsb"], sync=False)
        _lib.check(L.cds_onehot_argmax_i16(segment._ptr(onehot), meta["samples"], meta["classes"], segment._ptr(labels), sp))
        labels_host.copy_(labels, non_blocking=True)
        return ds.finish()

    e2e_steps = max(3, min(steps, 10))
    for _ in range(2):
        step_e2e()
    poison()
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(e2e_steps):
        st = step_e2e()
        assert st.status == 0
    e1.record(stream)
    torch.cuda.synchronize()
    ctx.barrier()
    e2e_ms = e0.elapsed_time(e1) / e2e_steps
    t_wall2 = time.time()
    gate("after the end-to-end steps")
    assert np.array_equal(labels_host.numpy(), meta["labels"])

    # H2D alone, against PCIe
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pin_t = torch.from_numpy(pt.array)
    d_toast.copy_(pin_t, non_blocking=True)
    torch.cuda.synchronize()
    ctx.barrier()
    h0.record(stream)
    for _ in range(3):
        d_toast.copy_(pin_t, non_blocking=True)
    h1.record(stream)
    torch.cuda.synchronize()
    h2d_gbs = 3 * seg.toast.size / (h0.elapsed_time(h1) / 1000.0) / 1e9

    # ---- reduce over ranks: max time, min rate, summed rows ------------------------------------------
    def mean(x):
        return float(np.mean(x)) if len(x) else 0.0
    t = torch.tensor([dev_ms, e2e_ms, mean(gather_ms), mean(stage_ms), -h2d_gbs], device="cuda", dtype=torch.float64)
    rows_total = torch.tensor([meta["rows"]], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(rows_total, op=dist.ReduceOp.SUM)
    dev_ms_max, e2e_ms_max, gather_ms_max, stage_ms_max, h2d_neg = [float(x) for x in t.cpu()]
    rows_total = float(rows_total.item())
    del d_heap, d_toast, out_ind, out_dep, labels, ph, pt
    ds.close()
    torch.cuda.empty_cache()
    if rank != 0:
        return None

    peak = ctx.peak
    # algorithmic bytes, SURVEY.md §8(d): every input page once + every output byte once
    ind_out = meta["ind_bytes"] // (2 if bf16 else 1)
    step_bytes = seg.heap.size + seg.toast.size + ind_out + meta["dep_bytes"] + 8 * meta["rows"]
    # the gather kernel: reads the TOAST file once; writes the values it unpacks itself (uncompressed and literal-only
    # ones, into the output) and the compressed payloads it stages for the pglz stage
    gather_bytes = seg.toast.size + meta["gather_out_bytes"] // (2 if bf16 else 1) + max(0, meta["pglz_in_bytes"])
    # the pglz stage: reads the compressed streams once, writes their decoded bytes once
    stage_bytes = meta["pglz_in_bytes"] + meta["pglz_out_bytes"]
    if meta["pglz_out_bytes"] == 0:
        dom, dom_ms, dom_bytes = "k_toast_gather", gather_ms_max, gather_bytes
    elif went_wide:
        dom, dom_ms, dom_bytes = ("wide pglz pipeline (k_wide_* : 16 kernels incl. k_wide_blocks + k_wide_finish, + k_pglz_decode for the small "
                                  "values; one CUDA-event pair around the stage)"), stage_ms_max, stage_bytes
        blocks_ms = float(np.mean(b1_ms) + np.mean(b3_ms)) if len(b1_ms) and len(b3_ms) else None
    else:
        dom, dom_ms, dom_bytes = "k_pglz_decode (+ the classification kernel in front of it, one CUDA-event pair around both)", stage_ms_max, stage_bytes
    achieved = dom_bytes / (dom_ms / 1000.0) / 1e9 if dom_ms > 0 else None
    res = {
        "value": rows_total / (dev_ms_max / 1000.0), "unit": "rows/s", "ms_per_step": dev_ms_max, "steps": steps, "warmup": warmup,
        "gb_per_s": world * step_bytes / (dev_ms_max / 1000.0) / 1e9,
        "hbm_frac_whole_step": step_bytes / (dev_ms_max / 1000.0) / 1e9 / peak,
        "config": workload_config(name, seg, meta, world),
        "launch": ("the %d timed steps replayed as one CUDA graph" % steps) if graph is not None else "eager launches",
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": (achieved / peak) if achieved else None, "peak_source": ctx.peak_src,
                     "frac_of_nominal_8_tb_s": (achieved / 8000.0) if achieved else None,
                     "algorithmic_bytes_per_launch": dom_bytes, "launch_ms": dom_ms,
                     "share_of_step": dom_ms / dev_ms_max if dev_ms_max else None,
                     "traffic": (load_profile_traffic(name) or {}).get("dram_bytes_per_launch"),
                     "stages_ms": {"gather": gather_ms_max, "pglz_stage": stage_ms_max, "k_wide_blocks_and_finish (rank 0)": blocks_ms}},
        "e2e": {"value": rows_total / (e2e_ms_max / 1000.0), "unit": "rows/s", "ms_per_step": e2e_ms_max,
                "h2d_bytes_per_step": int(seg.heap.size + seg.toast.size),
                "d2h_bytes_per_step": int(meta["samples"] * 8 + 96), "steps": e2e_steps,
                "api": "cds_decode_host (pinned host pages -> device tensors) + labels/status read back"},
        "h2d": {"gb_per_s": -h2d_neg, "note": "cudaMemcpyAsync from pinned memory alone, all ranks at once (slowest rank); PCIe Gen5 x16 ~63 GB/s theoretical"},
        "gpu_launches": launches,
        "checked": "64-bit checksums of both decoded buffers + labels against the generator's arrays: before the warm-up, after the timed "
                   "region and after the end-to-end steps",
        "_wall": (t_wall0, t_wall2),
    }
    if with_cpu and world == 1:
        ctx.restore_affinity()                                   # the CPU arm gets every core this process was given
        threads = host_threads()
        pool = PythonPortPool(name, args.toast_mib, threads)
        pool.step(1)                                         # warm the workers up
        dt = min(pool.step(1), pool.step(1))
        pool.close()
        c_rate, c_dt, c_threads = cpu_decode_rate(seg, meta, reps=2 if name == "imagenet_packed" else 3)
        res["cpu_baseline"] = {"value": pool.rows / dt, "unit": "rows/s", "cores": pool.procs, "kind": "port",
                               "sample": pool.describe(1) + " (best of 2, %.0f ms)" % (dt * 1000.0),
                               "c_port_openmp_rows_per_s": c_rate,
                               "c_port_note": "oracle/da_oracle.c, whole path in C + OpenMP (%d threads, %.0f ms per segment): much faster "
                                              "than the reference's Python reader; kept beside it for transparency" % (c_threads, c_dt * 1000.0)}
    return res


def mop_leg(ctx):
    """BASELINE config 5 / SURVEY.md §8 f-1 on the ranks of this run: ResNet-50-sized flat float32 states hop round
    the ring (ctq.py:448-508's rule with equal sub-epochs), every rank leaves its mark on the model it hosts, and
    after each epoch every element of every state must hold the analytically expected value."""
    from cerebro_ds_kpgi_b200 import mop
    torch, dist = ctx.torch, ctx.dist
    rank, world = ctx.rank, ctx.world
    numel = 25_557_032                                   # ResNet-50 parameters (madlib_keras_wrapper.py:119-131 layout: one flat f32 vector)
    hop = mop.ModelHopper(rank, world, numel, device="cuda", transport="auto")
    hop.add_local_model(rank, torch.full((numel,), float(rank), device="cuda"))
    marks = float(sum(range(1, world + 1)))
    hop_ms, ok = [], True
    epochs = 4
    for epoch in range(epochs):                            # epoch 0 warms the transport up
        for r in range(len(hop.schedule)):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ctx.barrier()
            e0.record()
            m = hop.hop_to(r)
            e1.record()
            hop.states[m].add_(float(rank + 1))           # the sub-epoch's stand-in: this partition's mark
            torch.cuda.synchronize()
            if epoch > 0 and (r > 0 or epoch > 1):          # (round 0 of an epoch only moves when the ring wraps)
                hop_ms.append(e0.elapsed_time(e1))
        m = hop.model_for_round(len(hop.schedule) - 1)
        want = float(m) + (epoch + 1) * marks
        ok = ok and bool((hop.states[m] == want).all().item())
    moved = [x for x in hop_ms if x > 0]
    t = torch.tensor([float(np.mean(moved)) if moved else 0.0, 0.0 if ok else 1.0], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, bad = float(t[0].item()), float(t[1].item())
    gbs = 4 * numel / (ms / 1000.0) / 1e9 if ms > 0 else None
    return {"state_bytes": 4 * numel, "schedule": "ring, %d rounds per epoch, %d epochs (the first untimed)" % (len(hop.schedule), epochs),
            "hop_ms": ms, "gb_per_s_per_gpu": gbs, "frac_of_nvlink": (gbs / 900.0) if gbs else None,
            "nvlink_peak_gb_per_s_per_direction": 900.0, "epoch_checksum_ok": bad == 0.0,
            "transport": hop.transport + (" (libcds_da cds_peer_copy into the destination's symmetric-memory inbox + barrier)" if hop.transport == "symm"
                                          else " (torch.distributed batch_isend_irecv over NCCL)"),
            "transport_note": hop.transport_note,
            "timing": "CUDA events around hop_to() on every rank after a barrier, mean over the timed hops, max over ranks"}


def run_ours(args):
    ctx = GpuContext(args)
    rank, world = ctx.rank, ctx.world
    sampler = ClockSampler(ctx.local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    head = gpu_leg(ctx, args.workload, args.steps, args.warmup, args.graph, with_cpu=not args.no_cpu)
    configs = {}
    if not args.only:
        for w in extra_workloads(args):
            sub_steps = max(3, min(args.steps, 5)) if w != "imagenet_packed" else 3
            sub = gpu_leg(ctx, w, sub_steps, 3, args.graph and w != "imagenet_packed", with_cpu=not args.no_cpu)
            if sub is not None:
                sub.pop("_wall", None)
                configs[w] = sub
    mop = None
    if world > 1 and not args.only:
        try:
            mop = mop_leg(ctx)
        except Exception as e:  # noqa
            mop = {"error": str(e).splitlines()[0][:300]}
    if rank == 0:
        time.sleep(0.2)
        sampler.stop()
        t_wall0, t_wall2 = head.pop("_wall")
        # nvidia-smi samples every 100 ms: the window spans both timed regions of the headline workload
        clocks = sampler.summary(t_wall0, t_wall2)
        clocks["window"] = "device-resident + e2e timed regions of the headline workload (%.0f ms)" % (1000.0 * (t_wall2 - t_wall0))
        bf16 = args.out_dtype == "bf16"
        line = {
            "metric": "decoded_rows_per_s", "value": head["value"], "unit": "rows/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8" if not bf16 else "u8->bf16", "data": "synthetic",
        }
        for k in ("gb_per_s", "hbm_frac_whole_step", "config", "launch", "roofline", "e2e", "h2d", "gpu_launches", "checked"):
            line[k] = head[k]
        line["clocks"] = clocks
        line["host"] = {"numa_bind_rank0": ctx.numa, "cores_visible": host_threads()}
        if "cpu_baseline" in head:
            head["cpu_baseline"]["reference_python_pglz_mb_per_s"] = python_pglz_rate()
            line["cpu_baseline"] = head["cpu_baseline"]
        if configs:
            line["configs"] = configs
        if mop is not None:
            line["mop"] = mop
        emit(line)
    if world > 1:
        ctx.dist.destroy_process_group()


def protect_stdout():
    """Libraries (NCCL's version banner, for one) write to fd 1.  The contract is ONE JSON line on stdout, so fd 1 is
    pointed at stderr for the whole run and the JSON line goes to the saved descriptor at the end."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


REAL_STDOUT = None


def emit(line):
    REAL_STDOUT.write(json.dumps(line) + "\n")
    REAL_STDOUT.flush()


def main():
    global REAL_STDOUT
    REAL_STDOUT = protect_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="imagenet_raw", choices=ALL_WORKLOADS, help="the headline workload of the JSON line")
    ap.add_argument("--only", action="store_true", help="run the headline workload alone (no `configs`, no `mop`)")
    ap.add_argument("--graph", dest="graph", action="store_true", default=True,
                    help="time the K steps as one CUDA graph replay (default; falls back to eager launches if capture fails)")
    ap.add_argument("--no-graph", dest="graph", action="store_false")
    ap.add_argument("--wide", default="auto", choices=["off", "auto", "force", "host"],
                    help="pglz stage: warp decoder only / automatic (default) / every large value through the wide decoder / "
                         "wide kernels always enqueued, cost model on")
    ap.add_argument("--criteo-buffers", type=int, default=0,
                    help="criteo: buffers per segment (default: from --toast-mib, at most 1024; a real partition holds ~6300)")
    ap.add_argument("--images-per-buffer", type=int, default=3210, help="imagenet_packed: images per packed buffer (the paper's tables: 3210)")
    ap.add_argument("--toast-mib", type=int, default=1024, help="TOAST file size per segment (config: 1 GiB)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--wide-store-limit", type=int, default=-1,
                    help="cds_set_wide_store_limit(bytes) before the run (0: the wide decoder always parses twice); default: the library's")
    ap.add_argument("--out-dtype", default="f32", choices=["f32", "bf16"],
                    help="independent_var output: f32 = the reference's bits (default); bf16 = fused conversion (extra, not the metric)")
    args = ap.parse_args()
    global IMAGES_PER_BUFFER
    IMAGES_PER_BUFFER = args.images_per_buffer
    os.environ["CDS_BENCH_IMAGES_PER_BUFFER"] = str(args.images_per_buffer)
    global CRITEO_BUFFERS
    CRITEO_BUFFERS = args.criteo_buffers
    os.environ["CDS_BENCH_CRITEO_BUFFERS"] = str(args.criteo_buffers)
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
