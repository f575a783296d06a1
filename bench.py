#!/usr/bin/env python
"""bench.py — Direct-Access decode throughput on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one pass of the hot path over one synthetic segment per GPU (heap scan, TOAST walk,
plan, gather, pglz, int64 labels).  Workloads (BASELINE.json `configs`):

    imagenet_raw   configs[1]: ImageNet-shaped rows, uncompressed external bytea  (default; the
                   configuration the metric is quoted on)
    imagenet_pglz  configs[2]: mixed pglz-compressed / incompressible rows
    criteo         configs[3]: Criteo-shaped buffers (256 x 7306 f32, pglz) + inline labels
    imagenet_packed  SURVEY.md 8 f-3: the paper's packed buffers (--images-per-buffer B, default 3210 = 483 MB per
                   value, pglz): decoded by the intra-value parallel decoder

One JSON line on stdout (rank 0).  `value` = rows/s over all ranks with pages resident in HBM;
`e2e` = the same through cds_decode_host with pinned HOST page buffers (H2D inside the timed
region, labels + status read back); `roofline` = the gather kernel against measured HBM copy
bandwidth; `cpu_baseline` = the CPU restatement (oracle/) on this host's cores.
"""
import argparse
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
NVSMI_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
               "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
               "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------------
# workloads
# ------------------------------------------------------------------------------------------------

def make_workload(name, seed, toast_mib):
    """Synthetic segment per SURVEY.md §8(d).  Returns (segment, meta)."""
    from cerebro_ds_kpgi_b200 import synth
    target = toast_mib << 20
    if name == "imagenet_raw":
        n = synth.rows_for_toast_bytes(target)
        ind, dep = synth.imagenet_rows(n, seed=seed, kind="norm_u8")
        seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="ext_raw", dist_key=seed - 2018,
                                  keep_expected=False)
        meta = dict(isb=IMAGENET_ROW_BYTES, dsb=LABEL_ROW_BYTES, samples=n, classes=1000,
                    desc="ImageNet-shaped rows (112x112x3 f32 + int16[1000] one-hot), uncompressed external bytea")
    elif name == "imagenet_pglz":
        n = int(synth.rows_for_toast_bytes(target) * 1.15)
        a_ind, a_dep = synth.imagenet_rows(n, seed=seed, kind="norm_u8")
        rng = np.random.default_rng(seed + 1000)
        noise = np.random.default_rng(seed + 2000)
        ind, st = [], []
        for i in range(n):
            if rng.random() < 0.5:
                ind.append(a_ind[i]); st.append("ext_pglz")
            else:
                ind.append(noise.standard_normal(a_ind[i].shape, dtype=np.float32)); st.append("ext_literal")
        seg = synth.build_segment(ind, a_dep, ind_storage=st, dep_storage="inline_pglz", dist_key=seed - 2018,
                                  keep_expected=False, effort=4)
        meta = dict(isb=IMAGENET_ROW_BYTES, dsb=LABEL_ROW_BYTES, samples=n, classes=1000,
                    desc="ImageNet-shaped rows, 50% pglz-compressed / 50% literal-only pglz, inline-compressed labels")
    elif name == "criteo":
        B = 256
        per_buf = B * 7306 * 4
        n = max(1, (target * 20) // per_buf)       # ~20x compression: size the TOAST file, not the output
        n = min(n, 1024)           # a real partition holds ~6 300 such buffers (1 624 157 rows / 256)
        if CRITEO_BUFFERS:
            n = CRITEO_BUFFERS     # --criteo-buffers: a partition-sized segment; distinct buffers repeat after 256
        ind, dep = synth.criteo_buffers(min(n, 256), seed=seed, rows_per_buffer=B)
        ind, dep = [ind[i % len(ind)] for i in range(n)], [dep[i % len(dep)] for i in range(n)]
        seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="inline_pglz", dist_key=seed - 2018,
                                  keep_expected=False, effort=4, toast_cap_pages=n * 16 + 64)
        meta = dict(isb=7306 * 4, dsb=4, samples=n * B, classes=2,
                    desc="Criteo-shaped buffers (256 x 7306 f32 pglz, int16[256,2] inline pglz)")
    elif name == "imagenet_packed":
        # the paper's table layout (load_imagenet.py:27-31): B images per buffer, one huge pglz value each
        B = IMAGES_PER_BUFFER
        n = max(1, int(target // (B * IMAGENET_ROW_BYTES * 0.67)))
        ind, dep = synth.imagenet_rows(min(n, 4), seed=seed, kind="norm_u8", images_per_row=B)     # distinct buffers repeat after 4
        ind, dep = [ind[i % len(ind)] for i in range(n)], [dep[i % len(dep)] for i in range(n)]
        seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="ext_pglz", dist_key=seed - 2018,
                                  keep_expected=False, effort=4)
        meta = dict(isb=IMAGENET_ROW_BYTES, dsb=LABEL_ROW_BYTES, samples=n * B, classes=1000,
                    desc="ImageNet-shaped packed buffers: %d images per buffer (%d MB per value), pglz-compressed" % (B, B * IMAGENET_ROW_BYTES // 10 ** 6))
    else:
        raise SystemExit("unknown workload " + name)
    meta["ind_bytes"] = meta["samples"] * meta["isb"]
    meta["dep_bytes"] = meta["samples"] * meta["dsb"]
    meta["rows"] = meta["samples"]
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
# CPU arm: the oracle's C restatement on the host cores
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


def host_threads():
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
_REF_PART = None          # (heap, toast, buffers) of THIS worker process


def _ref_pglz():
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import da_oracle
    try:
        return da_oracle.RefPglz().pglz_decompress, "oracle/_ref/pg_lzcompress.so (the reference's own C decoder)"
    except RuntimeError:
        return da_oracle.COracle().pglz_decompress, "oracle/da_oracle.c pglz (oracle/_ref not built on this machine)"


def _ref_init(counter, workload, part_mib):
    """Worker start-up (spawned, so nothing CUDA-related is inherited): build this worker's sub-segment."""
    global _REF_PART
    with counter.get_lock():
        idx = counter.value
        counter.value += 1
    seg, meta = make_workload(workload, 2018 + 100 * idx, part_mib)
    _REF_PART = (seg.heap, seg.toast, len(seg.rows), _ref_pglz()[0])


def _ref_worker(_):
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import da_oracle
    heap, toast, buffers, pglz = _REF_PART
    out = da_oracle.input_fn_bytes(heap, toast, None, pglz=pglz)
    assert len(out) == buffers
    return buffers


class PythonPortPool:
    """oracle/da_oracle.py (pure-Python page walk / detoast, the reference's loop structure) over all cores:
    `procs` spawned workers, each holding one sub-segment of 1/procs of the workload's rows."""

    def __init__(self, workload, toast_mib, procs):
        import multiprocessing as mp
        ctx = mp.get_context("spawn")
        self.procs = procs
        part_mib = max(1, toast_mib // procs)
        seg, meta = make_workload(workload, 2018, part_mib)
        self.rows, self.buffers = procs * meta["rows"], procs * len(seg.rows)
        self.pool = ctx.Pool(procs, initializer=_ref_init, initargs=(ctx.Value("i", 0), workload, part_mib))
        self.pglz_note = _ref_pglz()[1]
        self.step(1)                                              # every worker up and its sub-segment built

    def step(self, repeats=1):
        t0 = time.perf_counter()
        for _ in range(repeats):
            got = sum(self.pool.map(_ref_worker, range(self.procs), chunksize=1))
            assert got == self.buffers, (got, self.buffers)
        return time.perf_counter() - t0

    def one_process(self):
        """Seconds one worker alone needs for its own sub-segment (the reference is a single process per partition)."""
        t0 = time.perf_counter()
        got = sum(self.pool.map(_ref_worker, range(1), chunksize=1))
        assert got * self.procs == self.buffers
        return time.perf_counter() - t0

    def close(self):
        self.pool.close()
        self.pool.join()

    def describe(self, repeats):
        return ("%d x the %d-row workload per step as %d sub-segments, one process each: oracle/da_oracle.py (Python restatement "
                "of pg_page_reader.py's page walk / detoast) with pglz through %s" % (repeats, self.rows, self.procs, self.pglz_note))


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    n_seg = max(1, args.gpus)          # the whole job: one segment per GPU of our arm, decoded on this host's cores
    pool = PythonPortPool(args.workload, args.toast_mib, threads)
    times = []
    for i in range(args.warmup + args.steps):
        dt = pool.step(n_seg)
        if i >= args.warmup:
            times.append(dt)
    pool.close()
    ms = 1000.0 * float(np.mean(times))
    value = n_seg * pool.rows / (ms / 1000.0)
    # the much faster all-C restatement, for transparency (not what the reference ships)
    seg, meta = make_workload(args.workload, 2018, args.toast_mib)
    c_rate, c_dt, c_threads = cpu_decode_rate(seg, meta, 3)
    alg_bytes = seg.heap.size + seg.toast.size + meta["ind_bytes"] + meta["dep_bytes"] + 8 * meta["rows"]
    line = {
        "impl": "reference", "metric": "decoded_rows_per_s", "value": value, "unit": "rows/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "gb_per_s": value * alg_bytes / meta["rows"] / 1e9,
        "config": {"workload": args.workload, "description": meta["desc"], "rows_per_segment": pool.rows,
                   "toast_mib": seg.toast.size >> 20, "segments": n_seg},
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": threads, "kind": "port", "sample": pool.describe(n_seg)},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "host_cores": threads,
        "c_port_openmp": {"value": c_rate, "unit": "rows/s", "cores": c_threads, "ms_per_segment": 1000.0 * c_dt,
                          "note": "oracle/da_oracle.c, the whole path in C with OpenMP over values: far faster than anything the "
                                  "reference ships (its reader is a Python loop); reported so that nothing is hidden"},
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------

def load_profile_traffic():
    """dram bytes per gather launch from the committed ncu summary (profiles/), or None."""
    p = os.path.join(REPO, "profiles", "gather_traffic.json")
    try:
        with open(p) as f:
            return json.load(f)
    except Exception:  # noqa
        return None


def run_ours(args):
    import torch
    import torch.distributed as dist
    from cerebro_ds_kpgi_b200 import _lib, segment

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        log("note: WORLD_SIZE=%d, --gpus=%d" % (world, args.gpus))
    L = _lib.require_device()
    torch.cuda.set_device(local)
    from cerebro_ds_kpgi_b200 import hostnuma
    numa = hostnuma.bind_to_gpu_node(local)          # before any staging buffer is allocated or touched
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t_gen = time.time()
    seg, meta = make_workload(args.workload, 2018 + rank, args.toast_mib)
    heap_pages, toast_pages = seg.heap.size // 32768, seg.toast.size // 32768
    log("[rank %d] segment: %d rows, heap %d pages, TOAST %d pages (%.1f MiB), generated in %.1fs" % (
        rank, meta["rows"], heap_pages, toast_pages, seg.toast.size / 2 ** 20, time.time() - t_gen))

    ds = segment.DeviceSegment(heap_pages, toast_pages, device=local, stage=True)
    bf16 = args.out_dtype == "bf16"
    ds.set_output_format(bf16)
    _lib.check(L.cds_set_wide_pglz(ds._h, {"off": 0, "force": 1, "auto": 2}[args.wide]))
    d_heap = torch.from_numpy(seg.heap).cuda()
    d_toast = torch.from_numpy(seg.toast).cuda()
    out_ind = torch.empty(meta["ind_bytes"] // (2 if bf16 else 1), dtype=torch.uint8, device="cuda")
    out_dep = torch.empty(meta["dep_bytes"], dtype=torch.uint8, device="cuda")
    onehot = out_dep.view(torch.int16).view(-1, meta["classes"])
    labels = torch.empty(meta["samples"], dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream()
    sp = segment._stream_ptr(stream)

    def step_device():
        ds.decode_device(d_heap, d_toast, out_ind, out_dep, meta["isb"], meta["dsb"], sync=False)
        _lib.check(L.cds_onehot_argmax_i16(segment._ptr(onehot), meta["samples"], meta["classes"], segment._ptr(labels), sp))

    # ---- correctness gate before timing: counts + a checksum against the generator's arrays
    step_device()
    s = ds.finish()
    if s.status != 0 or s.ind_bytes != meta["ind_bytes"] or s.dep_bytes != meta["dep_bytes"]:
        raise SystemExit("decode failed: status %d, %d/%d bytes" % (s.status, s.ind_bytes, meta["ind_bytes"]))
    lab_ref = np.argmax(out_dep.cpu().numpy().view(np.int16).reshape(-1, meta["classes"]), axis=1)
    assert np.array_equal(labels.cpu().numpy(), lab_ref), "labels differ"

    # ---- device-resident timing (value) ---------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    ds.finish()
    ds.timing_enable(args.steps)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    barrier()
    launches0 = L.cds_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    graph = None
    if args.graph:
        # The K steps of the timed region as ONE CUDA graph (the same kernels and memsets; the gather's timing events become
        # event-record nodes): replay removes the host launch gaps between the short front-end kernels.
        try:
            graph = torch.cuda.CUDAGraph()
            cap_stream = torch.cuda.Stream()
            cap_stream.wait_stream(stream)
            with torch.cuda.graph(graph, stream=cap_stream):
                cur = torch.cuda.current_stream()
                for _ in range(args.steps):
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
            ds.timing_enable(args.steps)
            launches0 = L.cds_launch_count()
    barrier()
    t_wall0 = time.time()
    ev0.record(stream)
    if graph is not None:
        graph.replay()
    else:
        for _ in range(args.steps):
            step_device()
    ev1.record(stream)
    torch.cuda.synchronize()
    t_wall1 = time.time()
    barrier()
    if graph is None:
        launches = int(L.cds_launch_count() - launches0)
    s = ds.finish()
    assert s.status == 0
    dev_ms = ev0.elapsed_time(ev1) / args.steps
    gather_ms = ds.timing_read(args.steps)
    ds.timing_enable(0)

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

    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        step_e2e()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(e2e_steps):
        st = step_e2e()
        assert st.status == 0
    e1.record(stream)
    torch.cuda.synchronize()
    barrier()
    e2e_ms = e0.elapsed_time(e1) / e2e_steps
    t_wall2 = time.time()
    assert np.array_equal(labels_host.numpy(), lab_ref)

    # H2D alone, against PCIe
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pin_t = torch.from_numpy(pt.array)
    d_toast.copy_(pin_t, non_blocking=True)
    torch.cuda.synchronize()
    h0.record(stream)
    for _ in range(3):
        d_toast.copy_(pin_t, non_blocking=True)
    h1.record(stream)
    torch.cuda.synchronize()
    h2d_gbs = 3 * seg.toast.size / (h0.elapsed_time(h1) / 1000.0) / 1e9
    if rank == 0:
        time.sleep(0.2)
        sampler.stop()

    # ---- reduce over ranks: max time ---------------------------------------------------------------
    t = torch.tensor([dev_ms, e2e_ms, float(np.mean(gather_ms)) if gather_ms.size else 0.0], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms_max, e2e_ms_max, gather_ms_max = [float(x) for x in t.cpu()]
    rows_total = torch.tensor([meta["rows"]], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(rows_total, op=dist.ReduceOp.SUM)
    rows_total = float(rows_total.item())

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:  # noqa
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s (B200_PROFILING.md)"
        # algorithmic bytes, SURVEY.md §8(d): every input page once + every output byte once
        ind_out = meta["ind_bytes"] // (2 if bf16 else 1)
        step_bytes = seg.heap.size + seg.toast.size + ind_out + meta["dep_bytes"] + 8 * meta["rows"]
        # the gather kernel: reads the TOAST file once, writes every byte that is stored externally
        gather_out = ind_out + (meta["dep_bytes"] if args.workload == "imagenet_raw" else 0)
        if args.workload != "imagenet_raw":
            gather_out = int(seg.stored_bytes)          # compressed payloads go to the staging area
        gather_bytes = seg.toast.size + gather_out
        achieved = gather_bytes / (gather_ms_max / 1000.0) / 1e9 if gather_ms_max > 0 else None
        traffic = load_profile_traffic()
        # nvidia-smi samples every 100 ms: the window spans both timed regions (device-resident steps, then e2e steps)
        clocks = sampler.summary(t_wall0, t_wall2)
        clocks["window"] = "device-resident + e2e timed regions (%.0f ms)" % (1000.0 * (t_wall2 - t_wall0))
        # CPU baseline on this host (rank 0 only, N=1 only)
        cpu = None
        if world == 1 and not args.no_cpu:
            threads = host_threads()
            pool = PythonPortPool(args.workload, args.toast_mib, threads)
            pool.step(1)                                         # warm the workers up
            dt = min(pool.step(1), pool.step(1))
            dt_one = min(pool.one_process(), pool.one_process())
            pool.close()
            c_rate, c_dt, c_threads = cpu_decode_rate(seg, meta, reps=3)
            cpu = {"value": pool.rows / dt, "unit": "rows/s", "cores": threads, "kind": "port",
                   "sample": pool.describe(1) + " (best of 2, %.0f ms)" % (dt * 1000.0),
                   "one_process_rows_per_s": (pool.rows / threads) / dt_one,
                   "c_port_openmp_rows_per_s": c_rate,
                   "reference_python_pglz_mb_per_s": python_pglz_rate(),
                   "c_port_note": "oracle/da_oracle.c, whole path in C + OpenMP (%d threads, %.0f ms per segment): much faster "
                                  "than the reference's Python reader; kept beside it for transparency" % (c_threads, c_dt * 1000.0)}
        line = {
            "metric": "decoded_rows_per_s", "value": rows_total / (dev_ms_max / 1000.0), "unit": "rows/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms_max,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8" if not bf16 else "u8->bf16", "data": "synthetic",
            "gb_per_s": world * step_bytes / (dev_ms_max / 1000.0) / 1e9,
            "hbm_frac_whole_step": step_bytes / (dev_ms_max / 1000.0) / 1e9 / peak,
            "config": {"workload": args.workload, "description": meta["desc"], "rows_per_segment": meta["rows"],
                       "toast_mib": seg.toast.size >> 20, "heap_pages": heap_pages, "segments": world,
                       "partitioning": "one segment per GPU, no collective on the decode path",
                       "launch": ("the %d timed steps replayed as one CUDA graph" % args.steps) if graph is not None else "eager launches",
                       "l2": "inputs (%d MiB per step) larger than the 126 MB L2; no flush needed" % ((seg.toast.size + meta["ind_bytes"]) >> 20),
                       "algorithmic_bytes_per_row": step_bytes / meta["rows"]},
            "roofline": {"bound": "hbm", "kernel": "k_toast_gather", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": (achieved / peak) if achieved else None, "peak_source": peak_src,
                         "frac_of_nominal_8_tb_s": (achieved / 8000.0) if achieved else None,
                         "algorithmic_bytes_per_launch": gather_bytes, "launch_ms": gather_ms_max,
                         "share_of_step": gather_ms_max / dev_ms_max if dev_ms_max else None,
                         "traffic": (traffic or {}).get("dram_bytes_per_launch")},
            "e2e": {"value": rows_total / (e2e_ms_max / 1000.0), "unit": "rows/s", "ms_per_step": e2e_ms_max,
                    "h2d_bytes_per_step": int(seg.heap.size + seg.toast.size),
                    "d2h_bytes_per_step": int(meta["samples"] * 8 + 96), "steps": e2e_steps,
                    "api": "cds_decode_host (pinned host pages -> device tensors) + labels/status read back"},
            "h2d": {"gb_per_s": h2d_gbs, "note": "cudaMemcpyAsync from pinned memory alone; PCIe Gen5 x16 ~63 GB/s theoretical"},
            "gpu_launches": launches,
            "clocks": clocks,
            "host": {"numa_bind_rank0": numa, "cores_visible": host_threads()},
        }
        if cpu:
            line["cpu_baseline"] = cpu
        emit(line)
    if world > 1:
        dist.destroy_process_group()


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
    ap.add_argument("--workload", default="imagenet_raw", choices=["imagenet_raw", "imagenet_pglz", "criteo", "imagenet_packed"])
    ap.add_argument("--graph", dest="graph", action="store_true", default=True,
                    help="time the K steps as one CUDA graph replay (default; falls back to eager launches if capture fails)")
    ap.add_argument("--no-graph", dest="graph", action="store_false")
    ap.add_argument("--wide", default="auto", choices=["off", "auto", "force"],
                    help="intra-value parallel pglz decoder: never / decided per decode by the device-side cost model / always")
    ap.add_argument("--criteo-buffers", type=int, default=0,
                    help="criteo: buffers per segment (default: from --toast-mib, at most 1024; a real partition holds ~6300)")
    ap.add_argument("--images-per-buffer", type=int, default=3210, help="imagenet_packed: images per packed buffer (the paper's tables: 3210)")
    ap.add_argument("--toast-mib", type=int, default=1024, help="TOAST file size per segment (config: 1 GiB)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
