#!/usr/bin/env python
"""Text ETL ("next" row f-4) throughput on one GPU: ImageNet-shaped gpfdist TEXT unload lines
(id|{{{...}}}|label, 37632 values per row in PostgreSQL's float4 text form) -> float32 [rows,112,112,3].
Prints one JSON line; the CPU figure is the reference's own parser restated (oracle/da_oracle.py::text_parse,
ast.literal_eval + np.asarray) on a few rows, one core."""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main():
    from cerebro_ds_kpgi_b200 import _lib, segment
    import da_oracle
    L = _lib.require_device()
    rng = np.random.default_rng(2018)
    V = 112 * 112 * 3
    base_rows = 48
    lines = []
    for r in range(base_rows):
        vals = rng.standard_normal(V).astype(np.float32)
        toks = ["%.6g" % v for v in vals]
        pix = ["{" + ",".join(toks[3 * p:3 * p + 3]) + "}" for p in range(112 * 112)]
        rows = ["{" + ",".join(pix[112 * y:112 * y + 112]) + "}" for y in range(112)]
        lines.append("%d|{%s}|%d\n" % (r, ",".join(rows), r % 1000))
    chunk = "".join(lines).encode()
    reps = 24
    text = np.frombuffer(chunk * reps, dtype=np.uint8)
    rows = base_rows * reps
    d_text = torch.from_numpy(text.copy()).cuda()
    feats = torch.empty((rows, V), dtype=torch.float32, device="cuda")
    labels = torch.empty(rows, dtype=torch.int64, device="cuda")
    ids = torch.empty(rows, dtype=torch.int64, device="cuda")
    n_rows, err = ctypes.c_uint64(), ctypes.c_uint64()

    def run():
        rc = L.cds_text_etl(0, segment._ptr(d_text), text.size, V, 0, segment._ptr(feats), segment._ptr(labels), segment._ptr(ids),
                            rows, ctypes.byref(n_rows), ctypes.byref(err), segment._stream_ptr())
        assert rc == 0 and n_rows.value == rows, (rc, n_rows.value, err.value)
    for _ in range(3):
        run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    # parity spot check against the reference parser restated, on the first rows
    t0 = time.perf_counter()
    want = da_oracle.text_parse(chunk[:len(lines[0]) + len(lines[1])])
    cpu_s = time.perf_counter() - t0
    got = feats[:2].cpu().numpy()
    assert got.tobytes() == want[1].reshape(2, -1).tobytes()
    peak = 6540.8
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:  # noqa
        pass
    alg = 2 * text.size + rows * V * 4             # two passes over the text (count, parse) + the output
    print(json.dumps({"workload": "text_etl_imagenet", "text_mib": text.size >> 20, "rows": rows, "ms": ms,
                      "rows_per_s": rows / (ms / 1000), "text_gb_per_s": text.size / ms / 1e6,
                      "hbm_gb_per_s": alg / ms / 1e6, "frac_of_hbm_peak": alg / ms / 1e6 / peak,
                      "cpu_reference_parser_rows_per_s_1core": 2 / cpu_s,
                      "note": "ms includes the call's own cudaMalloc/cudaFree of its tile arrays and the final sync"}))


if __name__ == "__main__":
    main()
