#!/usr/bin/env python
"""The call a user of the reference actually makes: da.input_fn(path) on segment FILES, returning the numpy dict.
Times it for a 1 GiB synthetic segment (uncompressed and pglz variants) and prints one JSON line."""
import contextlib
import io
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import bench
    from cerebro_ds_kpgi_b200 import da, synth
    out = {}
    root = tempfile.mkdtemp(prefix="cds_dropin_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    for workload in ("imagenet_raw", "imagenet_pglz"):
        seg, meta = bench.make_workload(workload, 2018, 1024)
        d = seg.write(root, 0, 16384, 50000, 50001)
        shape_rows = [{"__dist_key__": 0, "buffer_id": i, "independent_var_shape": [1, 112, 112, 3],
                       "dependent_var_shape": [1, 1000], "table_name": "t"} for i in range(meta["rows"])]
        cat = os.path.join(root, "sys_cat.dill")
        synth.write_catalog(cat, {"train": [dict(gp_segment_id=0, relname="t", relfilenode=50000, toast_relfilenode=50001)]}, shape_rows)
        da.SYS_CAT_PATH = cat
        times = []
        for _ in range(3):
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(io.StringIO()):
                res = da.input_fn(os.path.join(d, "train"))
            times.append(time.perf_counter() - t0)
        assert len(res) == meta["rows"]
        k = next(iter(res))
        assert res[k]["independent_var"].shape == (1, 112, 112, 3)
        t0 = time.perf_counter()
        with contextlib.redirect_stdout(io.StringIO()):
            part = da.input_fn_device(os.path.join(d, "train"))
        import torch
        torch.cuda.synchronize()
        t_dev = time.perf_counter() - t0
        out[workload] = {"rows": meta["rows"], "input_fn_s": min(times), "rows_per_s": meta["rows"] / min(times),
                         "input_fn_device_s": t_dev, "all_runs_s": times}
        del res, part, seg
    print(json.dumps(out))


if __name__ == "__main__":
    main()
