#!/usr/bin/env python
"""Consumer-side kernels against the HBM roofline, on one GPU:
batch assembly (index gather + NHWC->NCHW + label lookup), whole-partition NHWC->NCHW, one-hot argmax,
float32->bfloat16.  Prints one JSON line.  Inputs are larger than L2 or rotated so that reads miss."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    from cerebro_ds_kpgi_b200 import consumers, da, segment
    peak = 6540.8
    try:
        peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:  # noqa
        pass
    n = 6971
    images = torch.randn(n, 112, 112, 3, device="cuda")
    onehot = torch.zeros(n, 1000, dtype=torch.int16, device="cuda")
    onehot[torch.arange(n), torch.randint(0, 1000, (n,))] = 1
    labels = segment.onehot_argmax(onehot)
    part = da.DevicePartition(images, onehot, labels, list(range(n)), [1] * n)
    ds = consumers.DeviceDADataset(part, length=160160, train=True)
    out = {}
    for B in (256, 1024):
        # different random indices every call: a batch (38 / 154 MB) never sits in L2 from the call before
        idxs = [torch.randint(0, 160160, (B,), device="cuda") for _ in range(8)]
        k = [0]

        def f():
            ds.batch(idxs[k[0] % 8]); k[0] += 1
        ms = timed(f)
        by = 2 * B * 150528 + 16 * B
        out["batch_assemble_f32_B%d" % B] = {"ms": ms, "gb_per_s": by / ms / 1e6, "frac_of_hbm_peak": by / ms / 1e6 / peak}

        def g():
            ds.batch(idxs[k[0] % 8], dtype=torch.bfloat16); k[0] += 1
        ms = timed(g)
        by = int(1.5 * B * 150528) + 16 * B
        out["batch_assemble_bf16_B%d" % B] = {"ms": ms, "gb_per_s": by / ms / 1e6, "frac_of_hbm_peak": by / ms / 1e6 / peak}
    ms = timed(lambda: segment.nhwc_to_nchw(images), reps=5)
    out["nhwc_to_nchw_partition"] = {"ms": ms, "gb_per_s": 2 * images.numel() * 4 / ms / 1e6, "frac_of_hbm_peak": 2 * images.numel() * 4 / ms / 1e6 / peak}
    ms = timed(lambda: segment.to_bf16(images), reps=5)
    out["f32_to_bf16_partition"] = {"ms": ms, "gb_per_s": 1.5 * images.numel() * 4 / ms / 1e6, "frac_of_hbm_peak": 1.5 * images.numel() * 4 / ms / 1e6 / peak}
    big = torch.zeros(8 * n, 1000, dtype=torch.int16, device="cuda")
    ms = timed(lambda: segment.onehot_argmax(big), reps=10)
    out["onehot_argmax_%d_rows" % (8 * n)] = {"ms": ms, "gb_per_s": big.numel() * 2 / ms / 1e6, "frac_of_hbm_peak": big.numel() * 2 / ms / 1e6 / peak}
    out["hbm_peak_gb_per_s"] = peak
    print(json.dumps(out))


if __name__ == "__main__":
    main()
