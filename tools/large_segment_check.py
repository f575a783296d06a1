#!/usr/bin/env python
"""One-off check that byte offsets beyond 4 GiB are handled: decode a ~6 GiB uncompressed segment (and its
bf16 variant) and compare every output byte with the generator's rows.  Not part of the test suite (needs
~20 GB of host RAM and 15 GB of HBM)."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from cerebro_ds_kpgi_b200 import segment, synth
    mib = int(sys.argv[1]) if len(sys.argv) > 1 else 6144
    n = synth.rows_for_toast_bytes(mib << 20)
    t0 = time.time()
    ind, dep = synth.imagenet_rows(n, seed=7, kind="normal")
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="ext_raw", keep_expected=False)
    gen_s = time.time() - t0
    ds = segment.DeviceSegment(seg.heap.size // 32768, seg.toast.size // 32768, scratch_bytes=0)
    h, t = torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda()
    out_i = torch.empty(n * 150528, dtype=torch.uint8, device="cuda")
    out_d = torch.empty(n * 2000, dtype=torch.uint8, device="cuda")
    s = ds.decode_device(h, t, out_i, out_d, 150528, 2000)
    assert s.status == 0 and s.n_rows == n and s.ind_bytes == n * 150528, (s.status, s.n_rows)
    ok = True
    step = 512
    for a in range(0, n, step):
        src = torch.from_numpy(np.concatenate([x.reshape(-1) for x in ind[a:a + step]])).cuda()
        ok &= bool(torch.equal(out_i.view(torch.float32)[a * 37632:(a + len(ind[a:a + step])) * 37632], src))
    lab = segment.onehot_argmax(out_d.view(torch.int16).view(n, 1000)).cpu().numpy()
    ok &= bool(np.array_equal(lab, np.concatenate([np.argmax(d, axis=1) for d in dep])))
    del out_i
    torch.cuda.empty_cache()
    ds.set_output_format(True)
    out_b = torch.empty(n * 150528 // 2, dtype=torch.uint8, device="cuda")
    s = ds.decode_device(h, t, out_b, out_d, 150528, 2000)
    assert s.status == 0
    okb = True
    for a in range(0, n, step):
        src = torch.from_numpy(np.concatenate([x.reshape(-1) for x in ind[a:a + step]])).cuda().to(torch.bfloat16)
        okb &= bool(torch.equal(out_b.view(torch.bfloat16)[a * 37632:(a + len(ind[a:a + step])) * 37632].view(torch.int16), src.view(torch.int16)))
    print(json.dumps({"toast_gib": seg.toast.size / 2 ** 30, "rows": n, "output_gib": n * 150528 / 2 ** 30,
                      "f32_bit_exact": ok, "bf16_bit_exact": okb, "generate_s": gen_s}))
    assert ok and okb


if __name__ == "__main__":
    main()
