"""Time the intra-value parallel pglz decoder (csrc/pglz_wide.cuh) on one huge value, beside the warp-per-value
decoder and the single-thread C oracle.  SURVEY.md §8 f-3: the paper's packed buffers (B = 3210 images = 483 MB).

    python tools/wide_pglz_bench.py [--mib 461] [--kind image|sparse|literal] [--warp-mib 32] [--reps 5]
"""
import argparse
import ctypes
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))


def make_raw(kind, n, seed=2018):
    rng = np.random.default_rng(seed)
    if kind == "image":
        lut = ((np.arange(256)[None, :] / 255.0 - np.array([0.485, 0.456, 0.406])[:, None]) /
               np.array([0.229, 0.224, 0.225])[:, None]).astype(np.float32)
        k = n // 4 + 1
        return lut[np.arange(k) % 3, rng.integers(0, 256, k)].view(np.uint8)[:n].copy()
    if kind == "sparse":
        x = np.zeros(n // 4 + 1, dtype=np.float32)
        x[rng.integers(0, x.size, x.size // 187)] = 1.0
        return x.view(np.uint8)[:n].copy()
    return rng.integers(0, 256, n, dtype=np.uint8)


def time_batch(L, torch, datum, reps):
    import cerebro_ds_kpgi_b200._lib as _lib
    raw = int(np.frombuffer(datum[4:8], dtype="<u4")[0])
    src_off = np.array([0, len(datum)], dtype=np.uint64)
    dst_off = np.array([0, (raw + 15) & ~15], dtype=np.uint64)
    d_src = torch.from_numpy(np.frombuffer(datum + b"\0" * 16, dtype=np.uint8).copy()).cuda()
    d_dst = torch.empty(int(dst_off[1]) + 16, dtype=torch.uint8, device="cuda")
    status = np.zeros(1, dtype=np.int32)
    times = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _lib.check(L.cds_pglz_decompress_batch(0, d_src.data_ptr(), src_off.ctypes.data, d_dst.data_ptr(), dst_off.ctypes.data, 1,
                                               status.ctypes.data, None))
        times.append(time.perf_counter() - t0)
        assert status[0] == 0
    return min(times), d_dst[:raw]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=float, default=3210 * 150528 / 2 ** 20)
    ap.add_argument("--kind", default="image")
    ap.add_argument("--warp-mib", type=float, default=32.0, help="size of the value timed on the warp decoder (it is slow)")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    import torch
    from cerebro_ds_kpgi_b200 import _lib, synth
    import da_oracle
    L = _lib.require_device()
    n = int(args.mib * 2 ** 20)
    raw = make_raw(args.kind, n)
    t0 = time.perf_counter()
    datum = synth.pglz_varlena(raw, "literal" if args.kind == "literal" else "lz")
    t_enc = time.perf_counter() - t0
    res = {"kind": args.kind, "raw_bytes": n, "stream_bytes": len(datum) - 8, "ratio": round((len(datum) - 8) / n, 4),
           "encode_s": round(t_enc, 2)}

    _lib.check(L.cds_set_wide_pglz_threshold(1 << 20, 0))
    before = L.cds_wide_decoded_count()
    t_wide, out = time_batch(L, torch, datum, args.reps)
    assert L.cds_wide_decoded_count() - before == args.reps
    assert out.cpu().numpy().tobytes() == raw.tobytes()
    res["wide_ms"] = round(t_wide * 1e3, 3)
    res["wide_out_GBps"] = round(n / t_wide / 1e9, 2)

    nw = min(n, int(args.warp_mib * 2 ** 20))
    small = synth.pglz_varlena(raw[:nw], "literal" if args.kind == "literal" else "lz")
    _lib.check(L.cds_set_wide_pglz_threshold(0, 0))
    t_warp, out = time_batch(L, torch, small, 2)
    assert out.cpu().numpy().tobytes() == raw[:nw].tobytes()
    _lib.check(L.cds_set_wide_pglz_threshold(1 << 20, 0))
    res["warp_bytes"] = nw
    res["warp_ms"] = round(t_warp * 1e3, 3)
    res["warp_out_GBps"] = round(nw / t_warp / 1e9, 3)

    co = da_oracle.COracle()
    t0 = time.perf_counter()
    got = co.pglz_decompress(small)
    t_cpu = time.perf_counter() - t0
    assert bytes(got) == raw[:nw].tobytes()
    res["cpu_c_1thread_out_GBps"] = round(nw / t_cpu / 1e9, 3)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
