#!/usr/bin/env python
"""Run the randomised differential GPU tests (tests/test_gpu_fuzz.py, parts of tests/test_gpu_wide_pglz.py and tests/test_gpu_stream.py) over a range of
seed shifts in one process:   python tools/soak_fuzz.py FIRST LAST     (prints the shifts that failed; exit code 1 if any)."""
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main():
    first, last = int(sys.argv[1]), int(sys.argv[2])
    import torch
    import da_oracle
    from cerebro_ds_kpgi_b200 import _lib, build
    build.build_oracle()
    import test_gpu_fuzz as f
    import test_gpu_wide_pglz as w
    import test_gpu_stream as sm
    co = da_oracle.COracle()
    L = _lib.require_device()

    def set_thr(n, force=1):
        _lib.check(L.cds_set_wide_pglz_threshold(int(n), int(force)))

    cases = [("pglz_fuzz", lambda: f.test_pglz_fuzz_structured_streams(torch, co)),
             ("layout", lambda: f.test_segment_layout_fuzz(torch, co, False)),
             ("layout_wide", lambda: f.test_segment_layout_fuzz(torch, co, True)),
             ("corruption", lambda: f.test_corruption_fuzz_never_disagrees_silently(torch, co)),
             ("wide_streams", lambda: w.test_wide_matches_oracle_on_structured_streams(torch, co, set_thr)),
             ("wide_corrupt", lambda: w.test_wide_corrupt_streams_agree_with_oracle(torch, co, set_thr)),
             ("streamed", lambda: sm.test_streamed_layout_fuzz(torch))]

    def two_parses(fn):                                   # the same case without the store of resolved blocks
        def run():
            _lib.check(L.cds_set_wide_store_limit(0))
            try:
                fn()
            finally:
                _lib.check(L.cds_set_wide_store_limit(2 ** 64 - 1))
        return run
    cases += [(name + "_two_parses", two_parses(fn)) for name, fn in cases if name.startswith("wide_") or name == "layout_wide"]
    failed = []
    for k in range(first, last + 1):
        f.SEED_SHIFT = w.SEED_SHIFT = sm.SEED_SHIFT = k
        for name, fn in cases:
            try:
                fn()
            except BaseException:  # noqa
                failed.append((k, name))
                print("seed shift %d, %s:\n%s" % (k, name, traceback.format_exc().splitlines()[-1]), flush=True)
            finally:
                set_thr(1 << 20, 0)
        print("shift %d done" % k, flush=True)
    print("shifts %d..%d: %d failures %s" % (first, last, len(failed), failed))
    sys.exit(1 if failed else 0)


if __name__ == "__main__":
    main()
