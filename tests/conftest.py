import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(REPO, "tests", "golden")
for p in (REPO, os.path.join(REPO, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # tests/test_zz_* files are meant to run last; keep pytest's default (alphabetical by file) order stable
    items.sort(key=lambda it: (os.path.basename(str(it.fspath)).startswith("test_zz"),))


SEGMENT_CASES = ["seg_small_mixed", "seg_imagenet_rows", "seg_criteo", "seg_many_shuffled",
                 "seg_multichunk_shuffled"]


def load_golden_segment(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    order = [int(k) for k in z["order"]]
    expected = {k: {"independent_var": z["ind_%d" % k], "dependent_var": z["dep_%d" % k]} for k in order}
    return z["heap"], z["toast"], order, expected


@pytest.fixture(scope="session")
def built():
    """Build the host-side libraries once (gcc only; the CUDA library is built by __graft_entry__.build)."""
    from cerebro_ds_kpgi_b200 import build
    build.build_synth()
    build.build_oracle()
    return build


@pytest.fixture(scope="session")
def coracle(built):
    import da_oracle
    return da_oracle.COracle()


def assert_dict_equal(got, expected, order=None):
    """Bit-exact comparison of reader results, including dict insertion order."""
    if order is not None:
        assert [int(k) for k in got.keys()] == list(order)
    assert set(int(k) for k in got.keys()) == set(int(k) for k in expected.keys())
    for k, v in expected.items():
        g = got[k]
        for att in ("independent_var", "dependent_var"):
            a, b = np.asarray(g[att]), np.asarray(v[att])
            assert a.dtype == b.dtype, (k, att, a.dtype, b.dtype)
            assert a.shape == b.shape, (k, att, a.shape, b.shape)
            assert a.tobytes() == b.tobytes(), (k, att)
