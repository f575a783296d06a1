"""Text ETL ("next" row f-4): oracle restatement vs the reference parsers' outputs (CPU), CUDA path vs both (GPU)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

import da_oracle


def golden():
    return np.load(os.path.join(GOLDEN, "text_etl.npz"))


def test_oracle_text_parse_matches_reference_parsers():
    z = golden()
    ids, feats, labels = da_oracle.text_parse(z["text_a"].tobytes())
    assert np.array_equal(ids, z["ids_a"]) and np.array_equal(labels, z["labels_a"])
    assert feats.dtype == np.float32 and feats.tobytes() == z["feats_a"].tobytes()
    ids, feats, labels = da_oracle.text_parse(z["text_b"].tobytes(), label_digits=2)
    assert np.array_equal(ids, z["ids_b"]) and np.array_equal(labels, z["labels_b"])
    assert feats.tobytes() == z["feats_b"].tobytes()


@pytest.mark.gpu
def test_gpu_text_etl_matches_reference_parsers():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    from cerebro_ds_kpgi_b200 import etl
    z = golden()
    ids, feats, labels = etl.parse_text(z["text_a"].tobytes(), 36, feature_shape=(4, 3, 3))
    assert np.array_equal(ids, z["ids_a"]) and np.array_equal(labels, z["labels_a"])
    assert feats.shape == z["feats_a"].shape and feats.tobytes() == z["feats_a"].tobytes()      # 0 ULP, signed zero included
    ids, feats, labels = etl.parse_text(z["text_b"].tobytes(), 64, label_digits=2)
    assert np.array_equal(ids, z["ids_b"]) and np.array_equal(labels, z["labels_b"])
    assert feats.tobytes() == z["feats_b"].tobytes()
    rid, f, lab = etl.parser("17|{{0.5,-1.25e-05},{3,1e10}}|42\n", 4)
    assert rid == "17" and lab == 42 and f.tobytes() == np.array([0.5, -1.25e-05, 3, 1e10], np.float32).tobytes()


@pytest.mark.gpu
def test_gpu_text_etl_random_lines_vs_oracle_and_refusals():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    from cerebro_ds_kpgi_b200 import etl
    rng = np.random.default_rng(4)
    V = 301
    lines = []
    for r in range(257):
        vals = rng.standard_normal(V).astype(np.float32)
        lines.append("%d|{%s}|%d\n" % (r, ",".join("%.6g" % v for v in vals), r % 1000))     # PostgreSQL float4 text form
    text = "".join(lines).encode()
    assert len(text) > 3 * 4096                                                              # several tiles, tokens straddle them
    want = da_oracle.text_parse(text)
    got = etl.parse_text(text, V)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[2], want[2])
    assert got[1].tobytes() == want[1].tobytes()
    # literal_eval raises on these; so does the GPU parser
    # the whole float4 range in PostgreSQL's text form, subnormals and extremes included, plus long mantissas
    vals = np.concatenate([np.float32(10.0) ** rng.uniform(-44.9, 38.5, 4000).astype(np.float32) * rng.choice([-1, 1], 4000),
                           np.array([1.17549435e-38, 3.40282347e+38, 1.4e-45, 0.1, 16777217.0], dtype=np.float64)])
    toks = ["%.9g" % v for v in vals] + ["%.17g" % float(v) for v in vals[:500]] + ["1234567890123456789", "0.000000000000000000012345"]
    line = ("5|{%s}|3\n" % ",".join(toks)).encode()
    want = da_oracle.text_parse(line)
    got = etl.parse_text(line, len(toks))
    assert got[1].tobytes() == want[1].tobytes()
    for bad in (b"1|{NaN,1}|0\n", b"1|{Infinity,1}|0\n", b"1|{1.5.2,1}|0\n", b"1|{1e-300,1}|0\n", b"1|{12345678901234567891,1}|0\n"):
        with pytest.raises(ValueError):
            etl.parse_text(bad, 2)
    with pytest.raises(ValueError):
        etl.parse_text(b"1|{1,2,3}|0\n2|{1,2}|0\n", 3)                                       # ragged row
    with pytest.raises(ValueError):
        etl.parse_text(b"1|{1,2,3}|0\n2|{1,2,3}\n", 3)                                       # missing label
