"""Round trips at the sizes BASELINE.json quotes: generator -> segment files -> CUDA decode must give
back the generator's arrays, row by row, for a 1 GiB uncompressed segment (configs[1]), a mixed pglz segment
(configs[2]) and a packed buffer of 1000 images in one value (§8 f-3).  The oracle is not involved (it is checked
against the same generator in the small tests): the property is encode -> decode == identity, compared on the GPU."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    return torch


def check_round_trip(torch, seg, ind, dep, isb, dsb):
    from cerebro_ds_kpgi_b200 import segment
    st, part = segment.decode_pages(torch.from_numpy(seg.heap).cuda(), torch.from_numpy(seg.toast).cuda(),
                                    ind_sample_bytes=isb, dep_sample_bytes=dsb)
    assert st == 0
    lay = part.layout
    assert sorted(int(r) for r in lay["order"]) == list(range(len(ind)))
    assert int(part.summary.ind_bytes) == sum(a.nbytes for a in ind) and int(part.summary.dep_bytes) == sum(a.nbytes for a in dep)
    step = max(1, (256 << 20) // max(1, ind[0].nbytes))                     # compare about 256 MB at a time
    for r0 in range(0, len(ind), step):
        rows = range(r0, min(len(ind), r0 + step))
        want_i = torch.from_numpy(np.concatenate([ind[r].reshape(-1).view(np.uint8) for r in rows])).cuda()
        want_d = torch.from_numpy(np.concatenate([dep[r].reshape(-1).view(np.uint8) for r in rows])).cuda()
        got_i = torch.cat([part.ind[int(lay["ind_off"][r]):int(lay["ind_off"][r]) + int(lay["ind_len"][r])] for r in rows])
        got_d = torch.cat([part.dep[int(lay["dep_off"][r]):int(lay["dep_off"][r]) + int(lay["dep_len"][r])] for r in rows])
        assert torch.equal(got_i, want_i), "independent_var differs in rows %d.." % r0
        assert torch.equal(got_d, want_d), "dependent_var differs in rows %d.." % r0
    return part


def test_one_gib_uncompressed_segment_round_trip(torch_cuda):
    from cerebro_ds_kpgi_b200 import synth
    n = synth.rows_for_toast_bytes(1 << 30)
    ind, dep = synth.imagenet_rows(n, seed=2018, kind="norm_u8")
    seg = synth.build_segment(ind, dep, ind_storage="ext_raw", dep_storage="ext_raw", keep_expected=False)
    assert seg.toast.size >= 1000 << 20
    check_round_trip(torch_cuda, seg, ind, dep, 150528, 2000)


def test_mixed_pglz_segment_round_trip(torch_cuda):
    from cerebro_ds_kpgi_b200 import synth
    n = 1200
    ind, dep = synth.imagenet_rows(n, seed=7, kind="norm_u8")
    noise = np.random.default_rng(8)
    st = []
    for i in range(n):
        if i % 2:
            ind[i] = noise.standard_normal(ind[i].shape, dtype=np.float32)
            st.append("ext_literal")
        else:
            st.append("ext_pglz")
    seg = synth.build_segment(ind, dep, ind_storage=st, dep_storage="inline_pglz", keep_expected=False, effort=4)
    check_round_trip(torch_cuda, seg, ind, dep, 150528, 2000)


def test_packed_buffer_of_1000_images_round_trip(torch_cuda):
    from cerebro_ds_kpgi_b200 import _lib, synth
    ind, dep = synth.imagenet_rows(2, seed=9, kind="norm_u8", images_per_row=1000)      # 150 MB per value
    seg = synth.build_segment(ind, dep, ind_storage="ext_pglz", dep_storage="ext_pglz", keep_expected=False, effort=4)
    before = int(_lib.load().cds_wide_decoded_count())
    part = check_round_trip(torch_cuda, seg, ind, dep, 150528, 2000)
    assert int(part.summary.n_wide_values) == 4                             # both attributes of both rows
    assert int(_lib.load().cds_wide_decoded_count()) - before == 4
