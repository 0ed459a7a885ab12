import numpy as np

from janus_b200.services.prosody import make_plan


def test_plan_ragged():
    p = make_plan([0, 1, 512, 513, 4096 * 3 + 7, 0], 512, 8, "cpu")
    assert p.n_streams == 6 and p.total_samples == 1 + 512 + 513 + 4096 * 3 + 7
    assert list(p.offsets) == [0, 0, 1, 513, 1026, 1026 + 12295]
    assert list(np.diff(p.hop_offsets)) == [0, 1, 1, 2, 25, 0]
    assert list(np.diff(p.group_offsets)) == [0, 1, 1, 1, 4, 0]
    assert p.total_hops == 29 and p.total_groups == 7
    raw = p.dev.numpy()
    assert list(raw[:6]) == list(p.offsets) and list(raw[6:12]) == list(p.lengths)
    assert list(raw[12:19]) == list(p.hop_offsets)
    assert list(raw[19:].view(np.int32)[:7]) == list(p.group_offsets)


def test_plan_empty():
    p = make_plan([], 512, 8, "cpu")
    assert p.n_streams == 0 and p.total_groups == 0
