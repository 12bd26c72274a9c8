"""CPU: the FreeNeRF mask schedule and resolution helpers (fewtensorf_b200/schedule.py, and the oracle's
restatement) against outputs of the unmodified reference functions (tests/golden/sched_free_mask.npz, made by
tests/golden/make_golden_schedule.py from utils.py:11-70, :117-125)."""
import os

import numpy as np
import pytest
import torch

from oracle import field_oracle as O
from tests.golden.make_golden_schedule import CASES, flatten

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sched_free_mask.npz"))


def _same(prefix, mask):
    got = {}
    flatten(prefix, mask, got)
    want = {k: GOLD[k] for k in GOLD.files if k.startswith(prefix + "/")}
    assert set(got) == set(want), (sorted(got), sorted(want))
    for k in want:
        assert np.array_equal(np.asarray(got[k]), want[k]), k


@pytest.mark.parametrize("i", range(len(CASES)))
def test_product_schedule_is_the_reference_schedule(i):
    from fewtensorf_b200.schedule import get_free_mask
    _same(f"c{i}", get_free_mask(**CASES[i]))


@pytest.mark.parametrize("i", [i for i, c in enumerate(CASES) if "max_visible" not in c])
def test_oracle_schedule_is_the_reference_schedule(i):
    kw = dict(CASES[i])
    _same(f"c{i}", O.free_mask(**kw))


def test_intended_schedule_gives_one_vector_per_length():
    from fewtensorf_b200.schedule import get_free_mask
    m = get_free_mask(pos_bl=[12], view_bl=[12], fea_bl=[108], den_bl=[16] * 3, app_bl=[48] * 3, step=100,
                      total_step=3000, ratio=0.8, intended=True)
    assert m["encoding"]["fea"].shape == (108,) and len(m["decomp"]["app"]) == 3
    o = O.free_mask(pos_bl=[12], view_bl=[12], fea_bl=[108], den_bl=[16] * 3, app_bl=[48] * 3, step=100,
                    total_step=3000, ratio=0.8, intended=True)
    assert torch.equal(m["encoding"]["fea"], o["encoding"]["fea"])
    for k in range(3):
        assert torch.equal(m["decomp"]["den"][k], o["decomp"]["den"][k])


def test_resolution_helpers():
    from fewtensorf_b200.schedule import N_to_reso, cal_n_samples
    aabb = torch.tensor([[-1.5, -1.5, -1.5], [1.5, 1.5, 1.5]])
    box2 = torch.tensor([[-1.0, -0.5, -2.0], [1.5, 1.5, 1.0]])
    for j, (n, b) in enumerate(((128 ** 3, aabb), (300 ** 3, aabb), (27000000, box2), (2097156, box2))):
        r = N_to_reso(n, b)
        assert r == GOLD[f"reso{j}"].tolist()
        assert cal_n_samples(r, 0.5) == int(GOLD[f"nsamp{j}"])
    assert cal_n_samples([300, 300, 300], 0.5) == 1039          # S at 300^3
