"""Pin the CPU oracle against outputs of the unmodified reference (tests/golden/*.npz)."""
import pytest
import torch

from oracle import field_oracle as O
from tests.util import CASES, Golden, rel_inf


@pytest.mark.parametrize("case", CASES)
def test_derived_constants(case):
    g = Golden(case)
    f = g.field()
    assert float(f.step_size) == g.meta["step_size"]
    assert f.n_samples == g.meta["nSamples"]


@pytest.mark.parametrize("case", CASES)
def test_stages_bit_exact(case):
    """Same torch ops in the same order => the oracle must reproduce the reference bit for bit."""
    g = Golden(case)
    f = g.field()
    keep = {}
    with torch.no_grad():
        rgb, depth, nv = O.render_chunk(f, g.rays, keep=keep, **g.call_kwargs())
    assert torch.equal(keep["z"], g.t["st_z"])
    assert torch.equal(keep["valid"], g.t["st_valid"])
    assert torch.equal(keep["app"], g.t["st_app"])
    assert torch.equal(keep["sigma"], g.t["st_sigma"])
    assert torch.equal(keep["weight"], g.t["st_weight"])
    assert int(nv) == int(g.t["n_valid"])
    assert torch.equal(rgb, g.t["rgb_map"])
    assert torch.equal(depth, g.t["depth_map"])


@pytest.mark.parametrize("case", CASES)
def test_gradients(case):
    g = Golden(case)
    loss, rgb, depth, nv, grads = O.loss_and_grads(g.field(), g.rays, g.target, **g.call_kwargs())
    assert float(loss) == pytest.approx(float(g.t["loss"]), rel=1e-6)
    assert set(grads) >= set(g.grads)
    for k, ref in g.grads.items():
        assert rel_inf(grads[k], ref) < 1e-6, k


def test_alpha_mask_rebuild_matches_reference():
    g = Golden("vm_fea_train_mask")
    f = g.field()
    f.alpha_mask = None
    O.build_alpha_mask(f, (24, 20, 22))
    assert torch.equal(f.alpha_mask.volume.bool(), g.t["am_volume"])
    assert torch.equal(f.alpha_mask.aabb, g.t["am_aabb"])


def test_free_mask_shipped_is_scalar_one():
    """as shipped the encoding masks are 0-dim ~1.0 and decomposition masks one vector."""
    m = O.free_mask(pos_bl=[12], view_bl=[12], fea_bl=[108], den_bl=[16], app_bl=[48], step=100,
                    total_step=3000, ratio=0.8)
    assert m["encoding"]["pos"].dim() == 0 and float(m["encoding"]["pos"]) == 1.0
    assert m["decomp"]["den"].shape == (16,)
    for k in range(3):
        assert float(m["decomp"]["den"][k]) == 1.0
    mi = O.free_mask(pos_bl=[12], view_bl=[12], fea_bl=[108], den_bl=[16] * 3, app_bl=[48] * 3, step=100,
                     total_step=3000, ratio=0.8, intended=True)
    assert mi["encoding"]["fea"].shape == (108,)
    assert len(mi["decomp"]["app"]) == 3 and mi["decomp"]["app"][2].shape == (48,)
    assert 0 < float(mi["encoding"]["fea"][-1]) < 1e-7


# ------------------------------------------------------------------ occupancy rebuild + ray pre-filter
@pytest.mark.parametrize("kind", ["vm", "cp"])
def test_occupancy_rebuild_and_ray_filter_match_reference(kind):
    from tests.util import OccGolden
    g = OccGolden(kind)
    f = g.field()
    box1 = O.build_alpha_mask(f, (24, 20, 22))                         # no mask yet: every lattice point evaluated
    assert torch.equal(f.alpha_mask.volume.bool(), g.t["am1_volume"])
    assert torch.equal(box1, g.t["am1_box"])
    assert torch.equal(O.filter_rays(f, g.rays, 48), g.t["keep1_s48"])
    assert torch.equal(O.filter_rays(f, g.rays, 24), g.t["keep1_s24"])
    assert torch.equal(O.filter_rays(f, g.rays, bbox_only=True), g.t["keep_bbox"])
    box2 = O.build_alpha_mask(f, (30, 26, 21))                         # rebuild under the first mask (compute_alpha :300-304)
    assert torch.equal(f.alpha_mask.volume.bool(), g.t["am2_volume"])
    assert torch.equal(box2, g.t["am2_box"])
    assert torch.equal(O.filter_rays(f, g.rays, 48), g.t["keep2_s48"])
