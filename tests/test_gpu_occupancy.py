"""`-m gpu`: occupancy rebuild (updateAlphaMask) and ray pre-filter (filtering_rays) kernels
against the reference-generated fixtures (tests/golden/occ_*.npz) and the oracle on larger seeded cases.
Bar: bit-exact masks, boxes and keep decisions (the planted box keeps every lattice alpha far from the
threshold, so fp32 summation order cannot flip a bit)."""
import pytest
import torch

from oracle import field_oracle as O
from tests.util import OccGolden

pytestmark = pytest.mark.gpu


def _volume_bool(m):
    return m.alphaMask.alpha_volume[0, 0].bool().cpu()


@pytest.mark.parametrize("kind", ["vm", "cp"])
def test_update_alpha_mask_and_filtering_match_reference_fixture(kind):
    from tests.gpu_util import model_from_field
    g = OccGolden(kind)
    m = model_from_field(g.field(), "cuda")
    rays = g.rays.cuda()
    tag = torch.arange(rays.shape[0], dtype=torch.float32)[:, None]

    def keep_of(**kw):
        kept_rays, kept = m.filtering_rays(g.rays, tag, **kw)          # host rays, as train.py:286 passes them
        k = torch.zeros(rays.shape[0], dtype=torch.bool)
        k[kept[:, 0].long()] = True
        assert torch.equal(kept_rays, g.rays[k])
        return k

    box1 = m.updateAlphaMask((24, 20, 22))
    assert torch.equal(_volume_bool(m), g.t["am1_volume"])
    assert torch.equal(box1.cpu(), g.t["am1_box"])
    assert m.alphaMask.gridSize.tolist() == [24, 20, 22]
    assert torch.equal(keep_of(N_samples=48, chunk=50), g.t["keep1_s48"])
    assert torch.equal(keep_of(N_samples=24), g.t["keep1_s24"])
    assert torch.equal(keep_of(bbox_only=True, chunk=64), g.t["keep_bbox"])
    # device-resident rays: one launch, same decisions
    kr, kc = m.filtering_rays(rays, tag.cuda(), N_samples=48)
    assert torch.equal(kc[:, 0].long().cpu(), torch.nonzero(g.t["keep1_s48"])[:, 0])
    box2 = m.updateAlphaMask((30, 26, 21))                              # rebuild under the first mask
    assert torch.equal(_volume_bool(m), g.t["am2_volume"])
    assert torch.equal(box2.cpu(), g.t["am2_box"])
    assert torch.equal(keep_of(N_samples=48), g.t["keep2_s48"])
    assert abs(m.last_alpha_rest - float(g.t["am2_volume"].double().mean())) < 1e-12


@pytest.mark.parametrize("grid,mask_grid", [((48, 40, 44), (37, 65, 50)), ((33, 33, 33), (64, 32, 96))])
def test_update_alpha_mask_matches_oracle_ragged_word_sizes(grid, mask_grid):
    """Lattice sizes around the 32-texel word granularity (x = 37, 64; partial last words, multi-word rows)."""
    from tests.gpu_util import model_from_field
    cfg = O.FieldConfig(grid=grid, density_n_comp=(16, 16, 16), app_n_comp=(8, 8, 8))
    p = O.init_params(cfg, seed=3)
    O.plant_box(cfg, p, half=0.7)
    f = O.Field(cfg, p)
    m = model_from_field(f, "cuda")
    box_o = O.build_alpha_mask(f, mask_grid)
    box = m.updateAlphaMask(mask_grid)
    assert torch.equal(_volume_bool(m), f.alpha_mask.volume.bool())
    assert torch.equal(box.cpu(), box_o)
    # the mask the kernels wrote drives the march exactly like one packed from the float volume
    rays = O.synthetic_rays(512, seed=4)
    want = O.filter_rays(f, rays, 100)
    got_rays, _ = m.filtering_rays(rays, torch.zeros(512, 3), N_samples=100)
    assert torch.equal(got_rays, rays[want])
    import fewtensorf_b200 as ft
    m2 = model_from_field(f, "cuda")                                    # float volume -> ftrf_pack_alpha_mask
    assert torch.equal(m2.alphaMask.packed()[0], m.alphaMask.packed()[0])
    assert isinstance(m.alphaMask, ft.AlphaGridMask)


def test_update_alpha_mask_empty_scene_raises_like_reference():
    from tests.gpu_util import model_from_field
    cfg = O.FieldConfig(grid=(16, 16, 16), density_n_comp=(8, 8, 8), app_n_comp=(8, 8, 8))
    f = O.Field(cfg, O.init_params(cfg, seed=1))                        # un-planted: sigma ~ softplus(-10) everywhere
    m = model_from_field(f, "cuda")
    with pytest.raises(RuntimeError, match="updateAlphaMask"):
        m.updateAlphaMask((16, 16, 16))
    with pytest.raises(RuntimeError, match="filter_rays"):
        import ctypes as C
        from fewtensorf_b200._abi import lib, check
        check(lib.ftrf_filter_rays(C.byref(m._fill_field(None)), None, 6, 4, 8, 1, None, None), "filter_rays")


def test_repeated_rebuilds_never_see_a_stale_mask():
    """The cached call descriptor is keyed on a serial number of the mask, not on id(): masks replaced in a loop
    reuse Python object addresses, and a stale descriptor would point at a freed bit field."""
    from tests.gpu_util import model_from_field
    cfg = O.FieldConfig(grid=(40, 40, 40), density_n_comp=(16, 16, 16), app_n_comp=(8, 8, 8))
    p = O.init_params(cfg, seed=5)
    O.plant_box(cfg, p)
    m = model_from_field(O.Field(cfg, p), "cuda")
    m.updateAlphaMask((64, 64, 64))
    m.updateAlphaMask((64, 64, 64))
    want = m.alphaMask.packed()[0].clone()
    for _ in range(8):
        junk = torch.randint(0, 2 ** 31 - 1, (1 << 16,), dtype=torch.int32, device="cuda")     # recycle freed blocks
        m.updateAlphaMask((64, 64, 64))
        assert torch.equal(m.alphaMask.packed()[0], want)
        del junk
