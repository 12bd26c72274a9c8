"""`-m gpu`: factor maintenance on kernels (second half): upsample_volume_grid
(models/tensoRF.py:267-288, :418-435 — bilinear, align_corners), shrink (:290-327, :437-466) and vector_comp_diffs
(:175-187) against the torch expressions the reference uses, on the same parameters."""
import pytest
import torch
import torch.nn.functional as F

from oracle import field_oracle as O
from tests.util import rel_inf

pytestmark = pytest.mark.gpu


def _model(kind="VM", grid=(20, 28, 36)):
    from tests.gpu_util import model_from_field
    comps = (8, 16, 4) if kind == "VM" else (12, 12, 12)
    cfg = O.FieldConfig(model=kind, grid=grid, density_n_comp=comps, app_n_comp=comps if kind == "CP" else (16, 8, 24))
    p = O.init_params(cfg, seed=7)
    return model_from_field(O.Field(cfg, p)), p


@pytest.mark.parametrize("kind", ["VM", "CP"])
@pytest.mark.parametrize("target", [(41, 33, 57), (20, 28, 36), (13, 64, 21)])
def test_upsample_matches_interpolate(kind, target):
    m, p = _model(kind)
    m.upsample_volume_grid(target)
    torch.cuda.synchronize()
    mat, vec = O.MAT_MODE, O.VEC_MODE
    for k in range(3):
        for name in ("density", "app"):
            want = F.interpolate(p[f"{name}_line.{k}"].cuda(), size=(target[vec[k]], 1), mode="bilinear", align_corners=True)
            got = getattr(m, f"{name}_line")[k]
            assert got.shape == want.shape and rel_inf(got, want) < 2e-6, (name, "line", k)
            if kind == "VM":
                want = F.interpolate(p[f"{name}_plane.{k}"].cuda(), size=(target[mat[k][1]], target[mat[k][0]]), mode="bilinear",
                                     align_corners=True)
                got = getattr(m, f"{name}_plane")[k]
                assert got.shape == want.shape and rel_inf(got, want) < 2e-6, (name, "plane", k)
    assert m.gridSize.tolist() == list(target)
    # the resampled field still renders (layout handed to the gather kernels is the one they expect)
    rays = O.synthetic_rays(64, seed=1).cuda()
    rgb, depth, nv = m(rays, None, white_bg=True, is_train=False, N_samples=64)
    assert torch.isfinite(rgb).all()


@pytest.mark.parametrize("kind", ["VM", "CP"])
def test_shrink_matches_slicing(kind):
    m, p = _model(kind, grid=(24, 30, 36))
    # a mask of the field's own resolution, so shrink takes the un-corrected aabb path like the reference's first shrink
    import fewtensorf_b200 as ft
    m.alphaMask = ft.AlphaGridMask("cuda", m.aabb, torch.ones(36, 30, 24, device="cuda"))
    new_aabb = torch.tensor([[-0.9, -0.6, -1.2], [0.7, 1.1, 0.4]], device="cuda")
    units = m.units.clone()
    t_l = torch.round(torch.round((new_aabb[0] - m.aabb[0]) / units)).long()
    b_r = torch.round((new_aabb[1] - m.aabb[0]) / units).long() + 1
    b_r = torch.stack([b_r, m.gridSize]).amin(0)
    m.shrink(new_aabb)
    torch.cuda.synchronize()
    mat, vec = O.MAT_MODE, O.VEC_MODE
    for k in range(3):
        for name in ("density", "app"):
            want = p[f"{name}_line.{k}"].cuda()[..., t_l[vec[k]]:b_r[vec[k]], :]
            assert torch.equal(getattr(m, f"{name}_line")[k].detach(), want)
            if kind == "VM":
                want = p[f"{name}_plane.{k}"].cuda()[..., t_l[mat[k][1]]:b_r[mat[k][1]], t_l[mat[k][0]]:b_r[mat[k][0]]]
                assert torch.equal(getattr(m, f"{name}_plane")[k].detach(), want)
    assert m.gridSize.tolist() == (b_r - t_l).tolist()


@pytest.mark.parametrize("kind", ["VM", "CP"])
def test_vector_comp_diffs_value_and_gradient(kind):
    m, p = _model(kind)
    loss = m.vector_comp_diffs()
    (loss * 3.0).backward()
    torch.cuda.synchronize()
    leaves = {}
    total = 0
    for name in ("density", "app"):
        lines = [p[f"{name}_line.{k}"].clone().double().cuda().requires_grad_(True) for k in range(3)]
        leaves[name] = lines
        total = total + O.vector_diffs(lines)
    (total * 3.0).backward()
    assert float(loss) == pytest.approx(float(total), rel=2e-6)
    for name in ("density", "app"):
        for k in range(3):
            got = getattr(m, f"{name}_line")[k].grad
            assert got is not None and rel_inf(got, leaves[name][k].grad) < 2e-5, (name, k)
