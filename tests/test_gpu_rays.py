"""`-m gpu`: device ray generation / view rendering / resident sampler against the oracle's
restatement of dataLoader/ray_utils.py (fp32 tolerance: the reference's `directions @ R^T` is a BLAS matmul
whose summation order is not specified, so directions are compared to 2 ulp-scale 3e-7, origins exactly)."""
import math

import pytest
import torch

from oracle import field_oracle as O
from tests.util import Golden, rel_inf

pytestmark = pytest.mark.gpu


def _pose(elev=30.0, azim=45.0, radius=4.0311):
    el, az = math.radians(elev), math.radians(azim)
    o = torch.tensor([math.cos(el) * math.cos(az), math.cos(el) * math.sin(az), math.sin(el)]) * radius
    fwd = -o / o.norm()
    right = torch.linalg.cross(fwd, torch.tensor([0.0, 0.0, 1.0]))
    right = right / right.norm()
    down = torch.linalg.cross(fwd, right)
    return torch.cat([torch.stack([right, down, fwd], 1), o[:, None]], 1)        # 3x4, OpenCV convention


@pytest.mark.parametrize("H,W,normalize,blender,center", [(800, 800, True, False, None), (37, 53, False, False, None),
                                                          (16, 9, True, True, (4.0, 9.5)), (1, 1, True, False, None)])
def test_generate_rays_matches_ray_utils(H, W, normalize, blender, center):
    import fewtensorf_b200 as ft
    c2w = _pose()
    focal = [0.5 * 800 / math.tan(0.5 * 0.6911112) * (W / 800), 1111.0 * (H / 800)]
    want = O.camera_rays(H, W, focal, c2w, center, normalize, blender)
    got = ft.generate_rays(H, W, focal, c2w, center, normalize, blender).cpu()
    assert got.shape == want.shape
    assert torch.equal(got[:, :3], want[:, :3])
    assert float((got[:, 3:] - want[:, 3:]).abs().max()) < 3e-7
    # interleaved rows (what each GPU of a multi-GPU render generates)
    if H >= 4:
        part = ft.generate_rays(H, W, focal, c2w, center, normalize, blender, row0=1, row_step=3).cpu()
        assert torch.equal(part, got.view(H, W, 6)[1::3].reshape(-1, 6))
    with pytest.raises(RuntimeError, match="generate_rays"):
        ft.generate_rays(H, W, focal, c2w, row0=H, row_step=1, n_rows=1)


def test_render_view_equals_rendering_oracle_rays():
    import fewtensorf_b200 as ft
    from tests.gpu_util import model_from_field
    g = Golden("vm_fea_train_mask")
    f = g.field()
    m = model_from_field(f)
    H, W, focal = 24, 32, 30.0
    c2w = _pose(20.0, -60.0, 4.0)
    rays = O.camera_rays(H, W, [focal, focal], c2w)
    want_rgb, want_depth, _ = O.render_chunk(f, rays, None, white_bg=True, is_train=False, n_samples=-1)[:3]
    rgb, depth, nv = ft.render_view(m, c2w, H, W, focal, white_bg=True)
    assert rgb.shape == (H, W, 3) and depth.shape == (H, W) and nv > 0
    assert rel_inf(rgb.cpu().reshape(-1, 3), want_rgb) < 1e-4
    assert rel_inf(depth.cpu().reshape(-1), want_depth) < 1e-4
    top, _, _ = ft.render_view(m, c2w, H, W, focal, white_bg=True, row0=0, row_step=2)
    assert rel_inf(top.cpu(), rgb.cpu()[0::2]) < 1e-6


def test_device_sampler_epoch_logic():
    """train.py:44-56: a fresh permutation whenever fewer than one batch is left; every id once per epoch."""
    import fewtensorf_b200 as ft
    n, b = 1000, 96
    rays = torch.arange(n, dtype=torch.float32)[:, None].expand(n, 6).contiguous()
    rgbs = torch.arange(n, dtype=torch.float32)[:, None].expand(n, 3).contiguous()
    s = ft.DeviceRaySampler(rays, rgbs, b, generator=torch.Generator(device="cuda").manual_seed(3))
    seen, shuffles = [], 0
    for it in range(25):
        r, c = s.next()
        assert r.is_cuda and r.shape == (b, 6) and c.shape == (b, 3)
        assert torch.equal(r[:, 0], c[:, 0])                       # rays and colours stay paired
        shuffles += int(s.curr == 0)
        seen.append(r[:, 0].long().cpu())
    per_epoch = n // b                                             # 10 full batches, then reshuffle
    first = torch.cat(seen[:per_epoch])
    assert first.unique().numel() == per_epoch * b                 # no repeats inside an epoch
    assert shuffles == 3                                           # 25 batches = 2 full epochs + 5
