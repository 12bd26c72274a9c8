"""CPU: `TensorBase.grad_regions()` — the parts of the flat gradient buffer a multi-GPU step exchanges — must cover
every entry the reference's autograd makes non-zero (oracle gradients on seeded rays under an alphaMask), must be
whole float4 runs, and must shrink with the mask's bounding box.  (The kernels that move them need GPUs:
tests/test_gpu_multi.py.)"""
import pytest
import torch

from oracle import field_oracle as O


def _setup(kind, mask_grid, half):
    import fewtensorf_b200 as ft
    comps = (8, 12, 16) if kind == "VM" else (12, 12, 12)
    cfg = O.FieldConfig(model=kind, grid=(40, 48, 56), density_n_comp=comps, app_n_comp=comps)
    p = O.init_params(cfg, seed=1)
    O.plant_box(cfg, p, half=half)
    f = O.Field(cfg, p)
    if mask_grid:
        O.build_alpha_mask(f, mask_grid)
    from tests.gpu_util import args_from_cfg
    aabb = torch.tensor(cfg.aabb)
    if kind == "VM":
        m = ft.TensorVMSplit(args_from_cfg(cfg), aabb, list(cfg.grid), list(cfg.near_far), "cpu")
    else:
        m = ft.TensorCP(args_from_cfg(cfg), aabb, list(cfg.grid), "cpu", near_far=list(cfg.near_far))
    m.load_state_dict(p)
    if mask_grid:
        m.alphaMask = ft.AlphaGridMask("cpu", f.alpha_mask.aabb, f.alpha_mask.volume)
    return f, m


@pytest.mark.parametrize("kind", ["VM", "CP"])
@pytest.mark.parametrize("mask_grid,half", [((30, 34, 38), 0.5), ((64, 64, 64), 0.3), (None, 0.5)])
def test_regions_cover_every_nonzero_gradient(kind, mask_grid, half):
    f, m = _setup(kind, mask_grid, half)
    m._mask_serial = lambda: 1                       # (the packed mask needs a GPU; the regions only need the volume)
    rays = O.synthetic_rays(2500, seed=2)
    target = torch.rand(2500, 3, generator=torch.Generator().manual_seed(3))
    jitter = torch.rand(2500, 1, generator=torch.Generator().manual_seed(4))
    _, _, _, nv, grads = O.loss_and_grads(f, rays, target, is_train=True, white_bg=True, n_samples=200, jitter=jitter)
    assert int(nv) > 1000
    params = m._all_params()
    names = [k for k, _ in m.named_parameters()]
    by_id = {id(p): k for k, p in m.named_parameters()}
    total = sum(p.numel() for p in params) + 4
    regions = m.grad_regions()
    covered = torch.zeros((total + 3) // 4 * 4, dtype=torch.bool)
    for off, row_len, stride, rows in regions:
        assert off % 4 == 0 and row_len % 4 == 0 and stride % 4 == 0 and rows >= 1
        for r in range(rows):
            assert not covered[off + r * stride: off + r * stride + row_len].any(), "regions overlap"
            covered[off + r * stride: off + r * stride + row_len] = True
    off = 0
    for p in params:
        g = grads[by_id[id(p)]]
        flat = g.permute(0, 2, 3, 1).reshape(-1) if g.dim() == 4 else g.reshape(-1)     # buffer order is channels-last
        nz = flat != 0
        assert bool(covered[off:off + p.numel()][nz].all()), by_id[id(p)]
        off += p.numel()
    frac = float(covered.sum()) / covered.numel()
    if mask_grid is None:
        assert frac == 1.0
    elif kind == "VM":
        assert frac < (0.5 if half == 0.5 else 0.35), frac      # ~ (box side / aabb side)^2 of the planes


def test_empty_mask_exchanges_only_the_dense_tail():
    import fewtensorf_b200 as ft
    f, m = _setup("VM", None, 0.5)
    m.alphaMask = ft.AlphaGridMask("cpu", torch.tensor(f.cfg.aabb), torch.zeros(8, 8, 8))
    regions = m.grad_regions()
    n_factor = sum(p.numel() for p in m._factor_params())
    assert all(r[0] >= n_factor or r[1] * r[3] <= 64 for r in regions)
