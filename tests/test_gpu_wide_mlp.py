"""`-m gpu`: BASELINE config 4's shading module (configs/free_tensorf_human.txt: shadingMode MLP, pos_pe 8 ->
198 input columns, FreeNeRF encoding / decomposition masks) against the oracle.  The first layer is too wide
for one resident tcgen05 weight operand, so forward, dgrad and wgrad each run as two launches
(k_shade_tc.cu: K-split with accumulating epilogue, N-split, column-split) — same 1e-4 bar as everywhere."""
import pytest
import torch

from oracle import field_oracle as O
from tests.util import rel_inf

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("masked", [False, True])
def test_mlp_pos_pe8_matches_oracle(masked):
    from tests.gpu_util import model_from_field, pin_rng
    cfg = O.FieldConfig(grid=(32, 32, 32), shading="MLP", pos_pe=8, view_pe=2, fea_pe=2, density_n_comp=(32, 32, 32),
                        app_n_comp=(48, 48, 48))
    assert O.mlp_in_dim(cfg) == 198
    p = O.init_params(cfg, seed=11)
    O.plant_box(cfg, p)
    f = O.Field(cfg, p)
    O.build_alpha_mask(f, (32, 32, 32))
    mask = None
    if masked:
        mask = O.free_mask(pos_bl=[48], view_bl=[12], fea_bl=[108], den_bl=[32] * 3, app_bl=[48] * 3, step=900,
                           total_step=3000, ratio=0.65, intended=True)
    n = 700                                     # several 128-row tiles with a ragged tail
    rays = O.synthetic_rays(n, seed=12)
    g = torch.Generator().manual_seed(13)
    target, jitter = torch.rand(n, 3, generator=g), torch.rand(n, 1, generator=g)
    loss_o, rgb_o, depth_o, nv_o, grads_o = O.loss_and_grads(f, rays, target, mask=mask, is_train=True, jitter=jitter,
                                                             n_samples=110)
    assert int(nv_o) > 2000
    m = model_from_field(f, "cuda")
    pin_rng(m, jitter, False, True)
    dmask = None
    if masked:
        to = lambda v: None if v is None else ([t.cuda() for t in v] if isinstance(v, (list, tuple)) else v.cuda())
        dmask = {"encoding": {k: to(v) for k, v in mask["encoding"].items()},
                 "decomp": {k: to(v) for k, v in mask["decomp"].items()}}
    rgb, depth, nv = m(rays.cuda(), dmask, white_bg=True, is_train=True, N_samples=110)
    loss = torch.mean((rgb - target.cuda()) ** 2)
    loss.backward()
    assert abs(int(nv) - int(nv_o)) <= max(2, int(nv_o) // 1000)
    assert rel_inf(rgb.detach().cpu(), rgb_o) < 1e-4
    assert rel_inf(depth.cpu(), depth_o) < 1e-4
    for k, pp in m.named_parameters():
        assert rel_inf(pp.grad.cpu(), grads_o[k]) < 1e-4, k
