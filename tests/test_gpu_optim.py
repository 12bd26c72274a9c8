"""`-m gpu`: factor regularisers (`ftrf_reg_factors`) and the one-launch Adam step (`ftrf_adam_step`) against
the reference-generated fixture tests/golden/optim_vm.npz, the CPU oracle on other shapes, and — at the
BASELINE 300^3 size — torch's own CUDA arithmetic on the same tensors."""
import pytest
import torch

from oracle import field_oracle as O
from tests.util import OptimGolden, pseudo_grads, rel_inf

pytestmark = pytest.mark.gpu
TOL = 1e-4            # <= 1e-4 relative for fp32 values and gradients


@pytest.fixture()
def gm():
    from tests.gpu_util import model_from_field
    g = OptimGolden()
    return g, model_from_field(g.field())


def test_regulariser_values_and_autograd_match_reference(gm):
    import fewtensorf_b200 as ft
    g, m = gm
    reg = ft.TVLoss()
    tv_d, tv_a, l1, ortho = m.TV_loss_density(reg), m.TV_loss_app(reg), m.density_L1(), m.vector_comp_diffs()
    assert float(tv_d) == pytest.approx(float(g.t["tv_density"]), rel=1e-5)
    assert float(tv_a) == pytest.approx(float(g.t["tv_app"]), rel=1e-5)
    assert float(l1) == pytest.approx(float(g.t["l1"]), rel=1e-5)
    assert float(ortho) == pytest.approx(float(g.t["ortho"]), rel=1e-5)
    (g.w_l1 * l1 + g.w_tvd * tv_d + g.w_tva * tv_a).backward()
    want = g.group("greg")
    for k, p in m.named_parameters():
        if k in want:
            assert rel_inf(p.grad.cpu(), want[k]) < TOL, k
        else:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k
    for p in m.parameters():
        p.grad = None
    (g.w_o * m.vector_comp_diffs()).backward()
    for k, w in g.group("gortho").items():
        assert rel_inf(dict(m.named_parameters())[k].grad.cpu(), w) < TOL, k


def test_fused_regularisers_add_into_existing_gradients(gm):
    g, m = gm
    params = dict(m.named_parameters())
    with torch.no_grad():
        for k, p in params.items():
            p.grad = torch.full_like(p, 0.25, memory_format=torch.preserve_format)
    vals = m.add_regularizers_(L1_weight=g.w_l1, TV_weight_density=g.w_tvd, TV_weight_app=g.w_tva)
    assert vals.dtype == torch.float64 and vals.shape == (3,)
    assert float(vals[0]) == pytest.approx(float(g.t["tv_density"]), rel=1e-6)
    assert float(vals[1]) == pytest.approx(float(g.t["l1"]), rel=1e-6)
    assert float(vals[2]) == pytest.approx(float(g.t["tv_app"]), rel=1e-6)
    want = g.group("greg")
    for k, p in params.items():
        got = p.grad.cpu() - 0.25
        if k in want:
            assert float((got - want[k]).abs().max()) < TOL * float(want[k].abs().max()) + 1e-7, k
        else:
            assert float(got.abs().max()) == 0.0, k


def test_regularisers_ragged_planes_vs_oracle():
    """Three different plane shapes and channel counts; a custom TVLoss weight; a foreign `reg` callable."""
    import fewtensorf_b200 as ft
    from tests.gpu_util import model_from_field
    cfg = O.FieldConfig(model="VM", shading="SH", grid=(7, 33, 18), density_n_comp=(4, 12, 8), app_n_comp=(8, 4, 20),
                        app_dim=27)
    f = O.Field(cfg, O.init_params(cfg, seed=3, scale=0.7), None)
    m = model_from_field(f)
    planes = [f.params[f"app_plane.{k}"] for k in range(3)]
    lines = [f.params[f"density_line.{k}"] for k in range(3)]
    dplanes = [f.params[f"density_plane.{k}"] for k in range(3)]
    tv = m.TV_loss_app(ft.TVLoss(2.5))
    assert float(tv) == pytest.approx(2.5 * float(O.tv_loss_planes(planes)), rel=1e-5)
    generic = m.TV_loss_app(lambda x: O.tv_loss(x) * 2.5)         # not a TVLoss instance: generic torch path
    assert float(generic) == pytest.approx(float(tv), rel=1e-5)
    (0.3 * tv + 2.0 * m.density_L1()).backward()
    gp, _ = O.reg_grads(planes, [], 0.3 * 2.5, 0.0)
    gd, gl = O.reg_grads(dplanes, lines, 0.0, 2.0)
    for k in range(3):
        assert rel_inf(m.app_plane[k].grad.cpu(), gp[k]) < TOL
        assert rel_inf(m.density_plane[k].grad.cpu(), gd[k]) < TOL
        assert rel_inf(m.density_line[k].grad.cpu(), gl[k]) < TOL


def test_cp_l1_and_tv_refusal():
    from tests.gpu_util import model_from_field
    import fewtensorf_b200 as ft
    cfg = O.FieldConfig(model="CP", shading="SH", grid=(16, 20, 24), density_n_comp=(8, 8, 8), app_n_comp=(16, 16, 16),
                        app_dim=27)
    f = O.Field(cfg, O.init_params(cfg, seed=4), None)
    m = model_from_field(f)
    lines = [f.params[f"density_line.{k}"] for k in range(3)]
    l1 = m.density_L1()
    assert float(l1) == pytest.approx(float(O.density_l1([], lines)), rel=1e-5)
    l1.backward()
    _, gl = O.reg_grads([], lines, 0.0, 1.0)
    for k in range(3):
        assert rel_inf(m.density_line[k].grad.cpu(), gl[k]) < TOL
    with pytest.raises(RuntimeError, match="0/0"):          # loss.py:129 with W = 1: NaN upstream
        m.TV_loss_density(ft.TVLoss())
    with pytest.raises(RuntimeError):
        m.add_regularizers_(TV_weight_app=1.0)


def test_fused_adam_matches_reference_three_steps(gm):
    import fewtensorf_b200 as ft
    g, m = gm
    opt = ft.FusedAdam(m.get_optparam_groups(0.02, 0.001), betas=(0.9, 0.99))
    params = dict(m.named_parameters())
    pg = pseudo_grads([(k, tuple(g.group("p0")[k].shape)) for k in g.order])
    for t in (1, 2, 3):
        for k in g.order:
            p = params[k]
            grad = pg[t][k].cuda()
            if t == 1:      # gradient laid out like the parameter (what the backward kernels produce)
                grad = torch.empty_like(p, memory_format=torch.preserve_format).copy_(grad)
            p.grad = grad   # t > 1: plain contiguous gradients for channel-last parameters (repacked by step())
        opt.step()
        for k, want in g.group(f"adam_p{t}").items():
            assert rel_inf(params[k].detach().cpu(), want) < 2e-6, (t, k)
        for group in opt.param_groups:
            group["lr"] = group["lr"] * g.lr_factor
    for k, want in g.group("adam_m").items():
        assert rel_inf(opt.state[params[k]]["exp_avg"].cpu(), want) < 2e-6, k
    for k, want in g.group("adam_v").items():
        assert rel_inf(opt.state[params[k]]["exp_avg_sq"].cpu(), want) < 2e-6, k
    assert float(opt.state[params["basis_mat.weight"]]["step"]) == 3.0
    sd = opt.state_dict()                       # same keys as torch.optim.Adam: checkpoints interchange
    ref = torch.optim.Adam(m.get_optparam_groups(0.02, 0.001), betas=(0.9, 0.99))
    ref.load_state_dict(sd)
    assert float(ref.state[params["basis_mat.weight"]]["step"]) == 3.0


def test_full_size_reg_and_adam_vs_torch_cuda():
    """BASELINE 300^3 planes (17.3 M elements): the kernels against torch's CUDA arithmetic on the same tensors,
    and an unaligned / odd-length segment table for Adam."""
    import bench
    import fewtensorf_b200 as ft
    dev = torch.device("cuda", 0)
    m = bench.build_model(dev)
    reg = ft.TVLoss()
    want_tv = sum(reg(p.detach()) * 1e-2 for p in m.app_plane)
    assert float(m.TV_loss_app(reg)) == pytest.approx(float(want_tv), rel=1e-5)
    leaves = [p.detach().clone().requires_grad_(True) for p in m.density_plane]
    lines = [p.detach().clone().requires_grad_(True) for p in m.density_line]
    loss = 0.4 * sum(reg(p) * 1e-2 for p in leaves) + 3.0 * sum(p.abs().mean() for p in leaves + lines)
    loss.backward()
    for p in m.parameters():
        p.grad = None
    vals = m.add_regularizers_(L1_weight=3.0, TV_weight_density=0.4)
    for k in range(3):
        assert rel_inf(m.density_plane[k].grad, leaves[k].grad) < TOL
        assert rel_inf(m.density_line[k].grad, lines[k].grad) < TOL
    assert float(vals[1]) == pytest.approx(float(sum(p.abs().mean() for p in leaves + lines)), rel=1e-5)
    # Adam: all parameters + an odd-length, 4-byte-aligned extra tensor, against torch.optim.Adam
    extra = torch.randn(1001, device=dev)[1:].clone().requires_grad_(True)
    odd = torch.nn.Parameter(torch.randn(4099, device=dev)[3:])            # data_ptr not 16-byte aligned
    mine = [p for p in m.parameters()] + [odd]
    theirs = [p.detach().clone().requires_grad_(True) for p in mine]
    a = ft.FusedAdam([{"params": mine[:5], "lr": 0.02}, {"params": mine[5:], "lr": 0.001}], betas=(0.9, 0.99))
    b = torch.optim.Adam([{"params": theirs[:5], "lr": 0.02}, {"params": theirs[5:], "lr": 0.001}], betas=(0.9, 0.99))
    gen = torch.Generator(device=dev).manual_seed(1)
    for t in range(2):
        for p, q in zip(mine, theirs):
            gr = torch.randn(p.shape, device=dev, generator=gen) * 1e-3
            p.grad = torch.empty_like(p, memory_format=torch.preserve_format).copy_(gr)
            q.grad = gr
        a.step()
        b.step()
    for p, q in zip(mine, theirs):
        assert rel_inf(p.detach(), q.detach()) < 2e-6
    del extra


def test_guarded_adam_drops_the_step_when_the_flag_is_set():
    """FusedAdam(guard=...): a non-zero device flag turns the launch into a no-op, zero leaves it the plain step."""
    import fewtensorf_b200 as ft
    gen = torch.Generator().manual_seed(3)
    shapes = [(1, 16, 40, 48), (1, 16, 56, 1), (27, 48), (128, 150), (128,), (5000,)]
    base = [torch.randn(s, generator=gen) for s in shapes]
    grads = [torch.randn(s, generator=gen) for s in shapes]

    def run(flag_value):
        ps = [torch.nn.Parameter(b.clone().cuda()) for b in base]
        flag = None if flag_value is None else torch.full((1,), flag_value, device="cuda")
        opt = ft.FusedAdam(ps, lr=0.01, betas=(0.9, 0.99), guard=(None if flag is None else (lambda: flag)))
        for _ in range(2):
            for p, g in zip(ps, grads):
                p.grad = g.clone().cuda()
            opt.step()
        torch.cuda.synchronize()
        return ps, opt

    plain, _ = run(None)
    open_, _ = run(0.0)
    shut, opt = run(2.0)
    for a, b, c, b0 in zip(plain, open_, shut, base):
        assert torch.equal(a, b)
        assert torch.equal(c.detach().cpu(), b0)
        assert float(opt.state[c]["exp_avg"].abs().max()) == 0.0 and float(opt.state[c]["exp_avg_sq"].abs().max()) == 0.0
    assert not torch.equal(plain[0].detach().cpu(), base[0])
    with pytest.raises(RuntimeError):
        ft.FusedAdam([torch.nn.Parameter(base[0].clone().cuda())], guard=lambda: torch.zeros(2, device="cuda"))._guard_tensor()


def test_overflow_flag_is_the_gradient_buffers_tail(gm):
    g, m = gm
    m.grad_buffer()
    f = m.overflow_flag()
    assert f.numel() == 1 and f.data_ptr() == m._gradbuf.data_ptr() + (m._gradbuf.numel() - 4) * 4
