"""`-m gpu`: the module methods SURVEY section 8(b) lists beside forward() — compute_densityfeature(xn, mask),
compute_appfeature(xn, mask), renderModule(pts, viewdirs, features, mask), SHRender, sample_ray_ndc — against the
oracle functions that restate them (oracle.density_feature / app_feature / shade follow models/tensoRF.py:207-263,
:358-415 and models/mlp.py:15-155), plus the re-entrancy the boundary promises: two models driven from two host
threads at once give the gradients they give alone.  Tolerance 1e-4 relative (‖Δ‖∞ / ‖ref‖∞) as everywhere."""
import threading

import pytest
import torch

from oracle import field_oracle as O
from tests.util import rel_inf

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _field(model="VM", shading="MLP_Fea", grid=(40, 48, 56), den=(16, 8, 24), app=(16, 8, 24), seed=3, **kw):
    if model == "CP":
        den, app = (den[0],) * 3, (app[0],) * 3
    cfg = O.FieldConfig(model=model, shading=shading, grid=grid, density_n_comp=den, app_n_comp=app, **kw)
    p = O.init_params(cfg, seed=seed)
    O.plant_box(cfg, p)
    return O.Field(cfg, p)


def _points(n, seed):
    g = torch.Generator().manual_seed(seed)
    xn = torch.rand(n, 3, generator=g) * 2.2 - 1.1           # some outside [-1, 1]: zero-padding taps
    xn[:5] = torch.tensor([[-1.0, -1.0, -1.0], [1.0, 1.0, 1.0], [0.0, 0.0, 0.0], [1.0, -1.0, 0.5], [-1.2, 0.3, 1.3]])
    return xn


@pytest.mark.parametrize("model,comps,masked", [("VM", (16, 8, 24), False), ("VM", (16, 8, 24), True), ("CP", (24, 24, 24), False),
                                                ("CP", (96, 96, 96), True)])
def test_compute_densityfeature_matches_oracle(model, comps, masked):
    from tests.gpu_util import model_from_field
    f = _field(model, den=comps)
    m = model_from_field(f)
    xn = _points(3001, 11)
    mask = None
    if masked:
        g = torch.Generator().manual_seed(5)
        mask = [torch.rand(c, generator=g) for c in (f.cfg.density_n_comp if model == "VM" else f.cfg.density_n_comp[:1])]
    want = O.density_feature(f, xn, mask)
    got = m.compute_densityfeature(xn.cuda(), None if mask is None else [t.cuda() for t in mask])
    assert got.shape == want.shape
    assert rel_inf(got.cpu(), want) < TOL


@pytest.mark.parametrize("model,comps,masked", [("VM", (48, 48, 48), False), ("VM", (16, 8, 24), True), ("VM", (12, 20, 4), False),
                                                ("CP", (48, 48, 48), False), ("CP", (288, 288, 288), True)])
def test_compute_appfeature_matches_oracle(model, comps, masked):
    """VM with n_comp % 8 == 0 runs app_fwd_tc_kernel (tcgen05), the other shapes the generic kernel."""
    from tests.gpu_util import model_from_field
    f = _field(model, app=comps)
    m = model_from_field(f)
    xn = _points(2500, 12)
    mask = None
    if masked:
        g = torch.Generator().manual_seed(6)
        mask = [torch.rand(c, generator=g) for c in (f.cfg.app_n_comp if model == "VM" else f.cfg.app_n_comp[:1])]
    want = O.app_feature(f, xn, mask)
    got = m.compute_appfeature(xn.cuda(), None if mask is None else [t.cuda() for t in mask])
    assert got.shape == want.shape == (2500, 27)
    assert rel_inf(got.cpu(), want) < TOL


@pytest.mark.parametrize("shading,pe,masked", [("MLP_Fea", (2, 2, 2), False), ("MLP_Fea", (2, 2, 2), True), ("MLP_PE", (2, 2, 2), True),
                                               ("MLP", (2, 2, 2), False), ("MLP", (8, 2, 2), True), ("MLP_Fea", (2, 4, 1), False)])
def test_render_module_call_matches_oracle(shading, pe, masked):
    """renderModule(pts, viewdirs, features, mask) — mlp.py:41-69, :88-107, :127-155 — on the fused tcgen05 forward
    (per-layer tcgen05 GEMMs for the widths it does not cover)."""
    from tests.gpu_util import model_from_field
    f = _field(shading=shading, pos_pe=pe[0], view_pe=pe[1], fea_pe=pe[2])
    m = model_from_field(f)
    n = 1777
    g = torch.Generator().manual_seed(21)
    xn = torch.rand(n, 3, generator=g) * 2 - 1
    view = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=-1)
    feat = torch.randn(n, 27, generator=g) * 0.7
    enc = None
    if masked:
        enc = {"pos": torch.rand(2 * pe[0] * 3, generator=g), "view": torch.rand((), generator=g),
               "fea": torch.rand(2 * pe[2] * 27, generator=g)}
    want = O.shade(f, xn, view, feat, enc)
    got = m.renderModule(xn.cuda(), view.cuda(), feat.cuda(), None if enc is None else {k: v.cuda() for k, v in enc.items()})
    assert got.shape == (n, 3)
    assert rel_inf(got.cpu(), want) < TOL


def test_sh_render_function_matches_oracle():
    import fewtensorf_b200 as ft
    f = _field(shading="SH")
    n = 4099
    g = torch.Generator().manual_seed(22)
    view = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=-1)
    feat = torch.randn(n, 27, generator=g)
    want = O.shade(f, torch.zeros(n, 3), view, feat)
    got = ft.SHRender(torch.zeros(n, 3).cuda(), view.cuda(), feat.cuda())
    assert rel_inf(got.cpu(), want) < 1e-5
    from tests.gpu_util import model_from_field
    m = model_from_field(f)
    assert m.renderModule is ft.SHRender


def test_sample_ray_ndc_follows_the_reference_ops():
    """tensorBase.py:178-187: same points / z / in-box mask as the oracle's march_ndc without jitter; with jitter the
    draw comes from the rays' device generator (as in the reference: rand_like of a tensor already moved `.to(rays_o)`)."""
    from tests.gpu_util import model_from_field
    f = _field()
    m = model_from_field(f)
    rays = O.synthetic_rays(64, seed=4)
    pts, z, ok = m.sample_ray_ndc(rays[:, :3].cuda(), rays[:, 3:6].cuda(), is_train=False, N_samples=33)
    pts_o, z_o, ok_o = O.march_ndc(f, rays[:, :3], rays[:, 3:6], None, 33)
    assert torch.equal(z.cpu(), z_o) and torch.equal(ok.cpu(), ok_o)
    assert float((pts.cpu() - pts_o).abs().max()) < 1e-6          # (mul + add may contract to an fma on the GPU)
    pts_t, z_t, ok_t = m.sample_ray_ndc(rays[:, :3].cuda(), rays[:, 3:6].cuda(), is_train=True, N_samples=33)
    near, far = m.near_far
    assert z_t.shape == (1, 33) and bool(((z_t.cpu() - z_o) >= 0).all()) and bool(((z_t.cpu() - z_o) < (far - near) / 33 + 1e-6).all())


def test_two_models_from_two_threads_interleave():
    """No process-wide call state: thread A trains a VM/MLP_Fea model through autograd while thread B trains a
    different one through train_step, 12 steps each; every step's gradients equal the ones each model produces
    alone with the same RNG draws."""
    from tests.gpu_util import model_from_field, pin_rng
    fa = _field(grid=(48, 48, 48), den=(16, 16, 16), app=(48, 48, 48), seed=41)
    fb = _field(shading="MLP_PE", grid=(40, 40, 40), den=(8, 8, 8), app=(16, 16, 16), seed=42)
    O.build_alpha_mask(fa, (32, 32, 32))
    ma, mb = model_from_field(fa), model_from_field(fb)
    rays_a, rays_b = O.synthetic_rays(1500, seed=43).cuda(), O.synthetic_rays(900, seed=44).cuda()
    tg_a, tg_b = torch.rand(1500, 3, generator=torch.Generator().manual_seed(45)).cuda(), torch.rand(900, 3, generator=torch.Generator().manual_seed(46)).cuda()
    ja, jb = torch.rand(1500, 1, generator=torch.Generator().manual_seed(47)), torch.rand(900, 1, generator=torch.Generator().manual_seed(48))
    pin_rng(ma, ja, False, True)
    pin_rng(mb, jb, False, True)

    def step_a():
        for p in ma.parameters():
            p.grad = None
        rgb, _, _ = ma(rays_a, None, white_bg=True, is_train=True, N_samples=150)
        torch.mean((rgb - tg_a) ** 2).backward()
        torch.cuda.current_stream().synchronize()
        return {k: p.grad.detach().clone() for k, p in ma.named_parameters()}

    def step_b():
        mb.train_step(rays_b, tg_b, None, white_bg=True, N_samples=120)
        torch.cuda.current_stream().synchronize()
        return {k: p.grad.detach().clone() for k, p in mb.named_parameters()}

    ref_a, ref_b = step_a(), step_b()
    assert max(float(v.abs().max()) for v in ref_a.values()) > 0
    out, errs = {"a": [], "b": []}, []

    def worker(name, fn):
        try:
            with torch.cuda.stream(torch.cuda.Stream()):
                for _ in range(12):
                    out[name].append(fn())
        except Exception as e:                      # noqa: BLE001
            errs.append((name, repr(e)))

    ta, tb = threading.Thread(target=worker, args=("a", step_a)), threading.Thread(target=worker, args=("b", step_b))
    ta.start(); tb.start(); ta.join(); tb.join()
    assert not errs, errs
    # atomics reorder sums from run to run: same bar as run-to-run noise of one model alone
    for name, ref in (("a", ref_a), ("b", ref_b)):
        assert len(out[name]) == 12
        for got in out[name]:
            for k in ref:
                assert rel_inf(got[k], ref[k]) < 2e-5, (name, k, rel_inf(got[k], ref[k]))


def test_split_backward_halves_from_different_threads():
    """ftrf_backward_part(1) and (2) issued by different host threads: what the first half left pending on its side
    stream travels in the FtrfCall, so the second half joins it (gradients equal the one-call backward's)."""
    from tests.gpu_util import model_from_field, pin_rng
    f = _field(grid=(48, 48, 48), den=(16, 16, 16), app=(48, 48, 48), seed=51)
    m = model_from_field(f)
    rays = O.synthetic_rays(1200, seed=52).cuda()
    target = torch.rand(1200, 3, generator=torch.Generator().manual_seed(53)).cuda()
    pin_rng(m, torch.rand(1200, 1, generator=torch.Generator().manual_seed(54)), False, True)
    m.train_step(rays, target, None, white_bg=True, N_samples=150)
    torch.cuda.synchronize()
    ref = {k: p.grad.detach().clone() for k, p in m.named_parameters()}

    buf, views = m.grad_buffer()
    jitter, add_bg = m._draw_rng(1200, 150, True, False, True, rays.device)
    st = m._run_forward(rays, jitter, None, 150, True, False, add_bg, save=True)
    rgb = st.keep[2]
    g_rgb = (2.0 / rgb.numel()) * (rgb - target)
    torch.cuda.synchronize()
    buf.zero_()
    stream = torch.cuda.current_stream()
    errs = []

    def half(part):
        try:
            with torch.cuda.stream(stream):
                m._run_backward(st, g_rgb, views, part=part)
        except Exception as e:                      # noqa: BLE001
            errs.append(repr(e))
    for part in (1, 2):
        t = threading.Thread(target=half, args=(part,))
        t.start(); t.join()
    torch.cuda.synchronize()
    assert not errs, errs
    name_of = {id(p): k for k, p in m.named_parameters()}
    for p, v in zip(m._all_params(), views):          # the views follow _all_params() order
        assert rel_inf(v, ref[name_of[id(p)]]) < 2e-5, name_of[id(p)]
