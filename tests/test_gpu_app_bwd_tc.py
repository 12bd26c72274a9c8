"""`-m gpu`: the tensor-core appearance backward (app_bwd_vm_tc_kernel: GQ = G Wb and dWb += G^T P as 3xTF32 tcgen05
batches) against the CUDA-core kernel it replaces, on the same call, at the benchmark's component counts (3 x 48)
and at shapes around its tile / operand granularities.  Summation order differs (TMEM accumulation over the CTA's
tiles vs register tiles): 2e-5 relative; the reference-generated fixtures of test_gpu_golden.py pin either to the
reference."""
import pytest
import torch

from oracle import field_oracle as O
from tests.util import rel_inf

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model,grid,comps,n_rays,masked", [("VM", (64, 64, 64), (48, 48, 48), 2048, False),
                                                            ("VM", (40, 48, 56), (16, 32, 48), 700, True),
                                                            ("VM", (32, 32, 32), (8, 8, 16), 333, False),
                                                            ("VM", (48, 48, 48), (64, 64, 64), 1500, False),
                                                            ("CP", (64, 64, 64), (288, 288, 288), 2048, True),
                                                            ("CP", (40, 48, 56), (96, 96, 96), 700, False),
                                                            ("CP", (32, 32, 32), (48, 48, 48), 333, False),
                                                            ("CP", (48, 48, 48), (160, 160, 160), 1500, True)])
def test_app_backward_tensor_core_matches_cuda_core(model, grid, comps, n_rays, masked):
    """(TensorCP: the same kernel over component blocks of <= 144, three line samples per product.)"""
    from tests.gpu_util import model_from_field, through_test_lib
    cfg = O.FieldConfig(model=model, grid=grid, density_n_comp=(16, 16, 16), app_n_comp=comps)
    p = O.init_params(cfg, seed=31)
    O.plant_box(cfg, p)
    m = model_from_field(O.Field(cfg, p), "cuda")
    m.updateAlphaMask((32, 32, 32))
    rays = O.synthetic_rays(n_rays, seed=32).cuda()
    target = torch.rand(n_rays, 3, generator=torch.Generator().manual_seed(33)).cuda()
    mask = None
    if masked:
        g = torch.Generator().manual_seed(34)
        mask = {"encoding": {"pos": None, "view": None, "fea": None},
                "decomp": {"den": None, "app": [torch.rand(cc, generator=g).cuda() for cc in (comps if model == "VM" else comps[:1])]}}
    grads = {}
    for tcore in (1, 0):
        with through_test_lib(lambda l: l.ftrf_debug_set_app_bwd_tc(tcore), lambda l: l.ftrf_debug_set_app_bwd_tc(1)):
            torch.manual_seed(5)
            loss, nv = m.train_step(rays, target, mask, white_bg=True, N_samples=200)
            torch.cuda.synchronize()
        grads[tcore] = {k: q.grad.detach().clone() for k, q in m.named_parameters()
                        if k.startswith(("app_plane", "app_line", "basis_mat"))}
        assert nv > 1000
    for k in grads[1]:
        assert float(grads[0][k].abs().max()) > 0, k
        assert rel_inf(grads[1][k], grads[0][k]) < 2e-5, k


@pytest.mark.parametrize("model,comps,n,masked", [("VM", (48, 48, 48), 128 * 300 + 77, False), ("VM", (16, 32, 48), 5000, True),
                                                  ("VM", (8, 8, 16), 129, False), ("VM", (64, 64, 64), 128 * 149, False),
                                                  ("CP", (288, 288, 288), 128 * 150 + 5, True), ("CP", (96, 96, 96), 4000, False),
                                                  ("CP", (72, 72, 72), 700, False), ("CP", (40, 40, 40), 128, True)])
def test_app_forward_builds_agree(model, comps, n, masked):
    """The three forward builds of the VM appearance stage — warp-specialised tcgen05 (shipped), two-CTA tcgen05, CUDA
    cores — on the same points, against each other and the oracle; tile counts above and below one wave of CTAs."""
    from tests.gpu_util import model_from_field, through_test_lib
    cfg = O.FieldConfig(model=model, grid=(40, 48, 56), density_n_comp=(16, 16, 16), app_n_comp=comps)
    p = O.init_params(cfg, seed=41)
    f = O.Field(cfg, p)
    m = model_from_field(f, "cuda")
    g = torch.Generator().manual_seed(42)
    xn = torch.rand(n, 3, generator=g) * 2.1 - 1.05
    mask = [torch.rand(c, generator=g) for c in (comps if model == "VM" else comps[:1])] if masked else None
    want = O.app_feature(f, xn, mask)
    mk = None if mask is None else [t.cuda() for t in mask]
    got = {}
    got["ws"] = m.compute_appfeature(xn.cuda(), mk)
    with through_test_lib(lambda l: l.ftrf_debug_set_app_ws(0), lambda l: l.ftrf_debug_set_app_ws(1)):
        got["tc"] = m.compute_appfeature(xn.cuda(), mk)
    with through_test_lib(lambda l: l.ftrf_debug_set_app_tc(0), lambda l: l.ftrf_debug_set_app_tc(1)):
        got["cuda_cores"] = m.compute_appfeature(xn.cuda(), mk)
    for k, v in got.items():
        assert rel_inf(v.cpu(), want) < 1e-4, k
    if model == "VM":
        assert torch.equal(got["ws"], got["tc"])                  # same products in the same order
    assert rel_inf(got["ws"], got["cuda_cores"]) < 2e-5
