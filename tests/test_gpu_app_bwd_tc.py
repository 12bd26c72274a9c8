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


@pytest.mark.parametrize("grid,comps,n_rays,masked", [((64, 64, 64), (48, 48, 48), 2048, False),
                                                      ((40, 48, 56), (16, 32, 48), 700, True),
                                                      ((32, 32, 32), (8, 8, 16), 333, False),
                                                      ((48, 48, 48), (64, 64, 64), 1500, False)])
def test_app_backward_tensor_core_matches_cuda_core(grid, comps, n_rays, masked):
    from tests.gpu_util import model_from_field, through_test_lib
    cfg = O.FieldConfig(grid=grid, density_n_comp=(16, 16, 16), app_n_comp=comps)
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
                "decomp": {"den": None, "app": [torch.rand(cc, generator=g).cuda() for cc in comps]}}
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
