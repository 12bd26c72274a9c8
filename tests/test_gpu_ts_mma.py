"""`-m gpu`: the TS form of tcgen05.mma.kind::tf32 (A operand written to TMEM with tcgen05.st, B in shared memory) that
the fused shading forward uses — operand placement check against a plain matmul on tf32-representable inputs."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _tf32(t):
    return (t.view(torch.int32) & ~0x1FFF).view(torch.float32)


def test_ts_form_gemm_matches_matmul():
    from tests.gpu_util import test_lib
    lib = test_lib()
    lib.ftrf_test_ts_gemm.argtypes = [C.c_void_p] * 4
    g = torch.Generator().manual_seed(1)
    a = _tf32(torch.randn(128, 32, generator=g)).cuda()
    b = _tf32(torch.randn(128, 32, generator=g)).cuda()
    out = torch.zeros(128, 128, device="cuda")
    assert lib.ftrf_test_ts_gemm(a.data_ptr(), b.data_ptr(), out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    want = a.double() @ b.double().T
    assert float((out.double() - want).abs().max()) < 1e-5 * float(want.abs().max())


@pytest.mark.parametrize("shading,pe", [("MLP_Fea", (2, 2, 2)), ("MLP", (8, 2, 2)), ("MLP_Fea", (2, 4, 1))])
def test_tmem_operand_forward_matches_the_product_kernel_and_the_oracle(shading, pe):
    """shade_fwd_ts_kernel (test build only: the measured-and-rejected TMEM-operand forward) against the oracle and the
    shipped shared-memory-operand kernel, static and dynamic column recipes, with a tile count that is not a multiple of
    the grid."""
    from oracle import field_oracle as O
    from tests.gpu_util import model_from_field, through_test_lib
    from tests.util import rel_inf
    cfg = O.FieldConfig(model="VM", shading=shading, grid=(40, 48, 56), density_n_comp=(16, 8, 24), app_n_comp=(16, 8, 24),
                        pos_pe=pe[0], view_pe=pe[1], fea_pe=pe[2], app_dim=18 if pe[0] == 8 else 27)
    p = O.init_params(cfg, seed=4)
    f = O.Field(cfg, p)
    m = model_from_field(f)
    n = 128 * 151 + 37
    g = torch.Generator().manual_seed(8)
    xn = torch.rand(n, 3, generator=g) * 2 - 1
    view = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=-1)
    feat = torch.randn(n, cfg.app_dim, generator=g) * 0.7
    want = O.shade(f, xn, view, feat)
    ref = m.renderModule(xn.cuda(), view.cuda(), feat.cuda())
    with through_test_lib(lambda lib: lib.ftrf_debug_set_fwd_ts(1), lambda lib: lib.ftrf_debug_set_fwd_ts(0)):
        got = m.renderModule(xn.cuda(), view.cuda(), feat.cuda())
    assert rel_inf(ref.cpu(), want) < 1e-4
    assert rel_inf(got.cpu(), want) < 1e-4
    assert rel_inf(got.cpu(), ref.cpu()) < 2e-5


def test_wgrad_bulk_copy_ring_matches_direct_loads():
    """wgrad_fused_kernel<RAW> (activation rows brought in by cp.async.bulk, transposed from shared memory) against the
    build whose producers load from global memory: same sums in the same order, on a sample count with a ragged tail."""
    from tests.gpu_util import test_lib
    lib = test_lib()
    M, kp, ldx = 16 * 300 + 11, 152, 152
    g = torch.Generator().manual_seed(3)
    g2 = torch.randn(M, 128, generator=g).cuda()
    h1 = torch.randn(M, 128, generator=g).relu().cuda()
    g1 = torch.randn(M, 128, generator=g).cuda()
    x = torch.randn(M, ldx, generator=g).cuda()
    lib.ftrf_test_wgrad_fused.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_int, C.c_int64] + [C.c_void_p] * 2 + [C.c_int] + [C.c_void_p] * 3
    out = {}
    for raw in (1, 0):
        lib.ftrf_debug_set_wgrad_raw(raw)
        dw2 = torch.zeros(128, 128, device="cuda")
        dw1 = torch.zeros(128, kp, device="cuda")
        db1 = torch.zeros(128, device="cuda")
        scratch = torch.zeros(128 * 48 * 4, device="cuda")
        rc = lib.ftrf_test_wgrad_fused(g2.data_ptr(), h1.data_ptr(), g1.data_ptr(), x.data_ptr(), ldx, kp, M, dw2.data_ptr(),
                                       dw1.data_ptr(), kp, db1.data_ptr(), scratch.data_ptr(), None)
        assert rc == 0
        torch.cuda.synchronize()
        out[raw] = (dw2, dw1, db1)
    lib.ftrf_debug_set_wgrad_raw(1)
    want2 = g2.double().T @ h1.double()
    want1 = g1.double().T @ x[:, :kp].double()
    for raw in (1, 0):
        assert float((out[raw][0].double() - want2).abs().max()) < 2e-5 * float(want2.abs().max())
        assert float((out[raw][1].double() - want1).abs().max()) < 2e-5 * float(want1.abs().max())
        assert float((out[raw][2].double() - g1.double().sum(0)).abs().max()) < 2e-5 * float(g1.double().sum(0).abs().max())
    for a, b in zip(out[1], out[0]):
        assert float((a - b).abs().max()) < 1e-5 * float(b.abs().max())
