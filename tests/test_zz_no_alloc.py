"""`ftrf_shade_points_static` — the allocation-free twin of `ftrf_shade_points` (include/fewtensorf.h; no counterpart in
the reference): same kernels, scratch in a file-scope `static __device__`
array of FTRF_STATIC_SCRATCH_BYTES inside the library.  Called through the C ABI and compared with the existing entry
point on calls in a row that share the array (different models, widths and point counts, so a stale weight image or
counter would show), and with the oracle.  Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

from tests.util import rel_inf

TOL = 1e-4


def _field(shading, pe, seed):
    from oracle import field_oracle as O
    cfg = O.FieldConfig(model="VM", shading=shading, grid=(40, 48, 56), density_n_comp=(16, 8, 24), app_n_comp=(16, 8, 24),
                        pos_pe=pe[0], view_pe=pe[1], fea_pe=pe[2])
    p = O.init_params(cfg, seed=seed)
    O.plant_box(cfg, p)
    return O.Field(cfg, p)


def _inputs(n, seed):
    g = torch.Generator().manual_seed(seed)
    xn = torch.rand(n, 3, generator=g) * 2 - 1
    view = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=-1)
    feat = torch.randn(n, 27, generator=g) * 0.7
    return xn, view, feat


def _both(m, xn, view, feat):
    """(rgb of ftrf_shade_points with a torch-allocated scratch, rgb of ftrf_shade_points_static), same stream."""
    from fewtensorf_b200._abi import check, lib
    from fewtensorf_b200.module import _points_workspace, _stream
    n = xn.shape[0]
    with torch.cuda.device(xn.device):
        field = m._fill_field(None)
        ws = _points_workspace(field, n, xn.device)
        a = torch.full((n, 3), -7.0, device=xn.device)
        b = torch.full((n, 3), -9.0, device=xn.device)
        check(lib.ftrf_shade_points(C.byref(field), xn.data_ptr(), view.data_ptr(), feat.data_ptr(), n, a.data_ptr(),
                                    ws.data_ptr(), ws.numel(), _stream()), "shade_points")
        check(lib.ftrf_shade_points_static(C.byref(field), xn.data_ptr(), view.data_ptr(), feat.data_ptr(), n, b.data_ptr(),
                                           _stream()), "shade_points_static")
        torch.cuda.synchronize()
    return a, b


@pytest.mark.gpu
def test_static_scratch_matches_the_allocating_entry_point_on_calls_in_a_row():
    from oracle import field_oracle as O
    from fewtensorf_b200._abi import STATIC_MAX_POINTS
    from tests.gpu_util import model_from_field
    # (shading, (pos_pe, view_pe, fea_pe), points): fused tcgen05 forward at the limit, the 153-column few-shot shape,
    # SH (counters only), a shape on the dynamic-spec build, and the first model again with a handful of points
    calls = [("MLP_Fea", (2, 2, 2), STATIC_MAX_POINTS), ("MLP", (8, 2, 2), 1777), ("SH", (2, 2, 2), 4099),
             ("MLP_Fea", (2, 4, 1), 3000), ("MLP_Fea", (2, 2, 2), 5)]
    for k, (shading, pe, n) in enumerate(calls):
        f = _field(shading, pe, seed=3 + k)
        m = model_from_field(f)
        xn, view, feat = _inputs(n, 40 + k)
        a, b = _both(m, xn.cuda(), view.cuda(), feat.cuda())
        want = O.shade(f, xn, view, feat)
        assert rel_inf(a.cpu(), want) < TOL, (shading, pe, n)
        assert rel_inf(b.cpu(), want) < TOL, (shading, pe, n)
        assert torch.equal(a, b), (shading, pe, n, float((a - b).abs().max()))       # same kernels, same inputs
    # the same model twice in a row, second call on fewer points: rows beyond n stay untouched, the result is the first n rows'
    f = _field("MLP_Fea", (2, 2, 2), seed=11)
    m = model_from_field(f)
    xn, view, feat = (t.cuda() for t in _inputs(2048, 77))
    a1, b1 = _both(m, xn, view, feat)
    a2, b2 = _both(m, xn[:100].contiguous(), view[:100].contiguous(), feat[:100].contiguous())
    assert torch.equal(a1, b1) and torch.equal(a2, b2) and torch.equal(b2, b1[:100])


@pytest.mark.gpu
def test_static_entry_point_allocates_nothing():
    """Device memory in use (driver's view and torch's) is the same before and after calls through the static entry
    point, once the kernels' modules are loaded by a first call."""
    from fewtensorf_b200._abi import check, lib
    from fewtensorf_b200.module import _stream
    from tests.gpu_util import model_from_field
    m = model_from_field(_field("MLP_Fea", (2, 2, 2), seed=5))
    xn, view, feat = (t.cuda() for t in _inputs(4096, 9))
    out = torch.empty(4096, 3, device="cuda")
    field = m._fill_field(None)

    def call():
        check(lib.ftrf_shade_points_static(C.byref(field), xn.data_ptr(), view.data_ptr(), feat.data_ptr(), 4096,
                                           out.data_ptr(), _stream()), "shade_points_static")
    call()
    torch.cuda.synchronize()
    free0, torch0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_reserved()
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    assert torch.cuda.memory_reserved() == torch0
    assert torch.cuda.mem_get_info()[0] >= free0


def test_static_entry_point_rejects_sizes_beyond_its_limits():
    """Argument checks run before any CUDA call (host only): too many points, and a field whose scratch would not fit."""
    from fewtensorf_b200 import _abi
    lib = _abi.lib
    f = _abi.FtrfField()
    f.model, f.shading, f.app_dim, f.feature_c, f.in_dim = 0, 0, 27, 128, 150           # the benchmark's MLP_Fea shape
    f.pos_pe, f.view_pe, f.fea_pe = 2, 2, 2
    for i in range(3):
        f.den.n_comp[i], f.app.n_comp[i] = 16, 48
    assert lib.ftrf_points_workspace_bytes(C.byref(f), _abi.STATIC_MAX_POINTS) <= _abi.STATIC_SCRATCH_BYTES
    rc = lib.ftrf_shade_points_static(C.byref(f), None, None, None, _abi.STATIC_MAX_POINTS + 1, None, None)
    assert rc == -1 and b"exceed the static scratch" in lib.ftrf_last_error()
    f.shading, f.pos_pe, f.in_dim = 2, 8, 198              # wider rows: the same point count no longer fits
    assert lib.ftrf_points_workspace_bytes(C.byref(f), _abi.STATIC_MAX_POINTS) > _abi.STATIC_SCRATCH_BYTES
    rc = lib.ftrf_shade_points_static(C.byref(f), None, None, None, _abi.STATIC_MAX_POINTS, None, None)
    assert rc == -2 and b"the static array holds" in lib.ftrf_last_error()
    rc = lib.ftrf_shade_points_static(C.byref(f), None, None, None, -1, None, None)
    assert rc == -1
    rc = lib.ftrf_shade_points_static(None, None, None, None, 8, None, None)
    assert rc == -1
