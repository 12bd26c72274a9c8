"""`-m gpu` checks at BASELINE.json's full size (TensorVMSplit 300^3, 4096 rays x S=1039), where the CPU
oracle is too slow to be the checker: the shading back ends (fused tcgen05, per-layer tcgen05, fp32
CUDA cores) must agree with each other on outputs and on every gradient, and the packed sample lists
must be identical between training passes with the same jitter (idempotence)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _step(model, rays, target, path):
    from tests.gpu_util import shade_backend
    with shade_backend(path):
        torch.manual_seed(7)
        loss, nv = model.train_step(rays, target, None, white_bg=True, N_samples=1039)
        torch.cuda.synchronize()
        grads = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
        return float(loss), int(nv), grads


def _rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-20))


@pytest.fixture(scope="module")
def workload():
    import bench
    dev = torch.device("cuda", 0)
    model = bench.build_model(dev)
    rays, target = bench.synthetic_rays(bench.RAYS_PER_GPU, seed=0)
    return model, rays.to(dev), target.to(dev)


def _mlp_grads_fp64(model, st, internal):
    """MLP weight gradients recomputed in float64 from the activations the kernels saved."""
    import ctypes as C
    from fewtensorf_b200 import _abi
    n = st.n_valid
    sec = lambda name, cols: model.workspace_section(st, name, torch.float32, n * cols).view(n, cols).double()
    offs = (C.c_int64 * _abi.WS_COUNT)()
    _abi.lib.ftrf_workspace_layout(C.byref(st.field), C.byref(st.call), offs)
    ldx = (st.field.in_dim + 1 + 7) // 8 * 8
    X, h1, h2 = (model.saved_activation(st, nm, cc).double() for nm, cc in (("A_X", ldx), ("A_H1", 128), ("A_H2", 128)))
    rgb, grgb = sec("A_RGB", 3), sec("A_GRGB", 3)
    m = model.renderModule.mlp
    W2, W3 = m[2].weight.double(), m[4].weight.double()
    go = grgb * rgb * (1 - rgb)
    g_h2 = (h2 > 0) * (go @ W3)
    g_h1 = (h1 > 0) * (g_h2 @ W2)
    out = {"renderModule.mlp.4.weight": go.t() @ h2, "renderModule.mlp.4.bias": go.sum(0),
           "renderModule.mlp.2.weight": g_h2.t() @ h1, "renderModule.mlp.2.bias": g_h2.sum(0),
           "renderModule.mlp.0.bias": g_h1.sum(0)}
    dW1 = g_h1.t() @ X
    if internal is not None:          # X columns are in the fused kernels' internal order
        full = torch.zeros(128, st.field.in_dim, dtype=torch.float64, device=X.device)
        for k, o in enumerate(internal):
            if o >= 0:
                full[:, o] = dW1[:, k]
        dW1 = full
    else:
        dW1 = dW1[:, :st.field.in_dim]
    out["renderModule.mlp.0.weight"] = dW1
    # forward chain and input gradient, from the saved X (validates the forward / dgrad contractions)
    W1 = m[0].weight.double()
    if internal is not None:
        W1i = torch.zeros(128, len(internal), dtype=torch.float64, device=X.device)
        for k, o in enumerate(internal):
            if o >= 0:
                W1i[:, k] = W1[:, o]
        Xk = X[:, :len(internal)]
    else:
        W1i, Xk = W1, X[:, :st.field.in_dim]
    h1_ref = torch.relu(Xk @ W1i.t() + m[0].bias.double())
    h2_ref = torch.relu(h1_ref @ W2.t() + m[2].bias.double())
    rgb_ref = torch.sigmoid(h2_ref @ W3.t() + m[4].bias.double())
    gx_ref = g_h1 @ W1i
    chain = {"h1": (h1, h1_ref), "h2": (h2, h2_ref), "rgb": (rgb, rgb_ref), "g_h1": (model.saved_activation(st, "A_GH1", 128).double(), g_h1)}
    if internal is None:
        chain["g_x"] = (sec("A_GX", ldx)[:, :gx_ref.shape[1]], gx_ref)
    else:
        # the fused dgrad kernel keeps g_X on chip: check what it hands on, g_feat, through the chain rule of
        # [feat | view | PE(feat) | PE(view)] (models/mlp.py:41-64; the bench field has no frequency masks)
        D, F = 27, 2
        feat = sec("A_FEAT", D)
        g_feat = gx_ref[:, :D].clone()
        c0 = D + 3 + ((D + 3) & 1)
        for ch in range(D):
            for fq in range(F):
                k = c0 + 2 * (ch * F + fq)
                a = feat[:, ch] * 2.0 ** fq
                g_feat[:, ch] += 2.0 ** fq * (torch.cos(a) * gx_ref[:, k] - torch.sin(a) * gx_ref[:, k + 1])
        chain["g_feat"] = (sec("A_GFEAT", D), g_feat)
    return out, chain


def _internal_order(D=27, fea_pe=2, view_pe=2):
    """fused.cuh: [feat | view | (pad) | (sin, cos) pairs of PE(fea) | pairs of PE(view) | pad to 8] for MLP_Fea."""
    orig = list(range(D + 3))
    if len(orig) & 1:
        orig.append(-1)
    col0 = D + 3
    for C, F in ((D, fea_pe), (3, view_pe)):
        half = C * F
        for q in range(half):
            orig += [col0 + q, col0 + half + q]
        col0 += 2 * half
    while len(orig) % 8:
        orig.append(-1)
    return orig


@pytest.mark.parametrize("path", [2, 1])
def test_mlp_weight_gradients_vs_float64_full_size(workload, path):
    """102 k shaded samples reduce into each MLP weight gradient: compare with a float64 contraction of
    the very activations the kernels saved (isolates the tensor-core wgrad / dgrad chain)."""
    model, rays, target = workload
    loss, nv, g = _step(model, rays, target, path)
    ref, chain = _mlp_grads_fp64(model, model._last_call, _internal_order() if path == 2 else None)
    for k, (got, want) in chain.items():
        assert _rel(got, want) < 5e-6, (k, _rel(got, want))
    for k, r in ref.items():
        assert _rel(g[k].double(), r) < 2e-5, (k, _rel(g[k].double(), r))


def test_shading_paths_agree_full_size(workload):
    model, rays, target = workload
    l2, n2, g2 = _step(model, rays, target, 2)
    l0, n0, g0 = _step(model, rays, target, 0)      # fp32 CUDA cores: exact fp32 products
    l1, n1, g1 = _step(model, rays, target, 1)
    assert n2 == n0 == n1 and n2 > 50000
    assert l2 == pytest.approx(l0, rel=1e-5) and l1 == pytest.approx(l0, rel=1e-5)
    # Back ends round differently, so among the 26 M hidden activations of a step a few sit within
    # rounding of the ReLU kink and flip between back ends; each flip changes dL/dfeat of that sample by
    # O(1) (measured: max |d g_feat| ~ 1e-2 of the largest entry, every other sample agreeing to 1e-6).
    # The gradients are therefore compared in the L2 norm, with a loose cap on single entries; the
    # accuracy of each back end on its own activations is pinned by the float64 test above.
    l2 = lambda a, b: float((a - b).norm() / b.norm().clamp_min(1e-30))
    for k in g0:
        for x, y in ((g2, g0), (g1, g0), (g2, g1)):
            assert l2(x[k], y[k]) < 3e-3, (k, l2(x[k], y[k]))
            assert _rel(x[k], y[k]) < 5e-2, (k, _rel(x[k], y[k]))


def test_eval_render_matches_training_forward_without_jitter(workload):
    """eval forward (no autograd, no saved activations) and the saving forward give the same image."""
    model, rays, target = workload
    # the first eval call overflows its guessed list capacities and is repeated: the overflowed pass must
    # stay inside its workspace even when that memory holds garbage (regression: unwritten list holes)
    model._ws_eval.buf = None
    model._cap_hint.clear()
    junk = torch.full((96 << 20,), 0x7f7f7f7f, dtype=torch.int32, device=rays.device)
    del junk
    with torch.no_grad():
        rgb_e, depth_e, nv_e = model(rays, None, white_bg=True, is_train=False, N_samples=1039)
    rgb_t, depth_t, nv_t = model(rays, None, white_bg=True, is_train=False, N_samples=1039)   # grad enabled: saves
    assert int(nv_e) == int(nv_t)
    assert torch.equal(depth_e, depth_t)
    assert _rel(rgb_t.detach(), rgb_e) < 1e-6


def test_coarse_occupancy_keeps_sample_lists_bit_exact(workload):
    """The march kernel's coarse-block shortcut (approximate texel coordinate + dilated 8^3 occupancy) must
    not change a single keep decision: packed (ray, index) lists with and without it are the same set."""
    import bench
    model, rays, target = workload
    cam = bench.camera_rays(rays.device)[::7][:60000].contiguous()      # image-wide subset of a test view
    keys = []
    for use in (True, False):
        model.use_coarse_mask = use
        try:
            with torch.no_grad():
                model(cam, None, white_bg=True, is_train=False, N_samples=-1)
        finally:
            model.use_coarse_mask = True
        st = model._last_call
        # the call really ran with / without the coarse levels (the field descriptor is cached per setting)
        assert bool(st.field.am_coarse) == use
        n = st.n_samples
        ray = model.workspace_section(st, "S_RAY", torch.int32, n).long()
        idx = model.workspace_section(st, "S_IDX", torch.int32, n).long()
        cnt = model.workspace_section(st, "RAY_CNT", torch.int32, cam.shape[0]).clone()
        keys.append((torch.sort(ray * 4096 + idx).values.clone(), cnt, n))
    assert keys[0][2] == keys[1][2] and keys[0][2] > 100000
    assert torch.equal(keys[0][1], keys[1][1])
    assert torch.equal(keys[0][0], keys[1][0])
