"""Helpers for the `-m gpu` parity tests: build the product model from a golden fixture / oracle field."""
import contextlib

import torch

import fewtensorf_b200 as ft
from oracle import field_oracle as O

_TEST_LIB = None


def test_lib():
    """libfewtensorf_b200_test.so: the superset build with the cross-check back ends, unit-test entry points and
    debug switches (the product library has none of them)."""
    global _TEST_LIB
    if _TEST_LIB is None:
        from fewtensorf_b200 import _abi
        _TEST_LIB = _abi.load_test_lib()
    return _TEST_LIB


test_lib.__test__ = False       # not a pytest test


@contextlib.contextmanager
def through_test_lib(setup=None, teardown=None):
    """Route the module's C calls through the test build for the duration (same ABI, same structs), optionally
    flipping one of its switches: `setup(lib)` on entry, `teardown(lib)` on exit."""
    from fewtensorf_b200 import module
    tl = test_lib()
    old = module.lib
    module.lib = tl
    if setup is not None:
        setup(tl)
    try:
        yield tl
    finally:
        if teardown is not None:
            teardown(tl)
        module.lib = old


def shade_backend(path):
    """0 = fp32 CUDA-core MLP kernels, 1 = one tcgen05 GEMM per layer, 2 = the fused tcgen05 kernels (product default)."""
    return through_test_lib(lambda l: l.ftrf_set_shade_path(path), lambda l: l.ftrf_set_shade_path(2))


def args_from_cfg(cfg: O.FieldConfig):
    return {'step_ratio': cfg.step_ratio, 'fea2denseAct': cfg.fea2dense, 'density_n_comp': list(cfg.density_n_comp),
            'app_n_comp': list(cfg.app_n_comp), 'app_dim': cfg.app_dim, 'density_shift': cfg.density_shift,
            'distance_scale': cfg.distance_scale, 'alphaMask_thres': cfg.alpha_mask_thres, 'shadingMode': cfg.shading,
            'pos_pe': cfg.pos_pe, 'view_pe': cfg.view_pe, 'fea_pe': cfg.fea_pe, 'featureC': cfg.feature_c}


def model_from_field(f: O.Field, device="cuda"):
    """Product module holding exactly the oracle field's parameters / alphaMask."""
    cfg = f.cfg
    aabb = torch.tensor(cfg.aabb, dtype=torch.float32, device=device)
    if cfg.model == "VM":
        m = ft.TensorVMSplit(args_from_cfg(cfg), aabb, list(cfg.grid), list(cfg.near_far), device)
    else:
        m = ft.TensorCP(args_from_cfg(cfg), aabb, list(cfg.grid), device, near_far=list(cfg.near_far))
    m.rayMarch_weight_thres = cfg.weight_thres
    missing = m.load_state_dict({k: v for k, v in f.params.items()}, strict=True)
    if f.alpha_mask is not None:
        m.alphaMask = ft.AlphaGridMask(device, f.alpha_mask.aabb.to(device), f.alpha_mask.volume.to(device))
    return m


def pin_rng(model, jitter, coin, white_bg):
    """Feed the model the exact RNG draws the reference / oracle consumed."""
    def fixed(n_rays, S, is_train, ndc_ray, wb, device):
        j = None
        if ndc_ray:
            near, far = model.near_far
            z = torch.linspace(near, far, S).unsqueeze(0)
            if is_train:
                z = z + jitter * ((far - near) / S)
            j = z.reshape(-1).to(device)
        elif is_train:
            j = jitter.reshape(-1).to(device)
        return j, bool(wb or (is_train and coin))
    model._draw_rng = fixed


def dense_from_packed(model, st, n_rays, S):
    """Scatter the packed per-sample sections of a finished call back to dense [N,S] tensors."""
    V = st.n_samples
    ray = model.workspace_section(st, "S_RAY", torch.int32, V).long().cpu()
    idx = model.workspace_section(st, "S_IDX", torch.int32, V).long().cpu()
    out = {"valid": torch.zeros(n_rays, S, dtype=torch.bool)}
    out["valid"][ray, idx] = True
    for name, key in (("S_SIGMA", "sigma"), ("S_W", "weight")):
        d = torch.zeros(n_rays, S)
        d[ray, idx] = model.workspace_section(st, name, torch.float32, V).cpu()
        out[key] = d
    app = model.workspace_section(st, "S_APP", torch.int32, V).cpu()
    a = torch.zeros(n_rays, S, dtype=torch.bool)
    a[ray, idx] = app >= 0
    out["app"] = a
    out["tmin"] = model.workspace_section(st, "TMIN", torch.float32, n_rays).cpu()
    out["ray"], out["idx"] = ray, idx
    return out


def relu_decisions_in_oracle_order(model, st, app_mask):
    """The shading MLP's unit on/off decisions of a finished (saving) call, reordered to the oracle's sample order
    (row-major over the [N,S] shaded mask): (m1, m2) boolean [n,128] CPU tensors for `oracle.shade(relu_masks=...)`.
    The kernels' shaded list is compacted ray by ray in launch order, not sorted; each entry knows its (ray, index)."""
    n = st.n_valid
    V = st.n_samples
    slot = model.workspace_section(st, "A_SLOT", torch.int32, n).long()
    ray = model.workspace_section(st, "S_RAY", torch.int32, V).long()[slot].cpu()
    idx = model.workspace_section(st, "S_IDX", torch.int32, V).long()[slot].cpu()
    pos = (torch.cumsum(app_mask.reshape(-1).long(), 0) - 1).reshape(app_mask.shape)[ray, idx]
    assert bool(app_mask[ray, idx].all()) and int(app_mask.sum()) == n, "shaded sets differ: decisions cannot be aligned"
    out = []
    for name in ("A_H1", "A_H2"):
        h = model.saved_activation(st, name, 128).cpu()
        m = torch.zeros(n, 128, dtype=torch.bool)
        m[pos] = h > 0
        out.append(m)
    return out[0], out[1]


def flip_aware_reference(f, loss_and_grads_kw, keep, masks):
    """(n_flips, worst |z| at a flip, grads) — if the GPU's ReLU decisions `masks` differ from the oracle's anywhere,
    the oracle is re-evaluated WITH those decisions (every differing pre-activation must be roundoff-sized) and its
    gradients are returned; otherwise (0, 0.0, None)."""
    d1, d2 = (keep["z1"] > 0) != masks[0], (keep["z2"] > 0) != masks[1]
    n_flips = int(d1.sum()) + int(d2.sum())
    if n_flips == 0:
        return 0, 0.0, None
    worst = max(float(keep["z1"][d1].abs().max()) if d1.any() else 0.0, float(keep["z2"][d2].abs().max()) if d2.any() else 0.0)
    res = O.loss_and_grads(f, relu_masks=(masks[0].float(), masks[1].float()), **loss_and_grads_kw)
    return n_flips, worst, res
