"""Diagnostic for tests/test_gpu_fullsize_oracle.py: per-parameter gradient error vs the oracle for each shading back
end, and how many ReLU decisions differ between the GPU's saved activations and an fp64 recomputation from its own
saved inputs (kink flips at |pre-activation| ~ 1e-7 are the only legitimate source of O(1/N) gradient differences)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import field_oracle as O  # noqa: E402
from tests.test_gpu_fullsize_oracle import CASES, N_RAYS, _build  # noqa: E402
from tests.gpu_util import pin_rng, shade_backend  # noqa: E402
from tests.util import rel_inf  # noqa: E402

case = sys.argv[1] if len(sys.argv) > 1 else "vm_128_s443"
m, f, rays, target, jitter, S, train = _build(case)
keep = {}
loss_o, rgb_o, depth_o, nv_o, grads_o = O.loss_and_grads(f, rays, target, is_train=train, white_bg=True, n_samples=S, jitter=jitter, keep=keep)
print(case, "oracle: loss", float(loss_o), "shaded", int(nv_o), "valid", int(keep["valid"].sum()))
for path in (2, 1, 0):
    try:
        with shade_backend(path):
            if train:
                pin_rng(m, jitter, False, True)
            for p in m.parameters():
                p.grad = None
            rgb, depth, nv = m(rays.cuda(), None, white_bg=True, is_train=train, N_samples=CASES[case][4])
            loss = torch.mean((rgb - target.cuda()) ** 2)
            loss.backward()
            torch.cuda.synchronize()
    except Exception as e:          # noqa: BLE001
        print("path", path, "failed:", repr(e)[:200])
        continue
    errs = {k: rel_inf(p.grad.cpu(), grads_o[k]) for k, p in m.named_parameters()}
    l2 = {k: float((p.grad.cpu() - grads_o[k]).norm() / grads_o[k].norm().clamp_min(1e-30)) for k, p in m.named_parameters()}
    print(f"path {path}: loss {float(loss):.9f} nv {int(nv)} rgb {rel_inf(rgb.detach().cpu(), rgb_o):.2e} depth {rel_inf(depth.cpu(), depth_o):.2e}")
    for k in errs:
        if errs[k] > 2e-5:
            print(f"    {k:32s} inf {errs[k]:.2e}  l2 {l2[k]:.2e}")
    st = m._last_call
    n = st.n_valid
    if path == 2 and m.shadingMode != "SH":
        import ctypes as C
        ldx = (st.field.in_dim + 1 + 7) // 8 * 8
        sec = lambda name, cols: m.workspace_section(st, name, torch.float32, n * cols).view(n, cols).double()
        X, h1, h2 = (m.saved_activation(st, nm, cc).double() for nm, cc in (("A_X", ldx), ("A_H1", 128), ("A_H2", 128)))
        # W1 in the kernels' internal column order is not needed: compare h2 from h1, and h1 sign statistics
        W2, b2 = m.renderModule.mlp[2].weight.double(), m.renderModule.mlp[2].bias.double()
        z2 = h1 @ W2.t() + b2
        flips2 = ((z2 > 0) != (h2 > 0))
        print(f"    layer-2 ReLU decisions that differ from fp64 on the GPU's own h1: {int(flips2.sum())} of {flips2.numel()};"
              f" |z| there max {float(z2[flips2].abs().max()) if flips2.any() else 0:.2e}; smallest |z2| overall {float(z2.abs().min()):.2e}")
        # oracle-side activations for the same shaded samples (same order: ray-major, sample index ascending)
        app = keep["app"]
        pts = O.march(f, rays[:, :3], rays[:, 3:6], jitter if train else None, S)[0]
        xn = O.normalize(f, pts)[app]
        view = rays[:, 3:6].view(-1, 1, 3).expand(pts.shape)[app]
        feat = O.app_feature(f, xn, None)
        gfeat = sec("A_FEAT", f.cfg.app_dim)
        print(f"    app features GPU vs oracle: {rel_inf(gfeat.cpu(), feat):.2e} (n {n} vs {feat.shape[0]})")
        cols = [feat, view]
        cfg = f.cfg
        if cfg.shading == "MLP_Fea":
            cols += [O.posenc(feat, cfg.fea_pe), O.posenc(view, cfg.view_pe)]
        elif cfg.shading == "MLP_PE":
            cols += [O.posenc(xn, cfg.pos_pe), O.posenc(view, cfg.view_pe)]
        else:
            cols += [O.posenc(xn, cfg.pos_pe), O.posenc(view, cfg.view_pe), O.posenc(feat, cfg.fea_pe)]
        x = torch.cat(cols, -1).double()
        p_ = {k: v.double() for k, v in f.params.items()}
        z1 = x @ p_["renderModule.mlp.0.weight"].t() + p_["renderModule.mlp.0.bias"]
        hh1 = torch.relu(z1)
        z2o = hh1 @ p_["renderModule.mlp.2.weight"].t() + p_["renderModule.mlp.2.bias"]
        if feat.shape[0] == n:
            d1 = (z1 > 0) != (h1.cpu() > 0)
            d2 = (z2o > 0) != (h2.cpu() > 0)
            print(f"    ReLU decisions GPU vs oracle(fp64): layer 1 {int(d1.sum())} (max |z| {float(z1[d1].abs().max()) if d1.any() else 0:.2e}),"
                  f" layer 2 {int(d2.sum())} (max |z| {float(z2o[d2].abs().max()) if d2.any() else 0:.2e})")
            print(f"    h1 err {rel_inf(h1.cpu(), hh1):.2e}  h2 err {rel_inf(h2.cpu(), torch.relu(z2o)):.2e}")
