"""Generate golden vectors from the UNMODIFIED reference (`/root/reference/models`).

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_golden.py
Writes tests/golden/<case>.npz with: the reference model's parameters, the rays, the RNG
draws the reference consumed, the stage intermediates obtained by calling the reference's own
methods in `TensorBase.forward` order, the outputs, and all parameter gradients of
`mean((rgb_map-target)**2)` (train.py:338).  Nothing from this script ships in the product.
"""
import contextlib
import io
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
with contextlib.redirect_stdout(io.StringIO()):
    from models.tensoRF import TensorVMSplit, TensorCP          # noqa: E402
    from models.tensorBase import TensorBase, raw2alpha           # noqa: E402
    from models.mlp import SHRender                               # noqa: E402


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def make_args(shading, dn, an, app_dim, pos_pe=2, view_pe=2, fea_pe=2, act="softplus", step_ratio=0.5):
    return {'step_ratio': step_ratio, 'fea2denseAct': act, 'density_n_comp': dn, 'app_n_comp': an,
            'app_dim': app_dim, 'density_shift': -10.0, 'distance_scale': 25.0, 'alphaMask_thres': 0.001,
            'shadingMode': shading, 'pos_pe': pos_pe, 'view_pe': view_pe, 'fea_pe': fea_pe, 'featureC': 128}


def build(kind, args, grid, seed, near_far=(2.0, 6.0)):
    torch.manual_seed(seed)
    aabb = torch.tensor([[-1.5, -1.5, -1.5], [1.5, 1.5, 1.5]])
    if kind == "VM":
        m = quiet(TensorVMSplit, args, aabb, list(grid), list(near_far), 'cpu')
    else:   # TensorCP's own ctor mis-forwards `device`
        m = TensorCP.__new__(TensorCP)
        quiet(TensorBase.__init__, m, args, aabb, list(grid), list(near_far), 'cpu')
    return m


def plant(m, kind, grid, half=0.6, amp=11.0):
    def ind(axis):
        return (torch.linspace(-1.5, 1.5, grid[axis]).abs() < half).float()
    with torch.no_grad():
        if kind == "VM":
            m.density_plane[0][0, 0] = amp * ind(1)[:, None] * ind(0)[None, :]
            m.density_line[0][0, 0, :, 0] = ind(2)
        else:
            m.density_line[0][0, 0, :, 0] = amp * ind(2)
            m.density_line[1][0, 0, :, 0] = ind(1)
            m.density_line[2][0, 0, :, 0] = ind(0)


def rays_for(n, seed):
    g = torch.Generator().manual_seed(seed)
    o = torch.randn(n, 3, generator=g)
    o = 4.0 * o / o.norm(dim=-1, keepdim=True)
    tgt = (torch.rand(n, 3, generator=g) - 0.5) * 2
    d = tgt - o
    d = d / d.norm(dim=-1, keepdim=True)
    rays = torch.cat([o, d], -1)
    rays[0, 3:] = torch.tensor([0.0, 0.0, -1.0]); rays[0, :3] = torch.tensor([0.1, -0.2, 4.0])   # exact-zero dir comps
    rays[1, :3] = torch.tensor([5.0, 5.0, 5.0]); rays[1, 3:] = torch.tensor([1.0, 0.0, 0.0])     # misses the box
    target = torch.rand(n, 3, generator=g)
    return rays, target


def stage_capture(m, rays, mask, is_train, ndc, S, seed):
    """Replay forward's stages with the reference's own methods (tensorBase.py:337-370)."""
    torch.manual_seed(seed)
    o, d = rays[:, :3], rays[:, 3:6]
    with torch.no_grad():
        if ndc:
            pts, z, valid = m.sample_ray_ndc(o, d, is_train=is_train, N_samples=S)
            dists = torch.cat((z[:, 1:] - z[:, :-1], torch.zeros_like(z[:, :1])), dim=-1)
            dists = dists * torch.norm(d, dim=-1, keepdim=True)
            z = z.expand(rays.shape[0], -1)
        else:
            pts, z, valid = m.sample_ray(o, d, is_train=is_train, N_samples=S)
            dists = torch.cat((z[:, 1:] - z[:, :-1], torch.zeros_like(z[:, :1])), dim=-1)
        if m.alphaMask is not None:
            keep = m.alphaMask.sample_alpha(pts[valid]) > 0
            inv = ~valid
            inv[valid] |= ~keep
            valid = ~inv
        sigma = torch.zeros(pts.shape[:-1])
        xn = m.normalize_coord(pts)
        if valid.any():
            dm = None if mask is None else mask['decomp']['den']
            sigma[valid] = m.feature2density(m.compute_densityfeature(xn[valid], dm))
        _, w, _ = raw2alpha(sigma, dists * m.distance_scale)
        app = w > m.rayMarch_weight_thres
        feat = torch.zeros(0, m.app_dim)
        if app.any():
            am = None if mask is None else mask['decomp']['app']
            feat = m.compute_appfeature(xn[app], am)
    return dict(z=z.contiguous(), valid=valid, sigma=sigma, weight=w, app=app, feat=feat)


def run_case(name, kind, args, grid, n_rays, S, *, seed=7, **kw):
    """Fixtures must not sit on the shading threshold: a sample whose weight is within fp32 noise of
    rayMarch_weight_thres may be shaded by one implementation and not by another, which
    moves rgb by ~1e-4 for reasons unrelated to kernel correctness.  Re-seed until the nearest weight
    is at least 0.2 % away from the threshold."""
    if len(sys.argv) > 1 and name not in sys.argv[1:]:        # python make_golden.py [case ...]: only those
        return
    for attempt in range(50):
        gap = _run_case(name, kind, args, grid, n_rays, S, seed=seed + 100 * attempt, **kw)
        if gap > 2e-3:
            return
        print(f"  {name}: weight within {gap:.1e} of the threshold, re-seeding")
    raise SystemExit(f"{name}: no seed without threshold straddlers")


def _run_case(name, kind, args, grid, n_rays, S, *, is_train, white_bg, ndc=False, mask=None,
              mask_grid=None, sh=False, seed=7, near_far=(2.0, 6.0)):
    m = build(kind, args, grid, seed, near_far)
    plant(m, kind, grid)
    if sh:   # SH is wired by hand (mlp.py:15-19)
        del m._modules['renderModule']
        m.renderModule = lambda pts, v, f, mask: SHRender(pts, v, f)
    if mask_grid is not None:
        quiet(m.updateAlphaMask, tuple(mask_grid))
    rays, target = rays_for(n_rays, seed + 1)
    if ndc:   # put the rays inside a unit-ish NDC style box so that samples land in the aabb
        rays[:, :3] = (torch.rand(n_rays, 3) - 0.5) * 2.0
        rays[:, 2] = -1.4
        rays[:, 3:] = torch.tensor([0.05, -0.03, 1.0]) * (1 + 0.2 * torch.rand(n_rays, 1))

    call_seed = seed + 2
    # RNG draws exactly as the reference will consume them (tensorBase.py:183/:201 then :380)
    torch.manual_seed(call_seed)
    jitter = None
    if is_train:
        jitter = torch.rand(1, S) if ndc else torch.rand(n_rays, 1)
    coin = bool(torch.rand((1,)) < 0.5) if (is_train and not white_bg) else False

    st = stage_capture(m, rays, mask, is_train, ndc, S, call_seed)
    thres = m.rayMarch_weight_thres
    gap = float(((st["weight"][st["valid"]] - thres).abs() / thres).min())
    if gap <= 2e-3:
        return gap
    torch.manual_seed(call_seed)
    for p in m.parameters():
        p.grad = None
    rgb, depth, nv = m(rays, mask, white_bg=white_bg, is_train=is_train, ndc_ray=ndc, N_samples=S)
    loss = torch.mean((rgb - target) ** 2)
    loss.backward()

    out = {"rays": rays, "target": target, "rgb_map": rgb.detach(), "depth_map": depth, "n_valid": nv,
           "loss": loss.detach(), "coin": torch.tensor(coin)}
    if jitter is not None:
        out["jitter"] = jitter
    for k, v in st.items():
        out["st_" + k] = v
    sd = m.state_dict()
    for k, v in sd.items():
        if sh and k.startswith("renderModule"):
            continue
        out["p:" + k] = v
    for k, p in m.named_parameters():
        if sh and k.startswith("renderModule"):
            continue
        out["g:" + k] = p.grad if p.grad is not None else torch.zeros_like(p)
    if m.alphaMask is not None:
        out["am_volume"] = m.alphaMask.alpha_volume[0, 0].bool()
        out["am_aabb"] = m.alphaMask.aabb
    if mask is not None:
        for grp in ("encoding", "decomp"):
            for k, v in mask[grp].items():
                if v is None:
                    continue
                if isinstance(v, (list, tuple)):
                    for i, t in enumerate(v):
                        out[f"m:{grp}:{k}:{i}"] = t
                else:
                    out[f"m:{grp}:{k}"] = v
    meta = dict(kind=kind, shading="SH" if sh else args['shadingMode'], grid=list(grid), S=S, is_train=is_train,
                white_bg=white_bg, ndc=ndc, near_far=list(near_far), args=args, step_size=float(m.stepSize),
                nSamples=int(m.nSamples))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), meta=np.array(repr(meta)),
                        **{k: (v.numpy() if torch.is_tensor(v) else v) for k, v in out.items()})
    print(f"{name}: seed {seed} rgb mean {rgb.mean():.4f} n_valid {int(nv)} valid {int(st['valid'].sum())} "
          f"loss {loss:.5f} coin {coin} threshold gap {gap:.1e}")
    return gap


def main():
    A = make_args
    run_case("vm_fea_train_mask", "VM", A("MLP_Fea", [16, 16, 16], [48, 48, 48], 27), (20, 24, 28), 96, 70,
             is_train=True, white_bg=True, mask_grid=(24, 20, 22))
    run_case("vm_fea_eval_nomask", "VM", A("MLP_Fea", [8, 8, 8], [12, 12, 12], 27), (16, 16, 16), 64, -1,
             is_train=False, white_bg=True)
    g = torch.Generator().manual_seed(5)

    def fm(n):   # FreeNeRF-like: leading entries 1, the rest partially visible
        m = 0.3 + 0.7 * torch.rand(n, generator=g)
        m[:2] = 1.0
        return m
    vec_mask = {"encoding": {"pos": torch.rand(12, generator=g), "view": torch.rand(12, generator=g),
                             "fea": torch.rand(72, generator=g)},
                "decomp": {"den": [fm(8) for _ in range(3)],
                           "app": [fm(12) for _ in range(3)]}}
    run_case("vm_mlp_vecmask_blackbg", "VM", A("MLP", [8, 8, 8], [12, 12, 12], 18), (18, 16, 14), 64, 50,
             is_train=True, white_bg=False, mask=vec_mask, mask_grid=(16, 16, 16), seed=14)
    sc_mask = {"encoding": {"pos": torch.tensor(0.7), "view": torch.tensor(0.4), "fea": None},
               "decomp": {"den": torch.tensor([0.9, 0.8, 0.6, 0.3, 1, 1, 1, 1.]),
                          "app": torch.tensor([0.5, 0.7, 0.95] + [1.0] * 9)}}
    run_case("vm_pe_scalarmask", "VM", A("MLP_PE", [8, 8, 8], [12, 12, 12], 27), (16, 16, 16), 64, 48,
             is_train=True, white_bg=True, mask=sc_mask, seed=13)
    run_case("vm_sh_eval_mask", "VM", A("MLP_Fea", [8, 8, 8], [12, 12, 12], 27), (16, 16, 16), 64, -1,
             is_train=False, white_bg=True, sh=True, mask_grid=(16, 16, 16), seed=17)
    run_case("cp_fea_train", "CP", A("MLP_Fea", [24], [48], 27), (34, 36, 30), 64, 100,
             is_train=True, white_bg=True, mask_grid=(16, 16, 16), seed=19)
    cp_mask = {"encoding": {"pos": None, "view": None, "fea": torch.tensor(0.6)},
               "decomp": {"den": [fm(24)], "app": [fm(48)]}}
    run_case("cp_fea_vecmask", "CP", A("MLP_Fea", [24], [48], 27), (32, 32, 32), 48, 90,
             is_train=True, white_bg=True, mask=cp_mask, seed=23)
    run_case("vm_fea_ndc_train", "VM", A("MLP_Fea", [8, 8, 8], [12, 12, 12], 27), (16, 16, 16), 48, 40,
             is_train=True, white_bg=False, ndc=True, seed=29, near_far=(0.0, 3.0))
    # unequal, multiple-of-8 appearance components with per-channel decomposition masks on a non-cubic grid:
    # the tensor-core appearance forward, the run-merging scatters and the fused MLP kernels under masks
    tc_mask = {"encoding": {"pos": None, "view": torch.rand(12, generator=g), "fea": torch.rand(108, generator=g)},
               "decomp": {"den": [fm(8) for _ in range(3)], "app": [fm(16), fm(8), fm(24)]}}
    run_case("vm_fea_vecmask_tc", "VM", A("MLP_Fea", [8, 8, 8], [16, 8, 24], 27), (20, 18, 22), 96, 60,
             is_train=True, white_bg=True, mask=tc_mask, mask_grid=(18, 20, 16), seed=37)
    run_case("vm_relu_eval", "VM", A("MLP_Fea", [8, 8, 8], [12, 12, 12], 27, act="relu"), (16, 16, 16), 48, 40,
             is_train=False, white_bg=False, seed=31)


if __name__ == "__main__":
    main()
