"""`-m gpu`: the CUDA path against the CPU oracle AT BASELINE.json's SHAPES (VERDICT r1 "What's weak" 1-2): the same
parameters, alphaMask, rays and jitter go through `oracle.loss_and_grads` (the restatement pinned bit-for-bit to the
reference's own outputs by tests/test_oracle_golden.py) and through the product's C ABI, 512 rays per case so the
oracle finishes in seconds:

  vm_300_mask256   BASELINE config 1 / the bench workload: TensorVMSplit 16/48, 300^3, S = 1039, alphaMask 256^3, train
  cp_96_288        BASELINE config 3: TensorCP 96/288 at 300^3, same mask recipe
  sh_view_tile     BASELINE config 2: SH shading, rays of an 800x800 test view, eval, N_samples = -1 (1036)
  vm_128_s443      the 128^3 stage of the schedule (S = 443, mask 128^3)
  vm_300_nomask    the first 2000 iterations (no alphaMask)
  fewshot_mlp153   BASELINE config 4's real shape (configs/free_tensorf_human.txt: app_dim 18, pos_pe 8, MLP => 153 inputs)

Bars: the alphaMask bits the product's updateAlphaMask builds equal the oracle's volume; the
packed (ray, i) sample set equals the oracle's ray_valid — bit-exact, including the 8^3 / 32^3 occupancy levels in
front of the exact test; sigma, weights, rgb, depth and EVERY parameter gradient within 1e-4 relative
(‖Δ‖∞ / ‖ref‖∞); shaded-sample count equal unless a weight straddles the threshold in fp32 noise."""
import pytest
import torch

from oracle import field_oracle as O
from tests.util import rel_inf

pytestmark = pytest.mark.gpu
TOL = 1e-4
N_RAYS = 512

CASES = {
    #                 model shading    grid mask  S     rays      train
    "vm_300_mask256": ("VM", "MLP_Fea", 300, 256, 1039, "random", True),
    "cp_96_288":      ("CP", "MLP_Fea", 300, 256, 1039, "random", True),
    "sh_view_tile":   ("VM", "SH",      300, 256, -1,   "camera", False),
    "vm_128_s443":    ("VM", "MLP_Fea", 128, 128, 443,  "random", True),
    "vm_300_nomask":  ("VM", "MLP_Fea", 300, 0,   1039, "random", True),
    "fewshot_mlp153": ("VM", "MLP",     300, 256, 519,  "random", True),
}


def _build(case):
    import bench
    import fewtensorf_b200 as ft
    kind, shading, grid, mask_grid, S, ray_kind, train = CASES[case]
    dev = torch.device("cuda", 0)
    args = dict(bench.ARGS, shadingMode=shading)
    cfg_kw = dict(model=kind, shading=shading, grid=(grid,) * 3)
    if case == "fewshot_mlp153":
        args.update(bench.FEWSHOT_ARGS)
        cfg_kw.update(density_n_comp=tuple(args["density_n_comp"]), app_n_comp=tuple(args["app_n_comp"]), app_dim=args["app_dim"],
                      pos_pe=args["pos_pe"], step_ratio=args["step_ratio"])
    torch.manual_seed(20211202)
    aabb = torch.tensor([[-1.5] * 3, [1.5] * 3], device=dev)
    ind = (torch.linspace(-1.5, 1.5, grid).abs() < 0.6).float().to(dev)
    if kind == "VM":
        m = ft.TensorVMSplit(args, aabb, [grid] * 3, [2.0, 6.0], dev)
        with torch.no_grad():
            m.density_plane[0][0, 0] = 11.0 * ind[:, None] * ind[None, :]
            m.density_line[0][0, 0, :, 0] = ind
    else:
        args.update(density_n_comp=[96] * 3, app_n_comp=[288] * 3)
        cfg_kw.update(density_n_comp=(96,) * 3, app_n_comp=(288,) * 3)
        m = ft.TensorCP(args, aabb, [grid] * 3, dev, near_far=[2.0, 6.0])
        with torch.no_grad():
            m.density_line[0][0, 0, :, 0] = 11.0 * ind
            m.density_line[1][0, 0, :, 0] = ind
            m.density_line[2][0, 0, :, 0] = ind
    f = O.Field(O.FieldConfig(**cfg_kw), {k: v.detach().cpu().contiguous() for k, v in m.state_dict().items()})
    if mask_grid:
        O.build_alpha_mask(f, (mask_grid,) * 3)
        m.updateAlphaMask((mask_grid,) * 3)
        got = m.alphaMask.alpha_volume[0, 0].cpu()
        assert got.shape == f.alpha_mask.volume.shape
        assert torch.equal(got > 0, f.alpha_mask.volume > 0), "updateAlphaMask bits differ from the oracle's volume"
        assert torch.equal(m.alphaMask.aabb.cpu(), f.alpha_mask.aabb)
    if ray_kind == "camera":
        cam = bench.camera_rays("cpu")
        rays = cam[torch.arange(0, cam.shape[0], cam.shape[0] // N_RAYS)[:N_RAYS] + 37 * 800 % 611].contiguous()
    else:
        rays, _ = bench.synthetic_rays(N_RAYS, seed=0)
    target = torch.rand(N_RAYS, 3, generator=torch.Generator().manual_seed(99))
    jitter = torch.rand(N_RAYS, 1, generator=torch.Generator().manual_seed(98)) if train else None
    return m, f, rays, target, jitter, (S if S > 0 else f.n_samples), train


@pytest.mark.parametrize("case", list(CASES))
def test_full_size_matches_oracle(case):
    from tests.gpu_util import dense_from_packed, flip_aware_reference, pin_rng, relu_decisions_in_oracle_order
    m, f, rays, target, jitter, S, train = _build(case)
    assert S == (m.nSamples if CASES[case][4] < 0 else CASES[case][4])
    keep = {}
    kw = dict(rays=rays, target=target, is_train=train, white_bg=True, n_samples=S, jitter=jitter)
    loss_o, rgb_o, depth_o, nv_o, grads_o = O.loss_and_grads(f, keep=keep, **kw)
    if train:
        pin_rng(m, jitter, False, True)
    rgb, depth, nv = m(rays.cuda(), None, white_bg=True, is_train=train, N_samples=CASES[case][4])
    loss = torch.mean((rgb - target.cuda()) ** 2)
    loss.backward()
    torch.cuda.synchronize()
    st = m._last_call
    d = dense_from_packed(m, st, N_RAYS, S)
    # stage 1: bit-exact sample set (aabb test + alphaMask decisions, with the coarse occupancy levels on)
    assert torch.equal(d["valid"], keep["valid"]), f"{int((d['valid'] != keep['valid']).sum())} sample decisions differ"
    assert int(keep["valid"].sum()) > 20 * N_RAYS
    assert rel_inf(d["sigma"], keep["sigma"]) < TOL
    assert rel_inf(d["weight"], keep["weight"]) < TOL
    # shaded set: equal except for weights within fp32 noise of the threshold
    diff = d["app"] != keep["app"]
    if diff.any():
        w = keep["weight"][diff]
        assert float(((w - f.cfg.weight_thres).abs() / f.cfg.weight_thres).max()) < 1e-3, "shaded sets differ away from the threshold"
    assert abs(int(nv) - int(nv_o)) <= max(1, int(1e-3 * int(nv_o)))
    assert int(nv_o) > 5 * N_RAYS
    assert rel_inf(rgb.detach().cpu(), rgb_o) < TOL
    assert rel_inf(depth.cpu(), depth_o) < TOL
    assert float(loss.detach()) == pytest.approx(float(loss_o), rel=5e-5)
    # ReLU kinks: a hidden pre-activation within ~1e-6 of zero may round to the other side on the tensor cores (3xTF32
    # keeps ~22 mantissa bits per operand); each such flip moves the gradients by O(1 / shaded samples), far above 1e-4
    # at 512 rays.  The decisions are read back, every differing pre-activation must be roundoff-sized, and the
    # gradients are then held to 1e-4 against the oracle evaluated with the kernels' decisions.
    if m.shadingMode != "SH" and not diff.any():
        n_flips, worst_z, res = flip_aware_reference(f, kw, keep, relu_decisions_in_oracle_order(m, st, keep["app"]))
        if n_flips:
            assert n_flips <= 64 and worst_z < 2e-5, (n_flips, worst_z)
            assert rel_inf(rgb.detach().cpu(), res[1]) < TOL
            grads_o = res[4]
            print(f"{case}: {n_flips} ReLU decisions differ from the oracle (largest |z| there {worst_z:.1e}); gradients "
                  "compared against the oracle evaluated with the kernels' decisions")
    worst = {}
    for k, p in m.named_parameters():
        assert p.grad is not None and p.grad.shape == grads_o[k].shape, k
        worst[k] = rel_inf(p.grad.cpu(), grads_o[k])
    bad = {k: v for k, v in worst.items() if not v < TOL}
    assert not bad, bad
