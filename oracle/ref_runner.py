"""Runs the reference's own `models/` package (staged unmodified under oracle/_ref by oracle/make_ref.py) on the
benchmark workload: the CPU arm of bench.py (`--impl reference`, `cpu_baseline.kind = "reference"`) and a cross-check
of the restatement in field_oracle.py.  TEST / MEASUREMENT INFRASTRUCTURE ONLY — the product never imports this.

The reference's renderer.py cannot be imported (imageio, matplotlib, ...: SURVEY.md section 0.2), so its 14-line
chunk loop (renderer.py:13-26) is restated in `render`; everything it calls is the reference's code."""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(HERE, "_ref")


def available() -> bool:
    return os.path.exists(os.path.join(REF_ROOT, "models", "tensoRF.py"))


def _models():
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    from models import tensoRF, tensorBase          # the reference's modules
    return tensoRF, tensorBase


def build_model(args: dict, grid, kind="VM", aabb=((-1.5,) * 3, (1.5,) * 3), near_far=(2.0, 6.0), seed=20211202):
    """Reference TensorVMSplit / TensorCP on the CPU (train.py:550 seed; TensorCP via the workaround of SURVEY 0.5)."""
    tensoRF, tensorBase = _models()
    torch.manual_seed(seed)
    box = torch.tensor(aabb, dtype=torch.float32)
    if kind == "VM":
        return tensoRF.TensorVMSplit(dict(args), box, list(grid), list(near_far), "cpu")
    m = tensoRF.TensorCP.__new__(tensoRF.TensorCP)
    tensorBase.TensorBase.__init__(m, dict(args), box, list(grid), list(near_far), "cpu")
    return m


def load_params(model, params: dict, alpha_mask=None):
    """Give the reference model the very parameters (and alphaMask volume) of a product model / oracle field."""
    _, tensorBase = _models()
    model.load_state_dict({k: v.detach().cpu().contiguous() for k, v in params.items()}, strict=True)
    if alpha_mask is not None:
        vol, aabb = alpha_mask
        model.alphaMask = tensorBase.AlphaGridMask("cpu", aabb.cpu(), vol.cpu().float())
    return model


def plant_box(model, grid: int, kind="VM", half=0.6, amp=11.0):
    ind = (torch.linspace(-1.5, 1.5, grid).abs() < half).float()
    with torch.no_grad():
        if kind == "VM":
            model.density_plane[0][0, 0] = amp * ind[:, None] * ind[None, :]
            model.density_line[0][0, 0, :, 0] = ind
        else:
            model.density_line[0][0, 0, :, 0] = amp * ind
            model.density_line[1][0, 0, :, 0] = ind
            model.density_line[2][0, 0, :, 0] = ind


def render(rays, model, chunk=4096, N_samples=-1, white_bg=True, is_train=False):
    """renderer.py:13-26."""
    rgbs, depths, n = [], [], 0.0
    for ci in range(rays.shape[0] // chunk + int(rays.shape[0] % chunk > 0)):
        rgb, depth, nv = model(rays[ci * chunk:(ci + 1) * chunk], None, is_train=is_train, white_bg=white_bg, ndc_ray=False,
                               N_samples=N_samples)
        rgbs.append(rgb)
        depths.append(depth)
        n += float(nv)
    return torch.cat(rgbs), torch.cat(depths), n


def train_step(model, rays, target, N_samples):
    """train.py:323-338, :374-376 without the optimiser: render, MSE, backward.  Returns (loss, rgb, n_valid)."""
    for p in model.parameters():
        p.grad = None
    rgb, depth, nv = render(rays, model, chunk=rays.shape[0], N_samples=N_samples, white_bg=True, is_train=True)
    loss = torch.mean((rgb - target) ** 2)
    loss.backward()
    return loss.detach(), rgb.detach(), nv
