"""Rays made and sampled on the device.

`generate_rays` replaces the data loader's per-image `get_ray_directions` + `get_rays`
(dataLoader/ray_utils.py:24-85, dataLoader/blender.py:57-59,:106) with one kernel launch from a 3x4 pose;
`render_view` feeds it to the render boundary, so an 800x800 frame needs 48 bytes of pose instead of the
15 MB host->device ray copy renderer.py:18 makes; `DeviceRaySampler` is train.py:44-56's `SimpleSampler`
with the ray / colour tables and the permutation resident on the GPU (train.py:290-298 gathers on the host
and copies the batch every iteration).
"""
import ctypes as C
from typing import Optional

import torch

from ._abi import FtrfCamera, check, lib


from ._abi import current_stream_handle as _stream  # noqa: E402


def generate_rays(H: int, W: int, focal, c2w, center=None, normalize: bool = True, blender: bool = False,
                  row0: int = 0, row_step: int = 1, n_rows: Optional[int] = None, device='cuda') -> torch.Tensor:
    """[n_rows * W, 6] fp32 (origin, direction) for image rows row0, row0 + row_step, ...

    focal: scalar or (fx, fy) as get_ray_directions takes it; center defaults to (W/2, H/2)
    (ray_utils.py:38); c2w: 3x4 (or 4x4) camera-to-world, any device / dtype; normalize: divide the
    camera-space direction by its norm first (blender.py:59)."""
    fx, fy = (focal if isinstance(focal, (tuple, list)) else (focal, focal))
    cx, cy = center if center is not None else (W / 2, H / 2)
    m = torch.as_tensor(c2w, dtype=torch.float32).detach().cpu().reshape(-1, 4)[:3]
    cam = FtrfCamera()
    cam.H, cam.W = int(H), int(W)
    cam.fx, cam.fy, cam.cx, cam.cy = float(fx), float(fy), float(cx), float(cy)
    for i, v in enumerate(m.reshape(-1).tolist()):
        cam.c2w[i] = v
    cam.normalize, cam.blender = int(bool(normalize)), int(bool(blender))
    if n_rows is None:
        n_rows = max(0, (H - row0 + row_step - 1) // row_step)
    out = torch.empty(n_rows * W, 6, dtype=torch.float32, device=device)
    with torch.cuda.device(out.device):
        check(lib.ftrf_generate_rays(C.byref(cam), int(row0), int(row_step), int(n_rows), out.data_ptr(), _stream()),
              "generate_rays")
    return out


def render_view(tensorf, c2w, H: int, W: int, focal, center=None, mask=None, N_samples=-1, ndc_ray=False,
                white_bg=True, normalize=True, blender=False, row0=0, row_step=1, n_rows=None):
    """One test view straight from its pose: (rgb [rows, W, 3], depth [rows, W], n_valid_samples).
    What renderer.py:67-75 does per image (rays from the dataset -> OctreeRender_trilinear_fast ->
    reshape to H x W), without host rays."""
    from .renderer import OctreeRender_trilinear_fast
    rays = generate_rays(H, W, focal, c2w, center, normalize, blender, row0, row_step, n_rows, device=tensorf.device)
    rgb, _, depth, _, _, nv = OctreeRender_trilinear_fast(rays, tensorf, mask, chunk=rays.shape[0], N_samples=N_samples,
                                                         ndc_ray=ndc_ray, white_bg=white_bg, is_train=False,
                                                         device=tensorf.device)
    rows = rays.shape[0] // W
    return rgb.reshape(rows, W, 3), depth.reshape(rows, W), nv


class DeviceRaySampler:
    """train.py:44-56 SimpleSampler over GPU-resident tables: `next()` returns (rays [B,6], rgbs [B,3]) device
    tensors, reshuffling with `torch.randperm` on the device when fewer than one batch is left — the same
    epoch logic (`curr += batch; if curr + batch > total: reshuffle, curr = 0`).  The reference shuffles
    with numpy's host RNG, so the order differs by construction; `ids` can be injected for tests."""

    def __init__(self, all_rays: torch.Tensor, all_rgbs: torch.Tensor, batch: int, device='cuda', generator=None):
        if all_rays.shape[0] != all_rgbs.shape[0]:
            raise ValueError("DeviceRaySampler: rays and colours differ in length")
        self.rays = all_rays.to(device=device, dtype=torch.float32).contiguous()
        self.rgbs = all_rgbs.to(device=device, dtype=torch.float32).contiguous()
        self.total, self.batch = int(self.rays.shape[0]), int(batch)
        self.curr = self.total
        self.ids = None
        self.generator = generator

    def nextids(self) -> torch.Tensor:
        self.curr += self.batch
        if self.ids is None or self.curr + self.batch > self.total:
            self.ids = torch.randperm(self.total, device=self.rays.device, generator=self.generator)
            self.curr = 0
        return self.ids[self.curr:self.curr + self.batch]

    def next(self):
        idx = self.nextids()
        return self.rays.index_select(0, idx), self.rgbs.index_select(0, idx)
