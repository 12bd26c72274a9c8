"""Golden vectors for the occupancy rebuild and the ray pre-filter, from the UNMODIFIED
reference: `TensorBase.updateAlphaMask` (models/tensorBase.py:233-256; first build without a mask, then a
rebuild at another resolution under the first mask, which exercises compute_alpha's alphaMask branch
:300-304) and `TensorBase.filtering_rays` (:259-288) in both modes.

Run in the build container only:   python tests/golden/make_golden_occupancy.py
Writes tests/golden/occ_<kind>.npz: parameters, the two mask volumes with the boxes updateAlphaMask
returned, the rays and the keep masks filtering_rays selected.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import build, make_args, plant, quiet, rays_for            # noqa: E402  (imports the reference models)


def filter_mask(m, rays, **kw):
    """filtering_rays returns the selected rays, not the mask: recover it by tagging each ray with its index."""
    tag = torch.arange(rays.shape[0], dtype=torch.float32)[:, None]
    _, kept = quiet(m.filtering_rays, rays, tag, **kw)
    keep = torch.zeros(rays.shape[0], dtype=torch.bool)
    keep[kept[:, 0].long()] = True
    return keep


def one(kind, grid, dn, an, seed):
    args = make_args("MLP_Fea", dn, an, 27)
    m = build(kind, args, grid, seed)
    plant(m, kind, grid, half=0.45)
    out = {"p:" + k: v.detach().clone() for k, v in m.state_dict().items()}
    box1 = quiet(m.updateAlphaMask, (24, 20, 22))
    out["am1_volume"] = m.alphaMask.alpha_volume[0, 0].bool().clone()
    out["am1_box"] = box1
    rays, _ = rays_for(192, seed + 1)
    g = torch.Generator().manual_seed(seed + 5)
    # a few rays that graze the occupied region / the box faces, and some that start inside the box
    rays[2:34, :3] = torch.tensor([0.0, 0.0, 4.0]) + 0.05 * torch.randn(32, 3, generator=g)
    tgt = torch.stack([torch.linspace(-1.6, 1.6, 32), torch.linspace(1.6, -1.6, 32), torch.zeros(32)], -1)
    d = tgt - rays[2:34, :3]
    rays[2:34, 3:] = d / d.norm(dim=-1, keepdim=True)
    rays[34:42, :3] = 0.3 * torch.randn(8, 3, generator=g)
    out["rays"] = rays
    out["keep1_s48"] = filter_mask(m, rays, N_samples=48, chunk=50)
    out["keep1_s24"] = filter_mask(m, rays, N_samples=24)
    out["keep_bbox"] = filter_mask(m, rays, bbox_only=True, chunk=64)
    box2 = quiet(m.updateAlphaMask, (30, 26, 21))
    out["am2_volume"] = m.alphaMask.alpha_volume[0, 0].bool().clone()
    out["am2_box"] = box2
    out["keep2_s48"] = filter_mask(m, rays, N_samples=48)
    meta = dict(kind=kind, shading="MLP_Fea", grid=list(grid), S=-1, is_train=False, white_bg=True, ndc=False,
                near_far=[2.0, 6.0], args=args, step_size=float(m.stepSize), nSamples=int(m.nSamples))
    np.savez_compressed(os.path.join(HERE, f"occ_{kind.lower()}.npz"), meta=np.array(repr(meta)),
                        **{k: v.numpy() for k, v in out.items()})
    print(kind, "mask1", int(out["am1_volume"].sum()), "mask2", int(out["am2_volume"].sum()), "box", box2.tolist(),
          "keep", int(out["keep1_s48"].sum()), int(out["keep1_s24"].sum()), int(out["keep_bbox"].sum()),
          int(out["keep2_s48"].sum()), "of", rays.shape[0])


if __name__ == "__main__":
    one("VM", (20, 24, 28), [8, 8, 8], [12, 12, 12], 21)
    one("CP", (22, 18, 20), [16, 16, 16], [24, 24, 24], 22)
