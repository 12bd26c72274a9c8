"""Profiling aid: one occupancy rebuild (256^3 lattice of the 300^3 benchmark field) and one filtering_rays pass over
an 800x800 frame, for `ncu -k regex:dense_alpha_bits|mask_dilate|filter_rays`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

dev = torch.device("cuda", 0)
model = bench.build_model(dev)                      # first rebuild (no mask yet)
model.updateAlphaMask((256, 256, 256))              # rebuild under the mask
rays = bench.camera_rays(dev)
kept, _ = model.filtering_rays(rays, torch.empty(rays.shape[0], 1, device=dev), N_samples=256)
torch.cuda.synchronize()
print("kept", kept.shape[0], "of", rays.shape[0])
