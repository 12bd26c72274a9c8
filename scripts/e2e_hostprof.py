"""Debug aid: cProfile (tottime) of the reference-facing training step's host side."""
import cProfile
import os
import pstats
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import fewtensorf_b200 as ft  # noqa: E402

dev = torch.device("cuda", 0)
model = bench.build_model(dev)
rays_h, target_h = bench.synthetic_rays(4096, seed=0)
rays_p, target_p = rays_h.pin_memory(), target_h.pin_memory()
params = list(model.parameters())


def step():
    r = rays_p.to(dev, non_blocking=True); tg = target_p.to(dev, non_blocking=True)
    for p in params:
        p.grad = None
    rgb, _, depth, _, _, nvs = ft.OctreeRender_trilinear_fast(r, model, None, chunk=4096, N_samples=1039, white_bg=True,
                                                              is_train=True, device=dev)
    loss = torch.mean((rgb - tg) ** 2)
    loss.backward()
    return loss.item()


for _ in range(20):
    step()
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(45)
