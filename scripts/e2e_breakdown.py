"""Debug aid: host-side breakdown of the reference-facing training step (wall clock, sync after each piece)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import fewtensorf_b200 as ft  # noqa: E402

dev = torch.device("cuda", 0)
model = bench.build_model(dev)
rays_h, target_h = bench.synthetic_rays(4096, seed=0)
rays_p, target_p = rays_h.pin_memory(), target_h.pin_memory()
acc = {}


def tick(name, t0):
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    acc[name] = acc.get(name, 0.0) + (t1 - t0)
    return t1


for it in range(60):
    if it == 10:
        acc.clear()
    torch.cuda.synchronize()
    t = time.perf_counter()
    r = rays_p.to(dev, non_blocking=True); tg = target_p.to(dev, non_blocking=True)
    t = tick("h2d", t)
    for p in model.parameters():
        p.grad = None
    t = tick("grad_none", t)
    rgb, _, depth, _, _, nvs = ft.OctreeRender_trilinear_fast(r, model, None, chunk=4096, N_samples=1039, white_bg=True, is_train=True, device=dev)
    t = tick("forward", t)
    loss = torch.mean((rgb - tg) ** 2)
    t = tick("loss", t)
    loss.backward()
    t = tick("backward", t)
    v = loss.item()
    t = tick("item", t)
print({k: round(v / 50 * 1e3, 4) for k, v in acc.items()}, "ms; sum", round(sum(acc.values()) / 50 * 1e3, 4))
# unsynchronised host time of each piece (how long the CPU is busy)
import cProfile, pstats
def step():
    r = rays_p.to(dev, non_blocking=True); tg = target_p.to(dev, non_blocking=True)
    for p in model.parameters():
        p.grad = None
    rgb, _, depth, _, _, nvs = ft.OctreeRender_trilinear_fast(r, model, None, chunk=4096, N_samples=1039, white_bg=True, is_train=True, device=dev)
    loss = torch.mean((rgb - tg) ** 2)
    loss.backward()
    return loss.item()
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    step()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
