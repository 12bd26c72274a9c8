"""Debug aid: host-side timeline of the end-to-end autograd step of bench.py (pinned host rays -> renderer -> torch MSE ->
autograd backward -> loss.item()), averaged over some steps: where the host is when.  python scripts/e2e_timeline.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import fewtensorf_b200 as ft  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
model = bench.build_model(dev)
rays_h, target_h = bench.synthetic_rays(bench.RAYS_PER_GPU, seed=0)
rays_p, target_p = rays_h.pin_memory(), target_h.pin_memory()
params = list(model.parameters())
names = ["h2d enqueued", "grads cleared", "forward returned", "loss enqueued", "backward returned", "loss.item()"]
acc = [0.0] * len(names)
K = 40
for it in range(K + 5):
    t = [time.perf_counter()]
    r = rays_p.to(dev, non_blocking=True)
    tg = target_p.to(dev, non_blocking=True)
    t.append(time.perf_counter())
    for p in params:
        p.grad = None
    t.append(time.perf_counter())
    rgb, _, depth, _, _, nvs = ft.OctreeRender_trilinear_fast(r, model, None, chunk=bench.RAYS_PER_GPU, N_samples=bench.N_SAMPLES,
                                                              white_bg=True, is_train=True, device=dev)
    t.append(time.perf_counter())
    loss = torch.mean((rgb - tg) ** 2)
    t.append(time.perf_counter())
    loss.backward()
    t.append(time.perf_counter())
    loss.item()
    t.append(time.perf_counter())
    if it >= 5:
        for i in range(len(names)):
            acc[i] += (t[i + 1] - t[0]) * 1e6 / K
print("host clock, us from the start of the step:", {n: round(a, 1) for n, a in zip(names, acc)})

if len(sys.argv) > 1 and sys.argv[1] == "profile":            # where the host's time goes (cProfile, 200 steps)
    import cProfile
    import pstats

    def one():
        r = rays_p.to(dev, non_blocking=True)
        tg = target_p.to(dev, non_blocking=True)
        for p in params:
            p.grad = None
        rgb, _, depth, _, _, nvs = ft.OctreeRender_trilinear_fast(r, model, None, chunk=bench.RAYS_PER_GPU, N_samples=bench.N_SAMPLES,
                                                                  white_bg=True, is_train=True, device=dev)
        loss = torch.mean((rgb - tg) ** 2)
        loss.backward()
        return loss.item()

    pr = cProfile.Profile()
    pr.enable()
    for _ in range(200):
        one()
    pr.disable()
    pstats.Stats(pr).sort_stats("tottime").print_stats(40)
