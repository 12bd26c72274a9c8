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

# un-synchronised wall-clock split of the same step (what the e2e number is made of)
acc2 = {}
def lap(name, t0):
    t1 = time.perf_counter()
    acc2[name] = acc2.get(name, 0.0) + (t1 - t0)
    return t1
for it in range(110):
    if it == 10:
        acc2.clear()
    t = time.perf_counter()
    r = rays_p.to(dev, non_blocking=True); tg = target_p.to(dev, non_blocking=True)
    t = lap("h2d", t)
    for p in model.parameters():
        p.grad = None
    t = lap("grad_none", t)
    rgb, _, depth, _, _, nvs = ft.OctreeRender_trilinear_fast(r, model, None, chunk=4096, N_samples=1039, white_bg=True, is_train=True, device=dev)
    t = lap("forward(call+wait)", t)
    loss = torch.mean((rgb - tg) ** 2)
    t = lap("loss", t)
    loss.backward()
    t = lap("backward(call)", t)
    v = loss.item()
    t = lap("item(wait)", t)
print("unsynced:", {k: round(v / 100 * 1e3, 4) for k, v in acc2.items()}, "ms; sum", round(sum(acc2.values()) / 100 * 1e3, 4))
# host time inside the forward call before the first kernel can start / until everything is enqueued
import fewtensorf_b200.module as M
orig_launch, orig_read = M.TensorBase._launch_forward, M.TensorBase._read_counters
marks = {}
def timed_launch(self, *a, **k):
    t0 = time.perf_counter(); out = orig_launch(self, *a, **k); marks["launch"] = marks.get("launch", 0.0) + time.perf_counter() - t0
    return out
def timed_read(self, *a, **k):
    t0 = time.perf_counter(); out = orig_read(self, *a, **k); marks["wait"] = marks.get("wait", 0.0) + time.perf_counter() - t0
    return out
M.TensorBase._launch_forward, M.TensorBase._read_counters = timed_launch, timed_read
t0 = time.perf_counter()
for it in range(100):
    step()
tot = time.perf_counter() - t0
print("inside forward:", {k: round(v / 100 * 1e3, 4) for k, v in marks.items()}, "ms; step", round(tot / 100 * 1e3, 4))

# finer: wall-clock of the pieces of the forward call (monkeypatched timers, un-synchronised)
M.TensorBase._launch_forward, M.TensorBase._read_counters = orig_launch, orig_read
marks.clear()
def wrap(obj, name, key):
    fn = getattr(obj, name)
    def w(*a, **k):
        t0 = time.perf_counter()
        try:
            return fn(*a, **k)
        finally:
            marks[key] = marks.get(key, 0.0) + time.perf_counter() - t0
    setattr(obj, name, w)
wrap(M.TensorBase, "_draw_rng", "draw_rng")
wrap(M.TensorBase, "_run_forward", "run_forward(launch+prepare+wait)")
wrap(M.TensorBase, "_launch_forward", "launch_forward")
wrap(M.TensorBase, "_finish_forward", "finish_forward(wait)")
wrap(M.TensorBase, "_fill_field", "fill_field")
wrap(M.TensorBase, "grad_buffer", "grad_buffer")
wrap(M.TensorBase, "_run_backward", "run_backward")
wrap(M.TensorBase, "forward", "module.forward")
wrap(M.TensorBase, "_all_params", "all_params")
for it in range(100):
    step()
print("pieces:", {k: round(v / 100 * 1e3, 4) for k, v in marks.items()})
