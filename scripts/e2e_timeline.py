"""Debug aid: host-side timeline of the reference-facing training step (mean over 200 steps, microseconds from the
step's start): where the host is when the GPU could start each kernel sequence.  `python scripts/e2e_timeline.py [ref]`
runs the step in the reference's order (zero_grad after the loss, train.py:374)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import fewtensorf_b200 as ft  # noqa: E402
import fewtensorf_b200.module as M  # noqa: E402

ref_order = len(sys.argv) > 1 and sys.argv[1] == "ref"
if len(sys.argv) > 1 and sys.argv[1] == "st":
    torch.autograd.set_multithreading_enabled(False)        # engine runs in the calling thread (no device-thread hop)
dev = torch.device("cuda", 0)
model = bench.build_model(dev)
rays_h, target_h = bench.synthetic_rays(4096, seed=0)
rays_p, target_p = rays_h.pin_memory(), target_h.pin_memory()
params = list(model.parameters())
marks, t0 = {}, [0.0]


def mark(name):
    marks.setdefault(name, []).append((time.perf_counter() - t0[0]) * 1e6)


def wrap(obj, name, key):
    fn = getattr(obj, name)

    def w(*a, **k):
        mark(key + ">")
        try:
            return fn(*a, **k)
        finally:
            mark(key + "<")
    setattr(obj, name, w)


wrap(M.TensorBase, "forward", "module.forward")
wrap(M.TensorBase, "_draw_rng", "draw_rng")
wrap(M.TensorBase, "_launch_forward", "launch_forward")
wrap(M.TensorBase, "_finish_forward", "finish_forward")
wrap(M.TensorBase, "_run_backward", "run_backward")
orig_bwd = M._RenderFn.backward


def bwd(ctx, *g):
    mark("RenderFn.backward>")
    try:
        return orig_bwd(ctx, *g)
    finally:
        mark("RenderFn.backward<")


M._RenderFn.backward = staticmethod(bwd)


def step():
    torch.cuda.synchronize()
    t0[0] = time.perf_counter()
    r = rays_p.to(dev, non_blocking=True); tg = target_p.to(dev, non_blocking=True)
    mark("h2d<")
    if not ref_order:
        for p in params:
            p.grad = None
        mark("zero_grad<")
    rgb, _, depth, _, _, nvs = ft.OctreeRender_trilinear_fast(r, model, None, chunk=4096, N_samples=1039, white_bg=True,
                                                              is_train=True, device=dev)
    mark("renderer<")
    loss = torch.mean((rgb - tg) ** 2)
    mark("loss<")
    if ref_order:
        for p in params:
            p.grad = None
        mark("zero_grad<")
    loss.backward()
    mark("backward()<")
    v = loss.item()
    mark("item<")
    return v


for _ in range(30):
    step()
marks.clear()
for _ in range(200):
    step()
print("order:", "reference (zero_grad after loss)" if ref_order else "zero_grad first")
for k, v in sorted(marks.items(), key=lambda kv: sum(kv[1]) / len(kv[1])):
    print(f"{k:24s} {sum(v) / len(v):8.1f} us")
