"""Debug aid: pairwise gradient differences between shading back ends / repeated runs at full size."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fewtensorf_b200._abi import load_test_lib  # noqa: E402
lib = load_test_lib()

dev = torch.device("cuda", 0)
m = bench.build_model(dev)
rays, target = bench.synthetic_rays(4096, seed=0)
rays, target = rays.to(dev), target.to(dev)


def step(path):
    lib.ftrf_set_shade_path(path)
    torch.manual_seed(7)
    loss, nv = m.train_step(rays, target, None, white_bg=True, N_samples=1039)
    torch.cuda.synchronize()
    return float(loss), nv, {k: p.grad.detach().clone() for k, p in m.named_parameters()}


runs = {"2a": step(2), "2b": step(2), "1": step(1), "0a": step(0), "0b": step(0)}
rel = lambda a, b: float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
for x, y in (("2a", "2b"), ("0a", "0b"), ("2a", "1"), ("2a", "0a"), ("1", "0a")):
    print(x, y, "loss", runs[x][0], runs[y][0], "nv", runs[x][1], runs[y][1])
    for k in runs[x][2]:
        r = rel(runs[x][2][k], runs[y][2][k])
        if r > 1e-5:
            print(f"    {k:32s} {r:.3e}")
