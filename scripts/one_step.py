"""Profiling target: three training steps of the benchmark field (python scripts/one_step.py [rays])."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

dev = torch.device("cuda", 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
m = bench.build_model(dev)
rays, target = bench.synthetic_rays(n, seed=0)
rays, target = rays.to(dev), target.to(dev)
for _ in range(3):
    loss, nv = m.train_step(rays, target, None, white_bg=True, N_samples=bench.N_SAMPLES)
    torch.cuda.synchronize()
print(float(loss), nv)
