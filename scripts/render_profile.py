"""Debug aid: stage times (CUDA events between launches) of one 800x800 render.  python scripts/render_profile.py [SH|MLP_Fea] [frames]"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import fewtensorf_b200 as ft  # noqa: E402
from fewtensorf_b200._abi import lib  # noqa: E402

shading = sys.argv[1] if len(sys.argv) > 1 else "SH"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda", 0)
model = bench.build_model(dev, shading=shading)
rays = bench.camera_rays(dev)
n = 13
names = [lib.ftrf_profile_stage_name(i).decode() for i in range(n)]
buf = (C.c_float * n)()
acc = [0.0] * n
tot = 0.0
S = int(model.nSamples)
step = (1 << 30) // S
counts = [0, 0]
for it in range(frames + 2):
    lib.ftrf_profile_enable(1 if it >= 2 else 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for lo in range(0, rays.shape[0], step):
        with torch.no_grad():
            model(rays[lo:lo + step], None, white_bg=True, is_train=False, N_samples=-1)
        if it >= 2:
            torch.cuda.synchronize()          # forward() returns once the counters are final; the stage events need the tail
            lib.ftrf_profile_read(buf)
            for i in range(n):
                acc[i] += max(buf[i], 0.0) / frames
            if it == 2:
                counts[0] += model._last_call.n_samples
                counts[1] += model._last_call.n_valid
    e1.record()
    torch.cuda.synchronize()
    if it >= 2:
        tot += e0.elapsed_time(e1) / frames
lib.ftrf_profile_enable(0)
print(shading, "S", S, "ms/frame (with stage events)", round(tot, 3), "valid samples", counts[0], "shaded", counts[1])
print({k: round(a, 4) for k, a in zip(names, acc) if a > 0}, "sum", round(sum(acc), 3))
