"""Debug aid: stage times of the gather backward kernels against the run length of the run-merging variants."""
import ctypes as C
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
n = 16
names = [lib.ftrf_profile_stage_name(i).decode() for i in range(n)]
buf = (C.c_float * n)()
flush = torch.empty(512 << 18, dtype=torch.float32, device=dev)
for which in sys.argv[1:] or ["density"]:
    for v in ((0, 8, 11, 12, 16) if which == "density" else (-2, 8, 10, 13, 16, 22, -1)):
        if which == "density":
            lib.ftrf_debug_set_runs(v)
        else:
            lib.ftrf_debug_set_app_runs(v)
        acc = [0.0] * n
        for it in range(13):
            flush.zero_()
            lib.ftrf_profile_enable(1 if it >= 3 else 0)
            m.train_step(rays, target, None, white_bg=True, N_samples=1039)
            torch.cuda.synchronize()
            if it >= 3:
                lib.ftrf_profile_read(buf)
                for i in range(n):
                    acc[i] += max(buf[i], 0.0) / 10
        lib.ftrf_profile_enable(0)
        print(which, "run", v, {k: round(a, 4) for k, a in zip(names, acc) if k in ("density_bwd", "app_bwd")})
