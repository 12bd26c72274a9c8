"""Debug aid: phase timeline (SM clocks) of CTA 0 of the fused shading forward kernel."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fewtensorf_b200._abi import load_test_lib  # noqa: E402
lib = load_test_lib()
import fewtensorf_b200.module as _M  # noqa: E402
_M.lib = lib          # the timeline hooks only exist in the test build: route the module's calls through it

dev = torch.device("cuda", 0)
train = len(sys.argv) > 1 and sys.argv[1] == "train"
m = bench.build_model(dev)
rays, target = bench.synthetic_rays(4096, seed=0)
rays, target = rays.to(dev), target.to(dev)
buf = torch.zeros(3 * 512, dtype=torch.int64, device=dev)
lib.ftrf_debug_fused_timeline.argtypes = [C.c_void_p]
for it in range(3):
    if it == 2:
        lib.ftrf_debug_fused_timeline(buf.data_ptr())
    if train:
        m.train_step(rays, target, None, white_bg=True, N_samples=1039)
    else:
        with torch.no_grad():
            m(rays, None, white_bg=True, is_train=False, N_samples=1039)
    torch.cuda.synchronize()
lib.ftrf_debug_fused_timeline(None)
b = buf.cpu().view(3, 512)
t0 = int(b[0, 0])
print("compute warp 0 (tile: start acc1_ready epi1_done x_done acc2_ready end), clocks since start")
for it in range(6):
    row = [int(v) - t0 for v in b[0, it * 8: it * 8 + 6]]
    if row[0] < 0 and it:
        break
    print(it, row)
for role, name in ((1, "MMA issuer"),):
    print(name, "(chunk: begin, operands ready -> waited)")
    for g in range(30):
        r = [int(v) - t0 for v in b[role, g * 2: g * 2 + 2]]
        if r[0] < -10**9:
            break
        print(" ", g, "begin", r[0], "waited", r[1] - r[0])
