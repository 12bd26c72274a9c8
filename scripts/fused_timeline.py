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
print("compute warp 0 (pair: src_start x_start x_done acc1_ready epi1_done acc2_ready epi2_done), clocks since start")
for it in range(4):
    row = [int(v) - t0 for v in b[0, it * 8: it * 8 + 7]]
    if row[0] < 0:
        break
    print(it, row, "deltas", [row[i + 1] - row[i] for i in range(6)])
print("mma thread (chunk: t_begin, wait_w, wait_a) and tma issue time")
for wc in range(30):
    r = [int(v) - t0 for v in b[1, wc * 3: wc * 3 + 3]]
    if r[0] < -10**9:
        break
    print(wc, "mma begin", r[0], "w_wait", r[1] - r[0], "a_wait", r[2] - r[1], "| tma issue", int(b[2, wc]) - t0)
