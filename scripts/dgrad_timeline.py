"""Debug aid: phase timeline (SM clocks) of CTA 0 of the fused dgrad-chain kernel."""
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
m = bench.build_model(dev)
rays, target = bench.synthetic_rays(4096, seed=0)
rays, target = rays.to(dev), target.to(dev)
buf = torch.zeros(3 * 512, dtype=torch.int64, device=dev)
lib.ftrf_debug_dgrad_timeline.argtypes = [C.c_void_p]
for it in range(3):
    if it == 2:
        lib.ftrf_debug_dgrad_timeline(buf.data_ptr())
    m.train_step(rays, target, None, white_bg=True, N_samples=1039)
    torch.cuda.synchronize()
lib.ftrf_debug_dgrad_timeline(None)
b = buf.cpu().view(3, 512)
t0 = int(b[0, 0])
for half in range(2):
    print("group", half, "(tile: start gh2_put acc1_ready gh1_put acc2_ready epi_done)")
    for it in range(7):
        row = [int(v) - t0 for v in b[half, it * 8: it * 8 + 6]]
        if row[0] < -10**9:
            break
        print(" ", it, row)
print("mma thread (chunk: begin, a_wait)")
for cc in range(32):
    r = [int(v) - t0 for v in b[2, cc * 2: cc * 2 + 2]]
    if r[0] < -10**9:
        break
    print(" ", cc, "begin", r[0], "a_wait", r[1] - r[0])
