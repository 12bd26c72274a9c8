"""Debug aid: cycles per tcgen05.mma.kind::tf32 dispatch (M=128, K=8) by operand placement."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fewtensorf_b200._abi import load_test_lib  # noqa: E402
lib = load_test_lib()

lib.ftrf_debug_mma_probe.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
out = torch.zeros(2, dtype=torch.int64, device="cuda")
names = ["SS N=128", "SS N=256", "TS N=128", "TS N=256", "SS N=128 B MN-major", "SS N=64", "SS N=128 SW128", "SS N=256 SW128", "TS N=128 B SW128"]
for ctas in (148,):
    for mode, name in enumerate(names):
        for it in range(2):
            lib.ftrf_debug_mma_probe(mode, 2000, ctas, out.data_ptr(), None)
            torch.cuda.synchronize()
        issue, total = [int(v) for v in out.cpu()]
        print(f"ctas={ctas:3d} {name:22s} issue {issue / 2000:7.1f} clk/mma   complete {total / 2000:7.1f} clk/mma", flush=True)
