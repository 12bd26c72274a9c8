"""Debug aid: what the MMA-issuing thread of a fused kernel pays per operand chunk besides the dispatches themselves."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fewtensorf_b200._abi import load_test_lib  # noqa: E402
lib = load_test_lib()
lib.ftrf_debug_issue_probe.argtypes = [C.c_int] * 6 + [C.c_void_p, C.c_void_p]
out = torch.zeros(2, dtype=torch.int64, device="cuda")
G = 400
out = torch.zeros(4, dtype=torch.int64, device="cuda")
print("noise (0 none, 1 ALU, 2 tcgen05.st/ld, 3 ld.shared, 4 sleeping warps) per_group | clk per group, per MMA")
for noise in (0, 1, 2, 3, 4):
    for per_group in (12, 24):
        for _ in range(2):
            assert lib.ftrf_debug_issue_probe(G, 0, 0, 1 | (noise << 4), per_group, 148, out.data_ptr(), None) == 0
            torch.cuda.synchronize()
        a, b = (int(v) for v in out.cpu()[:2])
        print(f"{noise} {per_group:3d} | {b / G:8.1f}  {b / G / per_group:6.1f}")
