"""Bring-up aid: which shared-memory unit the tensor core reads as A(m, k) for a descriptor."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fewtensorf_b200._abi import load_test_lib  # noqa: E402
lib = load_test_lib()

lib.ftrf_debug_layout_probe.argtypes = [C.c_int, C.c_uint, C.c_uint, C.c_int, C.c_void_p, C.c_void_p]
out = torch.zeros(128 * 8, device="cuda")
for name, mn, lbo, sbo in (("K-major lbo128 sbo1024", 0, 128, 1024), ("MN-major lbo4096 sbo128", 1, 4096, 128),
                           ("MN-major lbo128 sbo4096", 1, 128, 4096), ("MN-major lbo256 sbo512", 1, 256, 512)):
    res = []
    for within in (0, 1):
        lib.ftrf_debug_layout_probe(mn, lbo, sbo, within, out.data_ptr(), None)
        torch.cuda.synchronize()
        res.append(out.cpu().view(128, 8).clone())
    print("==", name)
    for m in (0, 1, 2, 3, 4, 5, 8, 9, 32, 127):
        print(f"  m={m:3d} unit:", [int(v) for v in res[0][m]], " elem:", [int(v) for v in res[1][m]])
