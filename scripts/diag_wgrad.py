"""Bring-up aid: fused wgrad kernel vs fp64 for descriptor variants."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fewtensorf_b200._abi import check, load_test_lib  # noqa: E402
lib = load_test_lib()

lib.ftrf_test_wgrad_fused.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64,
                                      C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
g = torch.Generator(device="cuda").manual_seed(1)
for variant in (0,):
    lib.ftrf_debug_wgrad_variant(variant)
    for M, kp in ((8, 152), (16,152), (1000, 152), (100000, 152), (5000, 56), (4097,168)):
        rnd = lambda *s: torch.randn(*s, device="cuda", generator=g)
        G2, G1, H1, X = rnd(M, 128), rnd(M, 128), rnd(M, 128), rnd(M, kp)
        K = kp - 2
        dW2 = torch.zeros(128, 128, device="cuda")
        dW1 = torch.zeros(128, K, device="cuda")
        db1 = torch.zeros(128, device="cuda")
        scratch = torch.empty(128 * 48 * 4, device="cuda")
        check(lib.ftrf_test_wgrad_fused(G2.data_ptr(), H1.data_ptr(), G1.data_ptr(), X.data_ptr(), kp, kp, M, dW2.data_ptr(),
                                        dW1.data_ptr(), K, db1.data_ptr(), scratch.data_ptr(), None), "wgrad_fused")
        torch.cuda.synchronize()

        def err(out, A, B):
            ref = A.double().T @ B.double()
            scale = (A.double().abs().T @ B.double().abs()).clamp_min(1e-30)
            return float(((out.double() - ref).abs() / scale).max())
        print(f"variant {variant} M={M}: dW2 err {err(dW2, G2, H1):.3e}  dW1 err {err(dW1, G1, X[:, :K]):.3e}  "
              f"db1 err {float((db1.double() - G1.double().sum(0)).abs().max()):.3e}")
        if M == 8:
            ref = (G2.double().T @ H1.double())
            print("   dW2[0,:6]", dW2[0, :6].tolist(), "\n   ref     ", ref[0, :6].tolist())
            print("   dW2[5,:6]", dW2[5, :6].tolist(), "\n   ref     ", ref[5, :6].tolist())
lib.ftrf_debug_wgrad_variant(0)
