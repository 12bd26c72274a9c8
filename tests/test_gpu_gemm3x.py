"""`-m gpu`: the tcgen05 3xTF32 contraction against an fp64 matmul (torch fp32 reference beside it).
Tolerance: 3xTF32 keeps ~2^-21 per product, so the result must be within 4e-6 of fp64 relative to
the row scale |A||B| — the same order as a plain fp32 matmul, and 100x tighter than TF32 (1e-3)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def run_gemm(A, B, bias=None, mask=None, relu=False, transposed_b=False, variant=0, ka=None):
    from fewtensorf_b200._abi import check
    from tests.gpu_util import test_lib
    lib = test_lib()
    M, lda = A.shape
    if transposed_b:          # B given as [K][N]; element (n,k) at B[k*N + n]
        K, N = B.shape
        sn, sk = 1, N
    else:
        N, K = B.shape
        sn, sk = K, 1
    out = torch.full((M, N), float("nan"), device=A.device)
    check(lib.ftrf_test_gemm3x(A.data_ptr(), lda, lda if ka is None else ka, M, B.data_ptr(), sn, sk, N, K,
                               None if bias is None else bias.data_ptr(),
                               None if mask is None else mask.data_ptr(), 0 if mask is None else mask.shape[1],
                               int(relu), out.data_ptr(), N, variant, torch.cuda.current_stream().cuda_stream), "gemm3x")
    torch.cuda.synchronize()
    return out


def scale_err(out, ref, A, B):
    scale = (A.double().abs() @ B.double().abs().T).clamp_min(1e-30)
    return float(((out.double() - ref).abs() / scale).max())


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (1000, 128, 152), (333, 128, 128), (4097, 160, 128), (77, 64, 56)])
def test_gemm3x_matches_fp64(M, N, K):
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    lda = (K + 3) // 4 * 4
    A = torch.randn(M, lda, device="cuda", generator=g)
    A[:, K:] = 0
    B = torch.randn(N, K, device="cuda", generator=g) * 0.1
    out = run_gemm(A, B)
    ref = A[:, :K].double() @ B.double().T
    assert torch.isfinite(out).all()
    assert scale_err(out, ref, A[:, :K], B) < 4e-6
    fp32 = scale_err(A[:, :K] @ B.T, ref, A[:, :K], B)
    assert scale_err(out, ref, A[:, :K], B) < max(20 * fp32, 4e-6)


def test_gemm3x_epilogues_and_strided_b():
    g = torch.Generator(device="cuda").manual_seed(5)
    M, N, K = 700, 128, 128
    A = torch.randn(M, K, device="cuda", generator=g)
    W = torch.randn(K, N, device="cuda", generator=g) * 0.1          # used transposed: out = A @ W
    bias = torch.randn(N, device="cuda", generator=g)
    out = run_gemm(A, W, bias=bias, relu=True, transposed_b=True)
    ref = torch.relu(A.double() @ W.double() + bias.double())
    assert float((out.double() - ref).abs().max()) < 2e-5
    mask = torch.randn(M, N, device="cuda", generator=g)
    out = run_gemm(A, W, mask=mask, transposed_b=True)
    ref = torch.where(mask > 0, (A.double() @ W.double()), torch.zeros((), dtype=torch.float64, device="cuda"))
    assert float((out.double() - ref).abs().max()) < 2e-5


@pytest.mark.parametrize("M,K,with_bias", [(32, 128, False), (1000, 150, True), (5000, 128, True), (100000, 150, True)])
def test_wgrad3x_matches_fp64(M, K, with_bias):
    from fewtensorf_b200._abi import check
    from tests.gpu_util import test_lib
    lib = test_lib()
    g = torch.Generator(device="cuda").manual_seed(M + K)
    lda = (K + 7) // 8 * 8
    G = torch.randn(M, 128, device="cuda", generator=g) * (torch.rand(M, 128, device="cuda", generator=g) > 0.5)
    A = torch.randn(M, lda, device="cuda", generator=g)
    dW = torch.zeros(128, K, device="cuda")
    db = torch.zeros(128, device="cuda") if with_bias else None
    check(lib.ftrf_test_wgrad3x(G.data_ptr(), 128, A.data_ptr(), lda, K, M, dW.data_ptr(), K,
                                None if db is None else db.data_ptr(), torch.cuda.current_stream().cuda_stream), "wgrad3x")
    torch.cuda.synchronize()
    ref = G.double().T @ A[:, :K].double()
    scale = (G.double().abs().T @ A[:, :K].double().abs()).clamp_min(1e-30)
    assert float(((dW.double() - ref).abs() / scale).max()) < 4e-6
    if with_bias:
        refb = G.double().sum(0)
        assert float((db.double() - refb).abs().max() / G.double().abs().sum(0).max()) < 4e-6


@pytest.mark.parametrize("M,kp", [(16, 152), (17, 152), (1000, 152), (5000, 56), (100000, 152), (4097, 168)])
def test_wgrad_fused_matches_fp64(M, kp):
    """Both MLP weight gradients + the bias gradient from one kernel (MN-major UMMA operands)."""
    from fewtensorf_b200._abi import check
    from tests.gpu_util import test_lib
    lib = test_lib()
    lib.ftrf_test_wgrad_fused.restype = C.c_int
    lib.ftrf_test_wgrad_fused.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64,
                                          C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    g = torch.Generator(device="cuda").manual_seed(M + kp)
    rnd = lambda *s: torch.randn(*s, device="cuda", generator=g)
    G2 = rnd(M, 128) * (torch.rand(M, 128, device="cuda", generator=g) > 0.5)
    G1 = rnd(M, 128) * (torch.rand(M, 128, device="cuda", generator=g) > 0.5)
    H1 = torch.relu(rnd(M, 128))
    X = rnd(M, kp)
    K = kp - 2                                  # the last two data columns are padding: dropped by the un-permute
    dW2 = torch.zeros(128, 128, device="cuda")
    dW1 = torch.zeros(128, K, device="cuda")
    db1 = torch.zeros(128, device="cuda")
    scratch = torch.empty(128 * 48 * 4, device="cuda")
    check(lib.ftrf_test_wgrad_fused(G2.data_ptr(), H1.data_ptr(), G1.data_ptr(), X.data_ptr(), kp, kp, M, dW2.data_ptr(),
                                    dW1.data_ptr(), K, db1.data_ptr(), scratch.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream), "wgrad_fused")
    torch.cuda.synchronize()

    def err(out, A, B):
        ref = A.double().T @ B.double()
        scale = (A.double().abs().T @ B.double().abs()).clamp_min(1e-30)
        return float(((out.double() - ref).abs() / scale).max())
    assert err(dW2, G2, H1) < 4e-6
    assert err(dW1, G1, X[:, :K]) < 4e-6
    assert float((db1.double() - G1.double().sum(0)).abs().max() / G1.double().abs().sum(0).max()) < 4e-6
