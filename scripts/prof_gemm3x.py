"""Stand-alone timing of the tcgen05 contractions at the benchmark's shapes (V_a ~ 102k shaded samples)."""
import sys, torch
sys.path.insert(0, '.')
from tests.test_gpu_gemm3x import run_gemm
from fewtensorf_b200._abi import check, load_test_lib
lib = load_test_lib()
M = 102186
g = torch.Generator(device="cuda").manual_seed(1)
X = torch.randn(M, 152, device="cuda", generator=g); X[:, 150:] = 0
W1 = torch.randn(128, 150, device="cuda", generator=g) * 0.1
H = torch.randn(M, 128, device="cuda", generator=g)
W2 = torch.randn(128, 128, device="cuda", generator=g) * 0.1
b = torch.randn(128, device="cuda", generator=g)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, n=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); return ts[len(ts) // 2]
print("L1  fwd (K=150,N=128, bias+relu) us:", timeit(lambda: run_gemm(X, W1, bias=b, relu=True, ka=152)))
print("L2  fwd (K=128,N=128)            us:", timeit(lambda: run_gemm(H, W2, bias=b, relu=True)))
print("dgrad2 (mask, transposed B)      us:", timeit(lambda: run_gemm(H, W2, mask=H, transposed_b=True)))
W1t = W1.t().contiguous()     # [150][128] viewed as transposed_b => out = H @ W1  -> [M][150]... needs K=128
print("dgrad1 (N=150, transposed B)     us:", timeit(lambda: run_gemm(H, W1, transposed_b=True)))
dW = torch.zeros(128, 150, device="cuda"); db = torch.zeros(128, device="cuda")
st = torch.cuda.current_stream().cuda_stream
print("wgrad1 (K=150)                   us:", timeit(lambda: check(lib.ftrf_test_wgrad3x(H.data_ptr(), 128, X.data_ptr(), 152, 150, M, dW.data_ptr(), 150, db.data_ptr(), st))))
dW2 = torch.zeros(128, 128, device="cuda")
print("wgrad2 (K=128)                   us:", timeit(lambda: check(lib.ftrf_test_wgrad3x(H.data_ptr(), 128, H.data_ptr(), 128, 128, M, dW2.data_ptr(), 128, None, st))))
