import sys, torch
sys.path.insert(0, '.')
from fewtensorf_b200._abi import check, load_test_lib
lib = load_test_lib()
M = 102186
g = torch.Generator(device="cuda").manual_seed(1)
X = torch.randn(M, 152, device="cuda", generator=g); X[:, 150:] = 0
W1 = torch.randn(128, 150, device="cuda", generator=g) * 0.1
b = torch.randn(128, device="cuda", generator=g)
out = torch.empty(M, 128, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def run(variant):
    check(lib.ftrf_test_gemm3x(X.data_ptr(), 152, 152, M, W1.data_ptr(), 150, 1, 128, 150, b.data_ptr(), None, 0, 1, out.data_ptr(), 128, variant, st))
def timeit(variant, n=10, fl=True):
    for _ in range(3): run(variant)
    ts = []
    for _ in range(n):
        if fl: flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(variant); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); return ts[len(ts) // 2]
for name, v in [("full", 0), ("no MMA", 2), ("no epilogue", 4), ("no A loads", 8), ("no B fill", 16), ("no MMA+epi", 6), ("no MMA+epi+A", 14), ("nothing", 30), ("only B fill", 14), ("only epilogue", 26)]:
    print(f"{name:16s} flushed {timeit(v):8.1f} us   warm {timeit(v, fl=False):8.1f} us", flush=True)
