import torch, sys
sys.path.insert(0, '.')
from tests.test_gpu_gemm3x import run_gemm, scale_err
g = torch.Generator(device="cuda").manual_seed(1)
for (M,N,K) in [(128,128,32),(128,128,8),(256,128,152)]:
    A = torch.randn(M, K, device="cuda", generator=g); B = torch.randn(N, K, device="cuda", generator=g)*0.1
    ref = A.double() @ B.double().T
    for variant in (0,1):
        out = run_gemm(A,B,variant=variant)
        print(M,N,K,'variant',variant,'err',scale_err(out,ref,A,B),'fp32 err',scale_err(A@B.T,ref,A,B), 'nan', int(torch.isnan(out).sum()), flush=True)
