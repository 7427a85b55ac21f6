import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_xpu_kernels_b200._xpu_C  # noqa
E4 = torch.float8_e4m3fn
M, N, K = 4096, 4096, 8192
a = (torch.randn(M, K, device="cuda") / 4).to(E4)
w = (torch.randn(N, K, device="cuda") / 4).to(E4)
sa = torch.rand(M, K // 128, device="cuda") + 0.5
sw = torch.rand(K // 128, N // 128, device="cuda") + 0.5
for _ in range(3):
    out = torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None)
torch.cuda.synchronize()
