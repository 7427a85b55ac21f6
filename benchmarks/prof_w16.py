import sys, torch
sys.path.insert(0, ".")
import vllm_xpu_kernels_b200._xpu_C  # noqa
M, N, K = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16) / 4
ws = [(torch.randn(N, K, device="cuda") / 4).to(torch.float8_e4m3fn) for _ in range(3)]
sw = torch.rand(N, device="cuda") * 0.01 + 0.001
for i in range(6):
    torch.ops._xpu_C.fp8_gemm_w8a16(a, ws[i % 3].t(), sw, None)
torch.cuda.synchronize()
