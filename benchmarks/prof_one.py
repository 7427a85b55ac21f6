#!/usr/bin/env python
"""Run ONE op of the hot path a few times on rotating buffers -- the target of the ncu captures.
    python benchmarks/prof_one.py <op> [M|T] [N|H] [K]
ops: fp8 mxfp4 blockfp8 blockue8 w8a16   (GEMMs: M N K)
     any name of benchmarks/bench_quant.py (quant ops: T H)
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
from vllm_xpu_kernels_b200 import ops  # noqa: E402

E4, E8 = torch.float8_e4m3fn, torch.float8_e8m0fnu
DEV = "cuda"


def main():
    op = sys.argv[1]
    a = [int(v) for v in sys.argv[2:]]
    reps = 6
    if op in ("fp8", "mxfp4", "blockfp8", "blockue8", "w8a16"):
        M, N, K = a
        torch.manual_seed(0)
        if op == "fp8":
            xq = (torch.randn(M, K, device=DEV) / 4).to(E4)
            wq = [(torch.randn(N, K, device=DEV) / 4).to(E4) for _ in range(2)]
            xs = torch.rand(M, 1, device=DEV) * 0.01 + 0.001
            ws = torch.rand(N, 1, device=DEV) * 0.01 + 0.001
            fn = lambda i: ops.fp8_gemm(xq, wq[i % 2].t(), torch.bfloat16, xs, ws, None)
        elif op == "mxfp4":
            xq = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, device=DEV).view(torch.float4_e2m1fn_x2)
            wq = [torch.randint(0, 256, (N, K // 2), dtype=torch.uint8, device=DEV).view(torch.float4_e2m1fn_x2)
                  for _ in range(2)]
            xs = torch.randint(118, 130, (M, K // 32), dtype=torch.uint8, device=DEV).view(E8)
            ws = torch.randint(118, 130, (N, K // 32), dtype=torch.uint8, device=DEV).view(E8)
            fn = lambda i: ops.fp4_gemm(xq, wq[i % 2].t(), xs, ws, torch.bfloat16)
        elif op in ("blockfp8", "blockue8"):
            xq = (torch.randn(M, K, device=DEV) / 4).to(E4)
            wq = [(torch.randn(N, K, device=DEV) / 4).to(E4) for _ in range(2)]
            if op == "blockue8":
                xs = torch.randint(118, 130, (M, K // 128), dtype=torch.uint8, device=DEV).view(E8)
                ws = torch.randint(118, 130, (K // 128, N // 128), dtype=torch.uint8, device=DEV).view(E8)
            else:
                xs = torch.rand(M, K // 128, device=DEV) * 0.01 + 0.001
                ws = torch.rand(K // 128, N // 128, device=DEV) * 0.01 + 0.001
            fn = lambda i: ops.fp8_gemm(xq, wq[i % 2].t(), torch.bfloat16, xs, ws, None)
        else:
            x = torch.randn(M, K, device=DEV, dtype=torch.bfloat16) / 4
            wq = [(torch.randn(N, K, device=DEV) / 4).to(E4) for _ in range(2)]
            ws = torch.rand(N, device=DEV) * 0.01 + 0.001
            fn = lambda i: ops.fp8_gemm_w8a16(x, wq[i % 2].t(), ws, None)
        for i in range(reps):
            fn(i)
    else:
        import bench_quant
        T, H = a
        fns, _ = bench_quant.case(op, T, H)
        for i in range(reps):
            fns[i % len(fns)]()
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
