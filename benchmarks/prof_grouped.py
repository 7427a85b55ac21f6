#!/usr/bin/env python
"""One MoE grouped-GEMM launch per recipe on the reference test's shape -- the target of the ncu captures.
    python benchmarks/prof_grouped.py <bf16|fp8|mxfp4> [M]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from benchmarks.bench_grouped_gemm import make_weights, rows_top1  # noqa: E402
from vllm_xpu_kernels_b200.fused_moe_interface import cutlass_grouped_gemm_xe2  # noqa: E402

kind = sys.argv[1]
M = int(sys.argv[2]) if len(sys.argv) > 2 else 256
E, N, K = 16, 5120, 8192
b, sc, _ = make_weights(kind, E, N, K)
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16) / 8
d = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for i in range(3):
    cutlass_grouped_gemm_xe2(a, b, sc, None, d, rows_top1(M, E, 100 + i).cuda(), N, K, E)
torch.cuda.synchronize()
