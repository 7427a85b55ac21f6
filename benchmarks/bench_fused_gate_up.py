#!/usr/bin/env python
"""Fused gate_up projection (fp8_gemm -> silu_and_mul -> 1x128 block FP8 quantisation in one kernel) against
the two reference-surface ops it replaces, on the Llama-3 MLP shapes.  CUDA-graph replay, weights rotated through > L2.
    python benchmarks/bench_fused_gate_up.py [--Ms 1024,8192]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import peaks  # noqa: E402
from benchmarks.bench_gemm_sweep import time_graph  # noqa: E402
from vllm_xpu_kernels_b200 import ops  # noqa: E402

E4 = torch.float8_e4m3fn
DEV = "cuda"
SHAPES = {"llama3-8b": (4096, 14336), "llama3-70b": (8192, 28672)}      # (K = hidden, inter)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--Ms", default="512,2048,8192")
    args = ap.parse_args()
    pk = peaks()
    for name, (K, inter) in SHAPES.items():
        for M in [int(m) for m in args.Ms.split(",")]:
            a = (torch.randn(M, K, device=DEV) / 4).to(E4)
            sa = torch.rand(M, 1, device=DEV) * 0.01 + 0.001
            copies = 3
            ws = [(torch.randn(2 * inter, K, device=DEV) / 4).to(E4) for _ in range(copies)]
            sw = torch.rand(2 * inter, 1, device=DEV) * 0.01 + 0.001
            two = [lambda w=w: ops.silu_and_mul_per_block_quant(
                torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None), 128, E4) for w in ws]
            gemm = [lambda w=w: torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None) for w in ws]
            fused = [lambda w=w: ops.fp8_gemm_silu_mul_quant(a, w.t(), sa, sw, None, torch.bfloat16) for w in ws]
            t2, tg, tf = time_graph(two, 10), time_graph(gemm, 10), time_graph(fused, 10)
            flops = 2.0 * M * 2 * inter * K
            print(f"{name:10s} M={M:5d} K={K} inter={inter}: gemm {tg*1e3:8.1f} us   gemm + silu_mul_quant {t2*1e3:8.1f} us   "
                  f"fused {tf*1e3:8.1f} us ({flops/tf/1e9:7.1f} TFLOP/s, {100*flops/tf/1e9/(2*pk['bf16']):4.1f}% of fp8 peak)   "
                  f"x{t2/tf:.2f}", flush=True)
            del ws


if __name__ == "__main__":
    main()
