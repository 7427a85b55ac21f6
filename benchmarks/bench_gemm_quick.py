#!/usr/bin/env python
"""GEMM-only benchmark (same code path as the `gemm` section of bench.py)."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import peaks            # noqa: E402
from bench_gemm import bench_gemms  # noqa: E402
import vllm_xpu_kernels_b200._xpu_C  # noqa: F401,E402
M = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
r = bench_gemms(torch.device("cuda", 0), 0, 1, peaks(), steps=10, m_tokens=M)
for row in r["rows"]:
    print(f"{row['kind']:9s} M={row['M']:5d} N={row['N']:6d} K={row['K']:6d} {row['ms']:8.3f} ms {row['tflops']:8.1f} TFLOP/s "
          f"{100*row['frac_of_dense_peak']:5.1f}% of {row['peak_tflops']:.0f}")
print("tc/fallback calls", r["tensor_core_calls"], r["fallback_calls"])
