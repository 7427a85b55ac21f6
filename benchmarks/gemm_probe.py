#!/usr/bin/env python
"""Sustained-load probe of one GEMM shape: TFLOP/s next to the SM clock and board power sampled
while it runs (the tensor roofline on this box is power-capped: clocks under load sit well below
the boost clock), plus cuBLAS bf16 / cuBLASLt FP8 on the same shape as the library yardstick.
    python benchmarks/gemm_probe.py [M N K] [--secs 1.5]
"""
import argparse
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import vllm_xpu_kernels_b200._xpu_C  # noqa: F401,E402

E4 = torch.float8_e4m3fn
E8 = torch.float8_e8m0fnu


class Sampler(threading.Thread):
    def __init__(self):
        super().__init__(daemon=True)
        import pynvml
        pynvml.nvmlInit()
        self.nv = pynvml
        self.h = pynvml.nvmlDeviceGetHandleByIndex(0)
        self.stop = threading.Event()
        self.clk, self.pw = [], []

    def run(self):
        while not self.stop.is_set():
            self.clk.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
            self.pw.append(self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            time.sleep(0.01)


def sustained(fn, secs):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s = Sampler()
    s.start()
    n = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    while time.perf_counter() - t0 < secs:
        for _ in range(10):
            fn()
        n += 10
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    s.stop.set()
    s.join()
    ms = e0.elapsed_time(e1) / n
    half = s.clk[len(s.clk) // 2:] or [0]
    halfp = s.pw[len(s.pw) // 2:] or [0]
    return ms, sorted(half)[len(half) // 2], sorted(halfp)[len(halfp) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("shape", nargs="*", type=int, default=[8192, 8192, 8192])
    ap.add_argument("--secs", type=float, default=1.5)
    ap.add_argument("--sf-const", type=int, default=-1, help="all MX scale bytes = this value (default: random 118..129)")
    ap.add_argument("--zero-data", action="store_true", help="all-zero FP4 operand bytes")
    ap.add_argument("--kinds", default="bf16_cublas,fp8_cublaslt,fp8_w8a8,mxfp8_block,mxfp4")
    args = ap.parse_args()
    M, N, K = args.shape
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    flops = 2.0 * M * N * K
    for kind in args.kinds.split(","):
        if kind == "bf16_cublas":
            a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
            w = torch.randn(N, K, device=dev, dtype=torch.bfloat16)
            fn = lambda: torch.matmul(a, w.t())  # noqa: E731
        elif kind == "fp8_cublaslt":
            a = (torch.randn(M, K, device=dev) / 4).to(E4)
            w = (torch.randn(N, K, device=dev) / 4).to(E4)
            sa = torch.rand(M, 1, device=dev) * 0.01 + 0.001
            sw = torch.rand(1, N, device=dev) * 0.01 + 0.001
            fn = lambda: torch._scaled_mm(a, w.t(), scale_a=sa, scale_b=sw, out_dtype=torch.bfloat16)  # noqa: E731
        elif kind == "fp8_w8a8":
            a = (torch.randn(M, K, device=dev) / 4).to(E4)
            w = (torch.randn(N, K, device=dev) / 4).to(E4)
            sa = torch.rand(M, 1, device=dev) * 0.01 + 0.001
            sw = torch.rand(N, 1, device=dev) * 0.01 + 0.001
            wt = w.t()
            fn = lambda: torch.ops._xpu_C.fp8_gemm(a, wt, torch.bfloat16, sa, sw, None)  # noqa: E731
        elif kind == "fp8_block_f32":
            a = (torch.randn(M, K, device=dev) / 4).to(E4)
            w = (torch.randn(N, K, device=dev) / 4).to(E4)
            sa = torch.rand(M, K // 128, device=dev) * 0.01 + 0.001
            sw = torch.rand(K // 128, (N + 127) // 128, device=dev) * 0.01 + 0.001
            wt = w.t()
            fn = lambda: torch.ops._xpu_C.fp8_gemm(a, wt, torch.bfloat16, sa, sw, None)  # noqa: E731
        elif kind == "mxfp8_block":
            a = (torch.randn(M, K, device=dev) / 4).to(E4)
            w = (torch.randn(N, K, device=dev) / 4).to(E4)
            sa = torch.randint(118, 130, (M, K // 128), dtype=torch.uint8, device=dev).view(E8)
            sw = torch.randint(118, 130, (K // 128, N // 128), dtype=torch.uint8, device=dev).view(E8)
            wt = w.t()
            fn = lambda: torch.ops._xpu_C.fp8_gemm(a, wt, torch.bfloat16, sa, sw, None)  # noqa: E731
        else:
            hi = 1 if args.zero_data else 256
            a = torch.randint(0, hi, (M, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
            w = torch.randint(0, hi, (N, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
            lo_s, hi_s = (118, 130) if args.sf_const < 0 else (args.sf_const, args.sf_const + 1)
            sa = torch.randint(lo_s, hi_s, (M, K // 32), dtype=torch.uint8, device=dev).view(E8)
            sw = torch.randint(lo_s, hi_s, (N, K // 32), dtype=torch.uint8, device=dev).view(E8)
            wt = w.t()
            fn = lambda: torch.ops._xpu_C.fp4_gemm(a, wt, sa, sw, torch.bfloat16, None)  # noqa: E731
        try:
            ms, clk, pw = sustained(fn, args.secs)
            print(f"{kind:13s} M={M} N={N} K={K} {ms:8.3f} ms {flops/ms/1e9:8.1f} TFLOP/s  sm_clk {clk} MHz  {pw:6.1f} W",
                  flush=True)
        except Exception as exc:
            print(f"{kind:13s} failed: {type(exc).__name__}: {str(exc)[:200]}", flush=True)
        del a, w
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
