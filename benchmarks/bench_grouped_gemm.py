#!/usr/bin/env python
"""MoE grouped GEMM sweep on the reference test's shapes
(tests/fused_moe/test_grouped_gemm.py:14-21: Llama-4-scout N = 5120, K = 8192, 16 experts, top-1) over the
token counts of that test and the prefill sizes between them, for the three recipes (bf16 / fp8 / mxfp4 weights,
bf16 activations).

Roofline time = max(2 * M * N * K / bf16 dense peak, algorithmic bytes / HBM peak); the bytes are the weights of
the experts that own at least one row (each read once) + activations + output.  rows_per_expert is drawn like the
reference test (top-1 of uniform scores) from a different seed per launch, so that consecutive launches hit
different experts and the small-M numbers are HBM numbers, not L2 numbers (16 experts x 84 / 42 / 21 MB).
    python benchmarks/bench_grouped_gemm.py [--graph] [--Ms 1,4,...] [--kinds bf16,fp8,mxfp4] [--json out.json]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import peaks  # noqa: E402
from benchmarks.bench_gemm_sweep import time_graph, time_rot  # noqa: E402
from vllm_xpu_kernels_b200.fused_moe_interface import cutlass_grouped_gemm_xe2  # noqa: E402

DEV = "cuda"
E4 = torch.float8_e4m3fn


def make_weights(kind, E, N, K):
    if kind == "bf16":
        return torch.randn(E, K, N, device=DEV, dtype=torch.bfloat16) / 8, None, 2.0
    if kind == "fp8":
        b = torch.empty(E, K, N, device=DEV, dtype=E4)
        for e in range(E):
            b[e] = (torch.randn(K, N, device=DEV) / 4).to(E4)
        return b, torch.rand(E, device=DEV) + 0.5, 1.0
    if kind == "fp8_nk":      # extension: fp8 weights stored [E, N, K], handed over as the transposed view
        b = torch.empty(E, N, K, device=DEV, dtype=E4)
        for e in range(E):
            b[e] = (torch.randn(N, K, device=DEV) / 4).to(E4)
        return b.transpose(1, 2), torch.rand(E, device=DEV) + 0.5, 1.0
    b = torch.randint(0, 256, (E, N, K // 2), dtype=torch.uint8, device=DEV).view(torch.float4_e2m1fn_x2)
    sc = torch.randint(118, 130, (E, N, K // 32), dtype=torch.uint8, device=DEV)
    return b, sc, 0.5 + 1.0 / 32


def rows_top1(m, E, seed):
    g = torch.Generator().manual_seed(seed)
    idx = torch.rand(m, E, generator=g).argmax(dim=1)
    return torch.bincount(idx, minlength=E).to(torch.int32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--Ms", default="1,4,16,64,256,1024,4096,8192")
    ap.add_argument("--kinds", default="bf16,fp8,mxfp4")
    ap.add_argument("--E", type=int, default=16)
    ap.add_argument("--N", type=int, default=5120)
    ap.add_argument("--K", type=int, default=8192)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    pk = peaks()
    E, N, K = args.E, args.N, args.K
    out = []
    for kind in args.kinds.split(","):
        b, sc, wbytes_per = make_weights(kind, E, N, K)
        bias = torch.randn(E, N, device=DEV, dtype=torch.bfloat16)
        for M in [int(m) for m in args.Ms.split(",")]:
            a = torch.randn(M, K, device=DEV, dtype=torch.bfloat16) / 8
            d = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
            copies = 8 if M < 256 else 2
            fns, act = [], 0.0
            for i in range(copies):
                rows = rows_top1(M, E, 100 + i)
                act += float((rows > 0).sum())
                r = rows.to(DEV)
                fns.append(lambda r=r: cutlass_grouped_gemm_xe2(a, b, sc, bias, d, r, N, K, E))
            act /= copies
            ms = (time_graph if args.graph else time_rot)(fns, max(2, 40 // copies))
            flops = 2.0 * M * N * K
            nbytes = act * N * K * wbytes_per + 2.0 * M * K + 2.0 * M * N
            t_roof = max(flops / (pk["bf16"] * 1e12), nbytes / (pk["hbm"] * 1e9)) * 1e3
            bound = "tensor" if flops / (pk["bf16"] * 1e12) > nbytes / (pk["hbm"] * 1e9) else "hbm"
            rec = dict(kind=kind, M=M, N=N, K=K, E=E, active_experts=act, us=ms * 1e3, tflops=flops / ms / 1e9,
                       gbps=nbytes / ms / 1e6, frac=t_roof / ms, bound=bound)
            out.append(rec)
            print(f"moe_{kind:6s} M={M:5d} N={N} K={K} E={E} active={act:4.1f}  {ms*1e3:9.1f} us  "
                  f"{rec['tflops']:8.1f} TFLOP/s  {rec['gbps']:7.1f} GB/s  {100*rec['frac']:5.1f}% of {bound} roofline",
                  flush=True)
        del b, sc
        torch.cuda.empty_cache()
    if args.json:
        json.dump(out, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
