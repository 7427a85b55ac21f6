#!/usr/bin/env python
"""M sweep of the scaled GEMMs over the BASELINE.json shape sets (configs 3, 4, 5):

  llama8b_fp8    per-token x per-channel FP8 W8A8, Llama-3-8B (K, N)         config 3
  llama70b_mxfp4 MXFP4 (E2M1 + E8M0 per 32), Llama-3-70B (K, N)              config 4
  dsv3_blockfp8  1x128 / 128x128 fp32 block-scaled FP8, DeepSeek-V3 (K, N)   config 5
  llama8b_w8a16  bf16 activations x per-channel FP8 weights (fp8_gemm_w8a16), Llama-3-8B (K, N);
                 tensor peak = bf16 dense, weight stream 1 B / element         SURVEY 8(f) #1

For every (shape, M) the roofline time is max(2MNK / tensor peak, algorithmic bytes / HBM peak)
(small M is bound by streaming the weight from HBM) and the line reports
TFLOP/s, GB/s and roofline time / measured time.  Weights rotate over enough copies to exceed the
126 MB L2, so the small-M numbers are HBM numbers.
    python benchmarks/bench_gemm_sweep.py [--sets a,b] [--Ms 1,16,...] [--json out.json]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import peaks  # noqa: E402
import vllm_xpu_kernels_b200._xpu_C  # noqa: F401,E402

E4 = torch.float8_e4m3fn
E8 = torch.float8_e8m0fnu
DEV = "cuda"

SETS = {
    "llama8b_fp8": [(4096, 6144), (4096, 4096), (4096, 28672), (14336, 4096)],
    "llama70b_mxfp4": [(8192, 10240), (8192, 8192), (8192, 57344), (28672, 8192)],
    "llama8b_w8a16": [(4096, 6144), (4096, 4096), (4096, 28672), (14336, 4096)],
    "dsv3_blockfp8": [(7168, 1536), (1536, 24576), (512, 32768), (16384, 7168), (7168, 36864), (18432, 7168)],
}


def make_case(kind, M, K, N, copies):
    """Returns (list of closures over `copies` independent weight sets, algorithmic bytes)."""
    fns = []
    if kind == "llama8b_fp8":
        a = (torch.randn(M, K, device=DEV) / 4).to(E4)
        sa = torch.rand(M, 1, device=DEV) * 0.01 + 0.001
        for _ in range(copies):
            w = (torch.randn(N, K, device=DEV) / 4).to(E4)
            sw = torch.rand(N, 1, device=DEV) * 0.01 + 0.001
            fns.append(lambda w=w, sw=sw: torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None))
        nbytes = M * K + K * N + 2 * M * N + 4 * (M + N)
    elif kind == "llama8b_w8a16":
        a = torch.randn(M, K, device=DEV, dtype=torch.bfloat16) / 4
        for _ in range(copies):
            w = (torch.randn(N, K, device=DEV) / 4).to(E4)
            sw = torch.rand(N, device=DEV) * 0.01 + 0.001
            fns.append(lambda w=w, sw=sw: torch.ops._xpu_C.fp8_gemm_w8a16(a, w.t(), sw, None))
        nbytes = 2 * M * K + K * N + 2 * M * N + 4 * N
    elif kind == "llama70b_mxfp4":
        a = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, device=DEV).view(torch.float4_e2m1fn_x2)
        sa = torch.randint(118, 130, (M, K // 32), dtype=torch.uint8, device=DEV).view(E8)
        for _ in range(copies):
            w = torch.randint(0, 256, (N, K // 2), dtype=torch.uint8, device=DEV).view(torch.float4_e2m1fn_x2)
            sw = torch.randint(118, 130, (N, K // 32), dtype=torch.uint8, device=DEV).view(E8)
            fns.append(lambda w=w, sw=sw: torch.ops._xpu_C.fp4_gemm(a, w.t(), sa, sw, torch.bfloat16, None))
        nbytes = M * K // 2 + K * N // 2 + 2 * M * N + (M + N) * K // 32
    else:
        a = (torch.randn(M, K, device=DEV) / 4).to(E4)
        sa = torch.rand(M, K // 128, device=DEV) + 0.5
        for _ in range(copies):
            w = (torch.randn(N, K, device=DEV) / 4).to(E4)
            sw = torch.rand(K // 128, N // 128, device=DEV) + 0.5
            fns.append(lambda w=w, sw=sw: torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None))
        nbytes = M * K + K * N + 2 * M * N + 4 * (M * K // 128 + (K // 128) * (N // 128))
    return fns, nbytes


def time_rot(fns, iters):
    for f in fns:
        f()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    n = 0
    for _ in range(iters):
        for f in fns:
            f()
            n += 1
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n


def time_graph(fns, iters):
    """The same rotation captured once in a CUDA graph and replayed: what a serving engine does with
    its decode step; takes the host-side launch path (dispatcher, output allocation, tensor-map
    encoding, ~10 us per call) out of the measurement."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        for f in fns:
            f()
        side.synchronize()
        with torch.cuda.graph(g, stream=side):
            for _ in range(iters):
                for f in fns:
                    f()
    torch.cuda.current_stream().wait_stream(side)
    g.replay()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    g.replay()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / (iters * len(fns))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--graph", action="store_true", help="time CUDA-graph replays instead of eager launches")
    ap.add_argument("--sets", default="llama8b_fp8,llama70b_mxfp4,dsv3_blockfp8")
    ap.add_argument("--Ms", default="1,16,64,256,1024,4096,8192")
    ap.add_argument("--json", default=None)
    ap.add_argument("--shapes", default=None, help="restrict every set to these K:N pairs, e.g. 28672:8192,8192:10240")
    args = ap.parse_args()
    pk = peaks()
    out = []
    for kind in args.sets.split(","):
        peak_tf = (4.0 if kind == "llama70b_mxfp4" else 1.0 if kind == "llama8b_w8a16" else 2.0) * pk["bf16"]
        for (K, N) in SETS[kind]:
            if args.shapes and f"{K}:{N}" not in args.shapes.split(","):
                continue
            wbytes = K * N // (2 if kind == "llama70b_mxfp4" else 1)
            for M in [int(m) for m in args.Ms.split(",")]:
                copies = max(1, min(8, int(300e6 // wbytes) + 1))
                fns, nbytes = make_case(kind, M, K, N, copies)
                ms = (time_graph if args.graph else time_rot)(fns, max(2, 40 // copies))
                flops = 2.0 * M * N * K
                t_roof = max(flops / (peak_tf * 1e12), nbytes / (pk["hbm"] * 1e9)) * 1e3
                bound = "tensor" if flops / (peak_tf * 1e12) > nbytes / (pk["hbm"] * 1e9) else "hbm"
                row = dict(set=kind, M=M, N=N, K=K, us=ms * 1e3, tflops=flops / ms / 1e9, gbs=nbytes / ms / 1e6,
                           bound=bound, frac=t_roof / ms)
                out.append(row)
                print(f"{kind:15s} M={M:5d} N={N:6d} K={K:6d} {ms*1e3:9.1f} us {row['tflops']:8.1f} TFLOP/s "
                      f"{row['gbs']:7.1f} GB/s  {100*row['frac']:5.1f}% of {bound} roofline", flush=True)
                del fns
                torch.cuda.empty_cache()
    if args.json:
        json.dump(out, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
