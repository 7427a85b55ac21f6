#!/usr/bin/env python
"""Per-kernel sweep of the quantisation ops: algorithmic GB/s vs the
measured HBM peak.  L2 is flushed between timed launches by rotating over enough
independent input/output buffer sets to exceed 2 x 126 MB.
    python benchmarks/bench_quant.py [--quick] [--json out.json]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_xpu_kernels_b200 import _loader, ops  # noqa: E402

E4 = torch.float8_e4m3fn
DEV = "cuda"


def time_rot(fns, iters=20):
    """fns: list of closures over different buffer sets; returns ms per launch."""
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


def nsets(bytes_per_set):
    return max(2, min(16, int(300e6 // max(bytes_per_set, 1)) + 1))


SCALE = 1.0


def case(name, T, H):
    bf = torch.bfloat16
    if name == "dynamic_per_token_fp8":
        nb = T * H * 3 + 4 * T
        def mk():
            x = torch.rand(T, H, dtype=bf, device=DEV); o = torch.empty(T, H, dtype=E4, device=DEV)
            s = torch.empty(T, 1, device=DEV)
            return lambda: torch.ops._C.dynamic_per_token_scaled_fp8_quant(o, x, s, None)
    elif name == "static_fp8":
        nb = T * H * 3
        def mk():
            x = torch.rand(T, H, dtype=bf, device=DEV); o = torch.empty(T, H, dtype=E4, device=DEV)
            s = torch.tensor([0.01], device=DEV)
            return lambda: torch.ops._C.static_scaled_fp8_quant(o, x, s, None)
    elif name == "group_fp8_128":
        nb = T * H * 3 + 4 * T * H // 128
        def mk():
            x = torch.rand(T, H, dtype=bf, device=DEV); o = torch.empty(T, H, dtype=E4, device=DEV)
            s = torch.empty(T, H // 128, device=DEV)
            return lambda: torch.ops._C.per_token_group_fp8_quant(x, o, s, 128, 1e-10, -448., 448., False, False, False)
    elif name == "group_mxfp4":
        nb = T * H * 2.5 + 4 * T * H // 32
        def mk():
            x = torch.rand(T, H, dtype=bf, device=DEV); o = torch.empty(T, H // 2, dtype=torch.uint8, device=DEV)
            s = torch.empty(T, H // 32, device=DEV)
            return lambda: torch.ops._C.per_token_group_quant_mxfp4(x, o, s, 32, 1e-10)
    elif name in ("rms_token_fp8", "rms_token_fp8_res"):
        res = name.endswith("res")
        nb = T * H * (7 if res else 3) + 4 * T
        def mk():
            x = torch.randn(T, H, dtype=bf, device=DEV); o = torch.empty(T, H, dtype=E4, device=DEV)
            w = torch.ones(H, dtype=bf, device=DEV); s = torch.empty(T, device=DEV)
            r = torch.randn(T, H, dtype=bf, device=DEV) if res else None
            return lambda: torch.ops._C.rms_norm_dynamic_per_token_quant(o, x, w, s, 1e-6, None, r)
    elif name == "rms_block_fp8_128":
        nb = T * H * 3 + 4 * T * H // 128
        def mk():
            x = torch.randn(T, H, dtype=bf, device=DEV); o = torch.empty(T, H, dtype=E4, device=DEV)
            w = torch.ones(H, dtype=bf, device=DEV); s = torch.empty(T, H // 128, device=DEV)
            return lambda: torch.ops._C.rms_norm_per_block_quant(o, x, w, s, 1e-6, None, None, 128, False, False)
    elif name == "rms_mxfp4":
        nb = T * H * 2.5 + 4 * T * H // 32
        def mk():
            x = torch.randn(T, H, dtype=bf, device=DEV); o = torch.empty(T, H // 2, dtype=torch.uint8, device=DEV)
            w = torch.ones(H, dtype=bf, device=DEV); s = torch.empty(T, H // 32, device=DEV)
            return lambda: torch.ops._C.rms_norm_mxfp4_quant(o, x, w, s, 1e-6, None, 32)
    elif name == "rms_static_fp8":
        nb = T * H * 3
        def mk():
            x = torch.randn(T, H, dtype=bf, device=DEV); o = torch.empty(T, H, dtype=E4, device=DEV)
            w = torch.ones(H, dtype=bf, device=DEV); s = torch.tensor([0.02], device=DEV)
            return lambda: torch.ops._C.rms_norm_static_fp8_quant(o, x, w, s, 1e-6)
    elif name == "silu_static_fp8":
        nb = T * H * 5
        def mk():
            x = torch.randn(T, 2 * H, dtype=bf, device=DEV) * SCALE; o = torch.empty(T, H, dtype=E4, device=DEV)
            s = torch.tensor([0.5], device=DEV)
            return lambda: torch.ops._C.silu_and_mul_quant(o, x, s)
    elif name == "silu_block_fp8_128":
        nb = T * H * 5 + 4 * T * H // 128
        def mk():
            x = torch.randn(T, 2 * H, dtype=bf, device=DEV) * SCALE; o = torch.empty(T, H, dtype=E4, device=DEV)
            s = torch.empty(T, H // 128, device=DEV)
            return lambda: torch.ops._C.silu_and_mul_per_block_quant(o, x, s, 128, None, False, False)
    elif name == "silu_mxfp4":
        nb = T * H * 4.5 + 4 * T * H // 32
        def mk():
            x = torch.randn(T, 2 * H, dtype=bf, device=DEV) * SCALE; o = torch.empty(T, H // 2, dtype=torch.uint8, device=DEV)
            s = torch.empty(T, H // 32, device=DEV)
            return lambda: torch.ops._C.silu_and_mul_mxfp4_quant(o, x, s, 32, 1e-10)
    elif name == "indexer_k_quant_cache":
        # (f)#3: T tokens x head_dim H (bf16) -> e4m3 codes + one fp32 scale per 128 into the paged cache
        nb = T * H * 3 + 4 * T * H // 128 + 8 * T
        def mk():
            bs = 64
            nblk = (T + bs - 1) // bs
            x = torch.randn(T, H, dtype=bf, device=DEV)
            cache = torch.zeros(nblk, bs, H + 4 * (H // 128), dtype=torch.uint8, device=DEV)
            slots = torch.randperm(nblk * bs, device=DEV)[:T]
            return lambda: torch.ops._C_cache_ops.indexer_k_quant_and_cache(x, cache, slots, 128, "ue8m0")
    else:
        raise KeyError(name)
    fns = [mk() for _ in range(nsets(nb))]
    return fns, nb


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--json", default=None)
    ap.add_argument("--silu-fast", type=int, default=-1, help="-1 auto (default), 0 accurate, 1 approximate")
    ap.add_argument("--only", default=None, help="comma-separated op names")
    ap.add_argument("--T", type=int, default=None)
    ap.add_argument("--H", default=None, help="comma-separated hidden sizes (overrides the per-op defaults)")
    ap.add_argument("--scale", type=float, default=1.0, help="multiply the random inputs (activation magnitude)")
    args = ap.parse_args()
    global SCALE
    SCALE = args.scale
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    _loader.cdll().vxk_set_silu_fast(args.silu_fast)
    row_ops = ["dynamic_per_token_fp8", "static_fp8", "group_fp8_128", "group_mxfp4", "rms_token_fp8",
               "rms_token_fp8_res", "rms_block_fp8_128", "rms_mxfp4", "rms_static_fp8"]
    act_ops = ["silu_static_fp8", "silu_block_fp8_128", "silu_mxfp4"]
    cache_ops = ["indexer_k_quant_cache"]
    Ts = [8192] if args.quick else [1, 16, 128, 1024, 4096, 8192]
    if args.T:
        Ts = [args.T]
    cache_Ts = [Ts[-1] * 64]    # tokens of a long-context prefill chunk (head_dim is only 128)
    rows = []
    for op in row_ops + act_ops + cache_ops:
        if args.only and op not in args.only.split(","):
            continue
        Hs = ([4096] if args.quick else [4096, 8192]) if op in row_ops else ([14336] if args.quick else [14336, 28672])
        if op in cache_ops:
            Hs = [128, 512]
        if args.H:
            Hs = [int(h) for h in args.H.split(",")]
        for H in Hs:
            for T in (cache_Ts if op in cache_ops else Ts):
                fns, nb = case(op, T, H)
                ms = time_rot(fns, iters=5 if T >= 1024 else 20)
                gbs = nb / ms / 1e6
                rows.append(dict(op=op, T=T, H=H, us=ms * 1e3, GBs=gbs, frac=gbs / peak, silu_fast=args.silu_fast))
                print(f"{op:24s} T={T:5d} H={H:6d} {ms*1e3:9.2f} us {gbs:8.1f} GB/s  {gbs/peak*100:5.1f}% of {peak:.0f}",
                      flush=True)
                del fns
                torch.cuda.empty_cache()
    if args.json:
        json.dump(rows, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
