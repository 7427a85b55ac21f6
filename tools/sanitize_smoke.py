"""One small call of every kernel family, for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python tools/sanitize_smoke.py
Shapes are chosen to reach each kernel (row / flat / generic quantisation kernels, single-CTA, CTA-pair,
fp32-block, skinny and w8a16 GEMMs, cache ops) with ragged edges; results are not checked here (tests/ do that).
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_xpu_kernels_b200 import ops  # noqa: E402

E4, E8 = torch.float8_e4m3fn, torch.float8_e8m0fnu
DEV = "cuda"


def main():
    torch.manual_seed(0)
    bf = torch.bfloat16
    n = 0
    for T, H in ((3, 4096), (130, 1000), (17, 7168)):          # vector path, generic path (H % 8 != 0 / odd), tail
        x = torch.randn(T, H, dtype=bf, device=DEV)
        w = torch.ones(H, dtype=bf, device=DEV)
        ops.scaled_fp8_quant(x, use_per_token_if_dynamic=True); n += 1
        ops.scaled_fp8_quant(x); n += 2
        ops.scaled_fp8_quant(x, torch.tensor([0.02], device=DEV)); n += 1
        ops.rms_norm_dynamic_per_token_quant(x, w, 1e-6, E4, None, torch.randn_like(x)); n += 1
        ops.rms_norm_static_fp8_quant(x, w, torch.tensor([0.02], device=DEV), 1e-6); n += 1
        if H % 128 == 0:
            ops.per_token_group_quant_fp8(x, 128); n += 1
            ops.rms_norm_per_block_quant(x, w, 1e-6, 128, E4); n += 1
            ops.per_token_group_quant_mxfp4_sf(x); n += 1
            ops.rms_norm_mxfp4_quant_sf(x, w, 1e-6); n += 1
            g = torch.randn(T, 2 * H, dtype=bf, device=DEV)
            out = torch.empty(T, H, dtype=E4, device=DEV)
            ops.silu_and_mul_quant(out, g, torch.tensor([0.5], device=DEV)); n += 1
            ops.silu_and_mul_per_block_quant(g, 128, E4); n += 1
            ops.silu_and_mul_mxfp4_quant_sf(g); n += 1
    # GEMMs: (M, N, K)
    for M, N, K in ((1, 512, 1024), (100, 384, 512), (300, 640, 1024), (520, 1000, 2048)):
        a = (torch.randn(M, K, device=DEV) / 4).to(E4)
        wq = (torch.randn(N, K, device=DEV) / 4).to(E4)
        ops.fp8_gemm(a, wq.t(), bf, torch.rand(M, 1, device=DEV), torch.rand(N, 1, device=DEV), None); n += 1
        ops.fp8_gemm_w8a16(torch.randn(M, K, dtype=bf, device=DEV), wq.t(), torch.rand(N, device=DEV), None); n += 1
        if N % 128 == 0 and K % 128 == 0:
            ops.fp8_gemm(a, wq.t(), bf, torch.rand(M, K // 128, device=DEV) + 0.5,
                         torch.rand(K // 128, N // 128, device=DEV) + 0.5, None); n += 1
            sa = torch.randint(120, 130, (M, K // 128), dtype=torch.uint8, device=DEV).view(E8)
            sw = torch.randint(120, 130, (K // 128, N // 128), dtype=torch.uint8, device=DEV).view(E8)
            ops.fp8_gemm(a, wq.t(), bf, sa, sw, None); n += 1
        x = torch.randn(M, K, dtype=bf, device=DEV)
        wb = torch.randn(N, K, dtype=bf, device=DEV) * 0.05
        xq, xs, xsf = ops.per_token_group_quant_mxfp4_sf(x)
        wq4, ws4 = ops.per_token_group_quant_mxfp4(wb)
        ops.fp4_gemm(xq, wq4.t(), xs, ws4, bf); n += 3
        ops.fp4_gemm_sf(xq, wq4.t(), xs, ws4, bf, None, xsf, ops.mx_sf_pack(ws4)); n += 2
    # cache ops
    T, H, bs = 100, 128, 64
    k = torch.randn(T, H, dtype=bf, device=DEV)
    cache = torch.zeros(4, bs, H + 4, dtype=torch.uint8, device=DEV)
    slots = torch.randperm(4 * bs, device=DEV)[:T]
    ops.indexer_k_quant_and_cache(k, cache, slots, 128, "ue8m0"); n += 1
    dst_k = torch.empty(T, H, dtype=torch.uint8, device=DEV)
    dst_s = torch.empty(T, 4, dtype=torch.uint8, device=DEV)
    bt = torch.arange(4, dtype=torch.int32, device=DEV).view(1, 4)
    ops.cp_gather_indexer_k_quant_cache(cache, dst_k, dst_s, bt, torch.tensor([0, T], dtype=torch.int32, device=DEV)); n += 1
    src = torch.randn(4, 64, 32, dtype=bf, device=DEV)
    dst = torch.empty(4, 64, 32, dtype=torch.uint8, device=DEV)
    ops.convert_fp8(dst, src, 2.0, "fp8"); n += 1
    # MoE grouped GEMM: 16-bit / fp8 (shared-memory operand) / mxfp4 (TMEM operand and, at 256-row tiles, shared memory)
    from vllm_xpu_kernels_b200.fused_moe_interface import cutlass_grouped_gemm_xe2
    for rows in ([5, 0, 40, 19], [300, 290, 310, 324]):
        E, N, K = 4, 264, 256
        r = torch.tensor(rows, dtype=torch.int32, device=DEV)
        a = torch.randn(sum(rows) + 2, K, dtype=bf, device=DEV)
        d = torch.empty(a.shape[0], N, dtype=bf, device=DEV)
        bias = torch.randn(E, N, dtype=bf, device=DEV)
        cutlass_grouped_gemm_xe2(a, torch.randn(E, K, N, dtype=bf, device=DEV), None, bias, d, r, N, K, E); n += 1
        cutlass_grouped_gemm_xe2(a, torch.randn(E, K, N, device=DEV).to(E4), torch.rand(E, device=DEV) + 0.5, bias, d, r,
                                 N, K, E); n += 1
        b4 = torch.randint(0, 255, (E, N, K // 2), dtype=torch.uint8, device=DEV).view(torch.float4_e2m1fn_x2)
        s4 = torch.randint(110, 130, (E, N, K // 32), dtype=torch.uint8, device=DEV)
        cutlass_grouped_gemm_xe2(a, b4, s4, bias, d, r, N, K, E); n += 1
    torch.cuda.synchronize()
    print(f"sanitize_smoke: {n} op calls completed")


if __name__ == "__main__":
    main()
