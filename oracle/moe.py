"""CPU oracle for the MoE grouped GEMM.  TEST INFRASTRUCTURE ONLY.

Restates the torch references the reference repository's own test uses to judge
``torch.ops._xpu_C.cutlass_grouped_gemm_interface`` (tests/fused_moe/test_grouped_gemm.py):
per expert, rows of A in expert order times the dequantised weight, fp32 matmul, + bias, rounded once to
the activation dtype.  Pinned by tests/golden/f4_grouped_gemm.pt (tests/test_oracle_golden.py).
"""
import torch

E2M1_TABLE = [0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0, -0.0, -0.5, -1.0, -1.5, -2.0, -3.0, -4.0, -6.0]


def dequantize_mxfp4(qweight: torch.Tensor, scales: torch.Tensor, group_size: int, dtype: torch.dtype) -> torch.Tensor:
    """tests/fused_moe/test_grouped_gemm.py:341-377.  qweight [N, K/2] uint8 (even k in the low nibble),
    scales [N, K/group] uint8 E8M0; table value and 2^(s-127) are both taken in `dtype`
    ((scale << 7) viewed as bf16, then cast) and multiplied in `dtype`."""
    n, kh = qweight.shape
    k = kh * 2
    q = qweight.to(torch.int32)
    codes = torch.stack([q & 0xF, (q >> 4) & 0xF], dim=-1).reshape(n, k)
    table = torch.tensor(E2M1_TABLE, dtype=dtype)
    vals = table[codes.long()]
    exp_scales = scales.repeat_interleave(group_size, dim=1)
    dst_scale = (exp_scales.to(torch.int32) << 7).to(torch.uint16).view(torch.bfloat16).to(dtype)
    return (vals * dst_scale).to(dtype)


def dequantize_fp8(w_fp8: torch.Tensor, scale: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """tests/fused_moe/test_grouped_gemm.py:186-189: (fp8 -> fp32) * scale -> dtype, per expert."""
    return (w_fp8.to(torch.float32) * scale).to(dtype)


def grouped_gemm_ref(a: torch.Tensor, weights_kn, bias, rows_per_expert) -> torch.Tensor:
    """tests/fused_moe/test_grouped_gemm.py:131-146 (and :202-218, :311-327, :424-442): `weights_kn[e]` is the
    dequantised [K, N] weight of expert e in the activation dtype; fp32 matmul, + bias in fp32, one rounding.
    Rows past sum(rows_per_expert) are not part of the reference output and come back as zeros."""
    out = torch.zeros(a.shape[0], weights_kn[0].shape[1], dtype=a.dtype)
    pre = 0
    for e, r in enumerate(int(x) for x in rows_per_expert):
        if r == 0:
            continue
        acc = a[pre:pre + r].to(torch.float32) @ weights_kn[e].to(torch.float32)
        if bias is not None:
            acc += bias[e]
        out[pre:pre + r] = acc.to(a.dtype)
        pre += r
    return out


def grouped_gemm(a, b, scales, bias, rows_per_expert, kind: str) -> torch.Tensor:
    """The op as the caller sees it: kind in {"bf16", "fp8", "mxfp4"} (fused_moe_interface.py:32-45).
    b: [E, K, N] (bf16 / fp8) or uint8 [E, N, K/2] (mxfp4)."""
    dtype = a.dtype
    E = b.shape[0]
    if kind == "bf16":
        w = [b[e] for e in range(E)]
    elif kind == "fp8":
        w = [dequantize_fp8(b[e], scales[e], dtype) for e in range(E)]
    elif kind == "mxfp4":
        w = [dequantize_mxfp4(b[e], scales[e], 32, dtype).t() for e in range(E)]
    else:
        raise ValueError(kind)
    return grouped_gemm_ref(a, w, bias, rows_per_expert)
