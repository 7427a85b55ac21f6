"""CPU reference for the scaled GEMMs.
TEST INFRASTRUCTURE ONLY (see oracle/__init__).

The arithmetic of the reference's GEMMs lives in oneDNN (uxlfoundation/oneDNN
pinned 80afa71049cd69a3df32adcccb623b12cd7baa22, rls-v3.12,
/root/reference/CMakeLists.txt:53-60), which is not under /root/reference.
Its published contract for ``matmul`` with scale attributes is
``dst = (src * s_src) @ (wei * s_wei) + bias`` with fp32 accumulation; the
reference pins it at its own call sites only through
``tests/test_fp8_gemm_onednn.py`` / ``tests/test_fp4_gemm_onednn.py``
(dequantise -> torch.matmul in the output dtype, atol = rtol = 5e-2..6e-2).
``*_ref`` below is the exact-arithmetic statement of that contract (float64
accumulate, one rounding to the output dtype) that the CUDA GEMMs are held to
at atol = rtol = 1e-2; ``*_literal`` is the reference test's own formula.
"""
from __future__ import annotations

from typing import Optional

import torch

from .formats import E4M3, E8M0, e2m1_decode, unpack_nibbles

F32 = torch.float32


# --------------------------------------------------------------------------
# weight-side quantisers used by the reference tests to build GEMM operands
# --------------------------------------------------------------------------
def fp8_block_quant_2d(w: torch.Tensor, bm: int, bn: int, eps: float = 1e-6,
                       use_ue8m0: bool = False):
    """tests/ops/fp8_quant_op.py:16-78, vectorised, evaluated in w.dtype like
    the reference: s = max(amax/448, eps) (optionally 2^ceil(log2 s));
    q = e4m3(block / s).  Returns (q [M,N] e4m3, scales [M/bm, N/bn] fp32 or e8m0)."""
    M, N = w.shape
    blk = w.reshape(M // bm, bm, N // bn, bn)
    amax = blk.abs().amax(dim=(1, 3))
    s = torch.clamp(amax / 448.0, min=eps)
    if use_ue8m0:
        s = torch.exp2(torch.ceil(torch.log2(s)))
    q = (blk / s[:, None, :, None]).to(E4M3).reshape(M, N)
    s = s.to(F32)
    return q, (s.to(E8M0) if use_ue8m0 else s)


def _expand(s: torch.Tensor, rows: int, cols: int) -> torch.Tensor:
    """Broadcast a scale tensor of shape [], [1], [rows,1], [rows], [1,cols],
    [r_blocks, c_blocks] to a dense [rows, cols] fp64 tensor."""
    s = s.to(torch.float32).to(torch.float64)
    if s.numel() == 1:
        return s.reshape(1, 1).expand(rows, cols)
    if s.dim() == 1:
        s = s.reshape(-1, 1) if s.numel() == rows else s.reshape(1, -1)
    r, c = s.shape
    assert rows % r == 0 and cols % c == 0, (s.shape, rows, cols)
    return s.repeat_interleave(rows // r, 0).repeat_interleave(cols // c, 1)


# --------------------------------------------------------------------------
# a13  fp8_gemm       csrc/xpu/onednn/onednn_matmul.cpp:54-79, fp8_gemm_w8a8.h:12-184
# --------------------------------------------------------------------------
def fp8_gemm_ref(A: torch.Tensor, B: torch.Tensor, out_dtype: torch.dtype,
                 A_scale: Optional[torch.Tensor] = None,
                 B_scale: Optional[torch.Tensor] = None,
                 bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """A [M,K] fp8, B [K,N] fp8 (any strides).  Scale modes by shape
    (fp8_gemm_w8a8.h:28-41,86-134): numel 1 = per tensor; A_scale [M,1] per
    token; B_scale [N] / [N,1] per channel; A_scale [M,K/gk] with
    B_scale [K/gk, N/gn] = block (fp32 or e8m0)."""
    M, K = A.shape
    N = B.shape[1]
    a = A.to(torch.float32).to(torch.float64)
    b = B.to(torch.float32).to(torch.float64)
    if A_scale is not None:
        a = a * _expand(A_scale, M, K)
    if B_scale is not None:
        block = (A_scale is not None and A_scale.dim() == 2 and A_scale.shape[1] > 1)
        if block:
            b = b * _expand(B_scale, K, N)
        else:  # per tensor or per channel (one scale per column of B)
            sb = B_scale.to(torch.float32).to(torch.float64).reshape(-1)
            b = b * (sb.reshape(1, -1) if sb.numel() > 1 else sb)
    d = a @ b
    if bias is not None and bias.numel() > 0:
        d = d + bias.to(torch.float64)
    return d.to(out_dtype)


def fp8_gemm_literal(A_hp: torch.Tensor, W_hp: torch.Tensor) -> torch.Tensor:
    """tests/test_fp8_gemm_onednn.py:155,199,305,340: operands dequantised
    into the output dtype, then torch.matmul(A_hp, W_hp.t()) in that dtype."""
    return torch.matmul(A_hp, W_hp.t())


# --------------------------------------------------------------------------
# a14  fp4_gemm       onednn_matmul.cpp:101-123, fp4_gemm_w4a4.h:12-167
# --------------------------------------------------------------------------
def fp4_gemm_ref(A: torch.Tensor, B: torch.Tensor, A_scale: torch.Tensor,
                 B_scale: torch.Tensor, out_dtype: torch.dtype,
                 bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """A [M,K/2] packed E2M1, B [K/2,N] packed (a transposed view of [N,K/2]),
    A_scale [M,K/32] e8m0, B_scale [N,K/32] e8m0 (tests/test_fp4_gemm_onednn.py
    :65-71 convention)."""
    M = A.shape[0]
    N = B.shape[1]
    a = e2m1_decode(unpack_nibbles(A.contiguous().view(torch.uint8))).to(torch.float64)
    w = e2m1_decode(unpack_nibbles(B.t().contiguous().view(torch.uint8))).to(torch.float64)
    K = a.shape[1]
    a = a * _expand(A_scale.to(F32), M, K)
    w = w * _expand(B_scale.to(F32), N, K)
    d = a @ w.t()
    if bias is not None and bias.numel() > 0:
        d = d + bias.to(torch.float64)
    return d.to(out_dtype)


# --------------------------------------------------------------------------
# (f)#1  fp8_gemm_w8a16   onednn_matmul.cpp:81-99, fp8_gemm_w8a16.h:13-145
# --------------------------------------------------------------------------
def fp8_gemm_w8a16_ref(A: torch.Tensor, B: torch.Tensor,
                       B_scale: Optional[torch.Tensor] = None,
                       bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """A [M,K] bf16/f16, B [K,N] fp8 (any strides).  Weight scale by shape
    (fp8_gemm_w8a16.h:29-35,103-128): 2-D with size(0) > 1 = block
    [K/gs, N/gs]; numel 1 = per tensor; else per output channel ([N] or [1,N]).
    Exact statement: float64 dequantise, float64 accumulate, one rounding to
    A.dtype."""
    M, K = A.shape
    N = B.shape[1]
    a = A.to(torch.float32).to(torch.float64)
    b = B.to(torch.float32).to(torch.float64)
    if B_scale is not None and B_scale.numel() > 0:
        if B_scale.dim() == 2 and B_scale.shape[0] > 1:
            b = b * _expand(B_scale, K, N)
        else:
            sb = B_scale.to(torch.float32).to(torch.float64).reshape(-1)
            b = b * (sb.reshape(1, -1) if sb.numel() > 1 else sb)
    d = a @ b
    if bias is not None and bias.numel() > 0:
        d = d + bias.to(torch.float64)
    return d.to(A.dtype)


def fp8_block_dequant_2d(q: torch.Tensor, s: torch.Tensor, bm: int, bn: int,
                         dtype: torch.dtype) -> torch.Tensor:
    """tests/ops/fp8_quant_op.py:81-106: weight [N,K] fp8 with scales
    [N/bm, K/bn]: fp32 multiply, one rounding to `dtype`."""
    gm, gn = s.shape
    M, N = q.shape
    return (q.to(F32).reshape(gm, bm, gn, bn) * s.to(F32).reshape(gm, 1, gn, 1)).reshape(M, N).to(dtype)


def fp8_gemm_w8a16_literal(A: torch.Tensor, W_fp8: torch.Tensor, scale_nk: torch.Tensor) -> torch.Tensor:
    """tests/test_fp8_gemm_onednn.py:105-111,246-248,393-396: the weight
    W_fp8 [N,K] is dequantised INTO the activation dtype (scale broadcastable
    to [N,K], also cast to that dtype), then torch.matmul in that dtype."""
    w = W_fp8.to(A.dtype) * scale_nk.to(A.dtype)
    return torch.matmul(A, w.t())
