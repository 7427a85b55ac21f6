"""Python-side operator wrappers, mirroring the reference's
tests/register_ops.py (:29-74, :297-415) and the allocation helpers of
tests/ops/fp8_quant_op.py (:135-276) and tests/ops/mxfp4_quant_op.py (:8-63):
same names, same arguments, same scale layouts -- on CUDA tensors.

All compute goes through torch.ops._C.* / torch.ops._xpu_C.* (registered by the
compiled shim); nothing here computes on its own.
"""
from __future__ import annotations

from typing import Optional, Union

import torch

from . import _C, _xpu_C  # noqa: F401  (registers the operators; raises if not built)


# ----------------------------------------------------------------- raw ops
def static_scaled_fp8_quant(out, input, scale, group_shape=None) -> None:
    torch.ops._C.static_scaled_fp8_quant(out, input, scale, group_shape)


def dynamic_scaled_fp8_quant(out, input, scale) -> None:
    torch.ops._C.dynamic_scaled_fp8_quant(out, input, scale)


def dynamic_per_token_scaled_fp8_quant(out, input, scales, scale_ub=None) -> None:
    torch.ops._C.dynamic_per_token_scaled_fp8_quant(out, input, scales, scale_ub)


def silu_and_mul_quant(out, input, scale) -> None:
    torch.ops._C.silu_and_mul_quant(out, input, scale)


def fp8_gemm(input, weight, out_dtype, scale_act, scale_wei, bias=None):
    return torch.ops._xpu_C.fp8_gemm(input, weight, out_dtype, scale_act, scale_wei, bias)


def fp8_gemm_w8a16(input, weight, scale_wei, bias=None):
    """tests/register_ops.py:412-415 (its 4th argument, named scale_act there, is the bias slot of the schema)."""
    return torch.ops._xpu_C.fp8_gemm_w8a16(input, weight, scale_wei, bias)


def fp4_gemm(input, weight, scale_act, scale_wei, out_dtype, bias=None):
    return torch.ops._xpu_C.fp4_gemm(input, weight, scale_act, scale_wei, out_dtype, bias)


# ----------------------------------------------------------- cache ops (tests/register_ops.py:226-294)
def indexer_k_quant_and_cache(k, kv_cache, slot_mapping, quant_block_size: int, scale_fmt: str) -> None:
    torch.ops._C_cache_ops.indexer_k_quant_and_cache(k, kv_cache, slot_mapping, quant_block_size, scale_fmt)


def cp_gather_indexer_k_quant_cache(kv_cache, dst_k, dst_scale, block_table, cu_seq_lens) -> None:
    torch.ops._C_cache_ops.cp_gather_indexer_k_quant_cache(kv_cache, dst_k, dst_scale, block_table, cu_seq_lens)


def convert_fp8(dst_cache, src_cache, scale: float, kv_dtype: str = "fp8") -> None:
    torch.ops._C_cache_ops.convert_fp8(dst_cache, src_cache, scale, kv_dtype)


# ------------------------------------------------- allocation helpers
# Same call signatures, defaults and output layouts as the reference's test-side helpers
# (tests/ops/fp8_quant_op.py:135-276) so that its tests and callers run unchanged; the bodies are built on
# the one scale allocator below.
def _alloc_group_scales(rows_shape: tuple, n_groups: int, device, layout: str) -> torch.Tensor:
    """fp32 scale tensor of logical shape rows_shape + (n_groups,).
    layout "row": contiguous; "col": groups-major storage viewed back (stride 1 along rows);
    "col_tma": like "col" with the row count padded to a multiple of 4 floats (16 bytes) per column, the pitch a
    TMA descriptor over the scale columns needs."""
    if layout == "row":
        return torch.empty(rows_shape + (n_groups,), device=device, dtype=torch.float32)
    m = rows_shape[-1]
    if layout == "col":
        return torch.empty(rows_shape[:-1] + (n_groups, m), device=device, dtype=torch.float32).transpose(-1, -2)
    pitch = get_tma_aligned_size(m, 4)
    batch = rows_shape[:-1]
    strides = (1, pitch) if not batch else (pitch * n_groups, 1, pitch)
    return torch.empty_strided(batch + (m, n_groups), strides, device=device, dtype=torch.float32)


def get_tma_aligned_size(x: int, element_size: int) -> int:
    """Smallest count >= x whose byte size is a multiple of 16 (tests/ops/fp8_quant_op.py:197-203)."""
    per16 = 16 // element_size
    return -(-x // per16) * per16


def scaled_fp8_quant(
    input: torch.Tensor,
    scale: Optional[torch.Tensor] = None,
    num_token_padding: Optional[int] = None,
    scale_ub: Optional[torch.Tensor] = None,
    use_per_token_if_dynamic: bool = False,
    output: Optional[torch.Tensor] = None,
    fp8_dtype: torch.dtype = torch.float8_e5m2,
    group_shape: tuple[int, int] | None = None,
) -> tuple[torch.Tensor, torch.Tensor]:
    """FP8 quantisation of a [T, H] tensor: static when `scale` is given (per tensor, or per `group_shape`),
    dynamic per token or per tensor otherwise.  Interface of tests/ops/fp8_quant_op.py:135-194."""
    if input.ndim != 2:
        raise AssertionError("scaled_fp8_quant takes a 2-D input")
    rows = input.shape[0] if not num_token_padding else max(int(num_token_padding), input.shape[0])
    if output is not None:
        assert num_token_padding is None, "padding not supported if output passed in"
        assert output.dtype == fp8_dtype
    else:
        output = torch.empty((rows, input.shape[1]), device=input.device, dtype=fp8_dtype)
    if scale is not None:
        static_scaled_fp8_quant(output, input, scale, group_shape)
        return output, scale
    if use_per_token_if_dynamic:
        scale = torch.zeros((rows, 1), device=input.device, dtype=torch.float32)
        dynamic_per_token_scaled_fp8_quant(output, input.contiguous(), scale, scale_ub)
    else:
        scale = torch.empty(1, device=input.device, dtype=torch.float32)
        dynamic_scaled_fp8_quant(output, input, scale)
    return output, scale


def per_token_group_quant_fp8(
    x: torch.Tensor,
    group_size: int,
    eps: float = 1e-10,
    dtype: torch.dtype = torch.float8_e4m3fn,
    out_q: torch.Tensor | None = None,
    column_major_scales: bool = False,
    tma_aligned_scales: bool = False,
    use_ue8m0: bool | None = None,
) -> tuple[torch.Tensor, torch.Tensor]:
    """1 x group_size FP8 quantisation along the last dimension; scales row-major, column-major or column-major with a
    TMA-aligned pitch; UE8M0 scales come back as float8_e8m0fnu.  Interface of tests/ops/fp8_quant_op.py:206-276."""
    if x.shape[-1] % group_size != 0 or x.stride(-1) != 1:
        raise AssertionError("the last dimension must be contiguous and a multiple of group_size")
    q = torch.empty_like(x, dtype=dtype) if out_q is None else out_q
    layout = "row" if not column_major_scales else ("col_tma" if tma_aligned_scales else "col")
    s = _alloc_group_scales(tuple(x.shape[:-1]), x.shape[-1] // group_size, x.device, layout)
    lim = torch.finfo(dtype)
    torch.ops._C.per_token_group_fp8_quant(x, q, s, group_size, eps, lim.min, lim.max, bool(use_ue8m0),
                                           column_major_scales, tma_aligned_scales)
    return q, (s.to(torch.float8_e8m0fnu) if use_ue8m0 else s)


def per_token_group_quant_mxfp4(
    x: torch.Tensor,
    group_size: int = 32,
    eps: float = 1e-10,
    column_major_scales: bool = False,
) -> tuple[torch.Tensor, torch.Tensor]:
    """tests/ops/mxfp4_quant_op.py:8-63: returns (float4_e2m1fn_x2 [M,N/2], float8_e8m0fnu [M,N/32])."""
    assert x.ndim == 2 and x.shape[-1] % group_size == 0 and x.stride(-1) == 1
    M, N = x.shape
    out_q = torch.empty(M, N // 2, device=x.device, dtype=torch.uint8)
    out_s = _alloc_group_scales((M,), N // group_size, x.device, "col" if column_major_scales else "row")
    torch.ops._C.per_token_group_quant_mxfp4(x.contiguous(), out_q, out_s, group_size, eps)
    out_q = out_q.view(torch.float4_e2m1fn_x2)
    out_s = out_s.to(dtype=torch.float8_e8m0fnu, memory_format=torch.preserve_format)
    return out_q, out_s


def silu_and_mul_per_block_quant(
    input: torch.Tensor,
    group_size: int,
    quant_dtype: torch.dtype,
    scale_ub: Optional[torch.Tensor] = None,
    is_scale_transposed: bool = False,
    scale_ue8m0: bool = False,
) -> tuple[torch.Tensor, torch.Tensor]:
    """tests/register_ops.py:34-74."""
    assert input.ndim == 2
    num_tokens = input.shape[0]
    hidden_size = input.shape[-1] // 2
    num_groups = hidden_size // group_size
    out = torch.empty((num_tokens, hidden_size), device=input.device, dtype=quant_dtype)
    scales = _alloc_group_scales((num_tokens,), num_groups, input.device, "col" if is_scale_transposed else "row")
    torch.ops._C.silu_and_mul_per_block_quant(out, input, scales, group_size, scale_ub,
                                              is_scale_transposed, scale_ue8m0)
    return out, scales


def silu_and_mul_mxfp4_quant(input: torch.Tensor, group_size: int = 32, eps: float = 1e-10):
    """tests/test_fused_silu_mul_mxfp4_quant.py:44-56: returns (uint8 [T,H/2], fp32 scales [T,H/32])."""
    T = input.shape[0]
    H = input.shape[-1] // 2
    out = torch.empty(T, H // 2, dtype=torch.uint8, device=input.device)
    scales = torch.empty(T, H // group_size, dtype=torch.float32, device=input.device)
    torch.ops._C.silu_and_mul_mxfp4_quant(out, input.contiguous(), scales, group_size, eps)
    return out, scales


def rms_norm_dynamic_per_token_quant(x, weight, eps, quant_dtype, scale_ub=None, residual=None):
    """tests/test_fused_norm_quant.py:128-144 (residual is updated in place)."""
    x = x.contiguous()
    out = torch.empty_like(x, dtype=quant_dtype)
    T = x.numel() // x.shape[-1]
    scales = torch.empty(T, dtype=torch.float32, device=x.device)
    torch.ops._C.rms_norm_dynamic_per_token_quant(out, x, weight, scales, eps, scale_ub, residual)
    return out, scales


def rms_norm_per_block_quant(x, weight, eps, group_size, quant_dtype, residual=None,
                             is_scale_transposed=False, scale_ue8m0=False):
    """tests/test_fused_norm_quant.py:147-176."""
    out = torch.empty(x.shape, dtype=quant_dtype, device=x.device)
    T = x.numel() // x.shape[-1]
    G = x.shape[-1] // group_size
    scales = _alloc_group_scales((T,), G, x.device, "col" if is_scale_transposed else "row")
    torch.ops._C.rms_norm_per_block_quant(out, x, weight, scales, eps, None, residual, group_size,
                                          is_scale_transposed, scale_ue8m0)
    return out, scales


def rms_norm_mxfp4_quant(x, weight, eps, residual=None, group_size: int = 32):
    """tests/test_fused_norm_quant.py:645-663."""
    x = x.contiguous()
    T = x.numel() // x.shape[-1]
    H = x.shape[-1]
    out = torch.empty(T, H // 2, device=x.device, dtype=torch.uint8)
    scales = torch.empty(T, H // group_size, device=x.device, dtype=torch.float32)
    torch.ops._C.rms_norm_mxfp4_quant(out, x, weight, scales, eps, residual, group_size)
    return out, scales


def rms_norm_static_fp8_quant(x, weight, scale, eps, residual=None, quant_dtype=torch.float8_e4m3fn):
    """tests/test_fused_norm_quant.py:424-445."""
    out = torch.empty(x.shape, dtype=quant_dtype, device=x.device)
    if residual is None:
        torch.ops._C.rms_norm_static_fp8_quant(out, x, weight, scale, eps)
    else:
        torch.ops._C.fused_add_rms_norm_static_fp8_quant(out, x, residual, weight, scale, eps)
    return out


# ------------------------------------------------- extensions: MX scale factors in the tensor core's layout
# (include/vxk_b200.h "(f) next #2"; no reference counterpart -- the reference API above stays the drop-in)
def mx_sf_buffer(rows: int, K: int, device, fp4: bool = True, is_weight: bool = False) -> torch.Tensor:
    """Packed scale-factor buffer for a [rows, K] operand, to be filled by one of the *_sf quantisation ops.
    Zero-filled only when K has a partial K slice (the kernels write real groups only)."""
    n = torch.ops._vxk.mx_sf_packed_bytes(rows, K, fp4, is_weight)
    slice_k = 256 if fp4 else 128
    return (torch.zeros if K % slice_k else torch.empty)(n, dtype=torch.uint8, device=device)


def mx_sf_pack(scale_e8m0: torch.Tensor, fp4: bool = True, is_weight: bool = True) -> torch.Tensor:
    """[rows, K/32] float8_e8m0fnu -> packed uint8 buffer (once per weight, after loading)."""
    return torch.ops._vxk.mx_sf_pack(scale_e8m0, fp4, is_weight)


def per_token_group_quant_mxfp4_sf(x: torch.Tensor, group_size: int = 32, eps: float = 1e-10):
    """per_token_group_quant_mxfp4 that also returns the packed scale factors: (q, e8m0 scales, packed)."""
    M, N = x.shape
    out_q = torch.empty(M, N // 2, device=x.device, dtype=torch.uint8)
    out_s = torch.empty(M, N // group_size, device=x.device, dtype=torch.float32)
    sf = mx_sf_buffer(M, N, x.device)
    torch.ops._vxk.per_token_group_quant_mxfp4_sf(x.contiguous(), out_q, out_s, sf, group_size, eps)
    return out_q.view(torch.float4_e2m1fn_x2), out_s.to(torch.float8_e8m0fnu), sf


def rms_norm_mxfp4_quant_sf(x, weight, eps, residual=None, group_size: int = 32):
    x = x.contiguous()
    H = x.shape[-1]
    T = x.numel() // H
    out = torch.empty(T, H // 2, device=x.device, dtype=torch.uint8)
    scales = torch.empty(T, H // group_size, device=x.device, dtype=torch.float32)
    sf = mx_sf_buffer(T, H, x.device)
    torch.ops._vxk.rms_norm_mxfp4_quant_sf(out, x, weight, scales, sf, eps, residual, group_size)
    return out, scales, sf


def silu_and_mul_mxfp4_quant_sf(input: torch.Tensor, group_size: int = 32, eps: float = 1e-10):
    T = input.shape[0]
    H = input.shape[-1] // 2
    out = torch.empty(T, H // 2, dtype=torch.uint8, device=input.device)
    scales = torch.empty(T, H // group_size, dtype=torch.float32, device=input.device)
    sf = mx_sf_buffer(T, H, input.device)
    torch.ops._vxk.silu_and_mul_mxfp4_quant_sf(out, input.contiguous(), scales, sf, group_size, eps)
    return out, scales, sf


def fp4_gemm_sf(input, weight, scale_act, scale_wei, out_dtype, bias=None, act_sf_packed=None, wei_sf_packed=None):
    return torch.ops._vxk.fp4_gemm_sf(input, weight, scale_act, scale_wei, out_dtype, bias, act_sf_packed, wei_sf_packed)


def fp8_gemm_silu_mul_quant(input, weight, scale_act, scale_wei, bias=None, mid_dtype=torch.bfloat16,
                            is_scale_transposed=False):
    """Gate_up projection + silu_and_mul + 1x128 block FP8 quantisation in one kernel (no reference
    counterpart).  `weight` is the [K, 2 * inter] transposed view of W[2 * inter, K] (gate rows, then up rows).
    Returns (codes [M, inter] float8_e4m3fn, scales [M, inter / 128] float32), bit-identical to
    fp8_gemm(..., mid_dtype) followed by silu_and_mul_per_block_quant(group 128)."""
    M, inter = input.shape[0], weight.shape[1] // 2
    out_q = torch.empty(M, inter, dtype=torch.float8_e4m3fn, device=input.device)
    if is_scale_transposed:
        out_s = torch.empty(inter // 128, M, dtype=torch.float32, device=input.device).t()
    else:
        out_s = torch.empty(M, inter // 128, dtype=torch.float32, device=input.device)
    torch.ops._vxk.fp8_gemm_silu_mul_quant(input, weight, scale_act, scale_wei, bias, mid_dtype, out_q, out_s)
    return out_q, out_s
