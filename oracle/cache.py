"""CPU restatement of the reference's torch references for the indexer-K cache ops and of the
convert_fp8 kernel formula.  TEST INFRASTRUCTURE ONLY (see oracle/__init__).

  indexer_k_quant_and_cache        tests/test_indexer_k_quant_and_cache.py:34-131
  cp_gather_indexer_k_quant_cache  tests/test_cp_gather_indexer_k_quant_cache.py:34-86
  convert_fp8                      csrc/cache.cpp:1393-1440 (the reference has no torch ref for it: it is
                                   itself the helper its cache tests use, tests/test_cache.py:396,452-478)

Cache block layout (bytes): [block_size * head_dim e4m3 codes][block_size * head_dim / G fp32 scales].
"""
from __future__ import annotations

import torch

from .formats import E4M3
from .quant import exp2_ceil_log2

F32 = torch.float32
EPS = 1e-4   # tests/test_indexer_k_quant_and_cache.py:8


def indexer_group_quant(x: torch.Tensor, group_size: int = 128, use_ue8m0: bool = False,
                        exact_exponent: bool = True):
    """_pytorch_group_quant (:34-92): fp32 amax per 1 x G group, floored at 1e-4, s = amax / 448
    (ue8m0: 2^ceil(log2 s)), q = e4m3(clamp(x / s)).  exact_exponent: the power of two from the bit
    pattern instead of fp32 log2 (they agree except a few ulp above 2^k, DESIGN.md 'ceil(log2)').
    Returns (q [T, H] e4m3, s [T, H/G] fp32)."""
    T, H = x.shape
    xg = x.reshape(T, H // group_size, group_size).float()
    amax = torch.maximum(xg.abs().amax(dim=-1), torch.tensor(EPS, dtype=F32))
    s = amax / 448.0
    if use_ue8m0:
        s = exp2_ceil_log2(s) if exact_exponent else torch.pow(2.0, torch.ceil(torch.log2(s)))
    q = torch.clamp(xg / s.unsqueeze(-1), -448.0, 448.0).to(E4M3).reshape(T, H)
    return q, s.float()


def indexer_k_quant_and_cache(k: torch.Tensor, kv_cache: torch.Tensor, slot_mapping: torch.Tensor,
                              quant_block_size: int, scale_fmt: str, exact_exponent: bool = True) -> None:
    """ref_indexer_k_quant_and_cache (:95-131), vectorised: writes codes and scales of every token with a
    non-negative slot into kv_cache [num_blocks, block_size, cache_stride] (uint8, contiguous) in place."""
    T, head_dim = k.shape
    groups = head_dim // quant_block_size
    block_size, cache_stride = kv_cache.shape[1], kv_cache.shape[2]
    q, s = indexer_group_quant(k, quant_block_size, scale_fmt == "ue8m0", exact_exponent)
    flat = kv_cache.view(-1)
    slots = slot_mapping.flatten()[:T].long()
    keep = slots >= 0
    blk = torch.div(slots, block_size, rounding_mode="floor")
    off = slots - blk * block_size
    base = blk * (block_size * cache_stride)
    # codes
    dst = (base + off * head_dim)[keep].unsqueeze(1) + torch.arange(head_dim)
    flat[dst.reshape(-1)] = q.view(torch.uint8)[keep].reshape(-1)
    # scales: float index = (base + block_size*head_dim + (off*head_dim + g*G)*4 // G) // 4
    g = torch.arange(groups)
    sidx = (base.unsqueeze(1) + block_size * head_dim +
            torch.div((off.unsqueeze(1) * head_dim + g * quant_block_size) * 4, quant_block_size,
                      rounding_mode="floor"))
    sidx = torch.div(sidx, 4, rounding_mode="floor")[keep]
    flat.view(F32)[sidx.reshape(-1)] = s[keep].reshape(-1)


def cp_gather_indexer_k_quant_cache(kv_cache: torch.Tensor, dst_k: torch.Tensor, dst_scale: torch.Tensor,
                                    block_table: torch.Tensor, cu_seq_lens: torch.Tensor) -> None:
    """ref_cp_gather_indexer_k_quant_cache (:34-86): for every token of every sequence copy its head_dim
    code bytes and its fp32 scales out of the paged cache (dst_k [tokens, head_dim] uint8, dst_scale
    [tokens, 4 * groups] uint8), in place."""
    head_dim = dst_k.shape[1]
    groups = dst_scale.shape[1] // 4
    G = head_dim // groups
    block_size, cache_stride = kv_cache.shape[1], kv_cache.shape[2]
    block_stride = block_size * cache_stride
    flat = kv_cache.reshape(-1)
    f32 = flat.view(F32)
    ds = dst_scale.view(-1).view(F32)
    for b in range(block_table.shape[0]):
        s0, s1 = int(cu_seq_lens[b]), int(cu_seq_lens[b + 1])
        for pos in range(s1 - s0):
            tok = s0 + pos
            blk = int(block_table[b, pos // block_size])
            off = pos % block_size
            src = blk * block_stride + off * head_dim
            dst_k[tok] = flat[src:src + head_dim]
            for g in range(groups):
                sb = blk * block_stride + block_size * head_dim + (off * head_dim + g * G) * 4 // G
                ds[tok * groups + g] = f32[sb // 4]


def convert_fp8(src: torch.Tensor, scale: float, kv_cache_dtype: str, out_dtype: torch.dtype) -> torch.Tensor:
    """csrc/cache.cpp:1411-1435.  "fp8" / "fp8_e4m3": float source -> uint8 holding e4m3(clamp(x / scale));
    uint8 source -> out_dtype(float(e4m3) * scale).  "auto": plain cast of float(src) * scale."""
    sc = torch.tensor(scale, dtype=F32)
    if kv_cache_dtype == "auto":
        v = src.float() * sc
        if out_dtype == torch.uint8:
            return v.trunc().clamp(0, 255).to(torch.uint8)
        return v.to(out_dtype)
    assert kv_cache_dtype in ("fp8", "fp8_e4m3")
    if src.dtype == torch.uint8:
        return (src.view(E4M3).float() * sc).to(out_dtype)
    return (src.float() / sc).clamp(-448.0, 448.0).to(E4M3).view(torch.uint8)
