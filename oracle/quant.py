"""CPU restatement of the reference's torch reference implementations of the
activation-quantisation ops.
TEST INFRASTRUCTURE ONLY (see oracle/__init__).

Every function is functional (returns new tensors) and runs on the CPU; the
citation on each one is the reference test code it follows.  Where the
reference's test reference evaluates in the input dtype (bf16 arithmetic) and
therefore only pins its kernel to a loose tolerance, two functions exist:
``*_literal`` follows the test code operation by operation, the plain name is
the fp32 evaluation of the same formula that the CUDA kernels are held to
bit-exactly.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from .formats import (E4M3, E5M2, FP8_MAX, mx_quantize, e8m0_to_f32,
                      min_scaling_factor)

F32 = torch.float32


def _f32(x) -> torch.Tensor:
    return torch.as_tensor(x, dtype=F32)


def _sat_cast(v: torch.Tensor, qdtype: torch.dtype) -> torch.Tensor:
    m = FP8_MAX[qdtype]
    return v.clamp(-m, m).to(qdtype)


# --------------------------------------------------------------------------
# a1  dynamic per-token FP8           tests/test_fp8_quant.py:48-83
# --------------------------------------------------------------------------
def dynamic_per_token_scaled_fp8_quant(x: torch.Tensor, qdtype=E4M3,
                                       scale_ub: Optional[torch.Tensor] = None):
    """amax per row -> min(amax, ub) -> s = max(amax / fp8max, 1/(fp8max*512))
    -> q = fp8(clamp(float(x) / s)).  TRUE division, as the reference test.
    Returns (q [T,H], scales [T,1] fp32)."""
    qmax = _f32(FP8_MAX[qdtype])
    amax = _f32(x.abs().max(dim=-1)[0])
    if scale_ub is not None:
        amax = amax.clamp(max=scale_ub)
    s = (amax / qmax)[:, None].clamp(min=min_scaling_factor(qdtype))
    q = _sat_cast(_f32(x) / s, qdtype)
    return q, s


# --------------------------------------------------------------------------
# a2 / a3  static and dynamic per-tensor FP8    tests/test_fp8_quant.py:28-45
# --------------------------------------------------------------------------
def dynamic_scaled_fp8_quant(x: torch.Tensor, qdtype=E5M2):
    """s = amax / fp8max (no floor); q = fp8(clamp(float(x) * (1/s))).
    Returns (q, s[1])."""
    s = _f32(x.abs().max()) / _f32(FP8_MAX[qdtype])
    q = _sat_cast(_f32(x) * (_f32(1.0) / s), qdtype)
    return q, s.view(1)


def static_scaled_fp8_quant(x: torch.Tensor, scale: torch.Tensor, qdtype=E4M3,
                            group_shape: Optional[tuple[int, int]] = None):
    """q = fp8(clamp(float(x) * (1 / s[gi, gj]))) with the scale broadcast rule
    of csrc/quantization/fp8/fp8_quant.cpp:481-623: 0-D / 1 element = per
    tensor; 1-D needs group_shape ((1,-1) per token, (-1,1) per channel);
    2-D [T/gm, H/gn] per block.  The multiply-by-reciprocal form is the one
    the exact reference uses (test_fp8_quant.py:41-43, test_fused_norm_quant.py
    :448-456)."""
    T, H = x.shape
    xf = _f32(x)
    sc = _f32(scale)
    if sc.numel() == 1:
        inv = (_f32(1.0) / sc.reshape(())).expand(T, H)
    elif sc.dim() == 1:
        assert group_shape is not None, "1-D scale needs group_shape"
        gm = T if group_shape[0] == -1 else group_shape[0]
        gn = H if group_shape[1] == -1 else group_shape[1]
        assert sc.numel() == (T // gm) * (H // gn)
        s2 = sc.reshape(T // gm, H // gn)
        inv = (1.0 / s2).repeat_interleave(gm, 0).repeat_interleave(gn, 1)
    else:
        gm, gn = T // sc.shape[0], H // sc.shape[1]
        inv = (1.0 / sc).repeat_interleave(gm, 0).repeat_interleave(gn, 1)
    return _sat_cast(xf * inv, qdtype)


# --------------------------------------------------------------------------
# a4  per-token-group FP8 (1 x G)     tests/test_fp8_quant.py:86-142
# --------------------------------------------------------------------------
def per_token_group_fp8_quant(x: torch.Tensor, group_size: int, eps: float = 1e-10,
                              qdtype=E4M3, scale_ue8m0: bool = False):
    """fp32 evaluation of ref_per_block_quant(x, 1, G):
        s = max(amax / fp8max, eps) ; q = fp8(clamp(float(x) / s))
    and, with scale_ue8m0, of to_mxfp(x, block_size=G) (mx_utils.py:277-353)
    whose power-of-two scale is returned as fp32 like the op does
    (tests/ops/fp8_quant_op.py:267-274 casts it to e8m0 afterwards).
    Returns (q [T,H], s [T,H/G] fp32)."""
    T, H = x.shape
    G = group_size
    if scale_ue8m0:
        xin = x if x.dtype in (F32, torch.bfloat16) else x.to(F32)
        e, q = mx_quantize(xin.contiguous(), "mxfp8", block=G, exact_exponent=True)
        assert qdtype == E4M3
        return q, e8m0_to_f32(e)
    xg = _f32(x).reshape(T, H // G, G)
    amax = xg.abs().amax(dim=-1)
    s = (amax / _f32(FP8_MAX[qdtype])).clamp(min=eps)
    q = _sat_cast(xg / s.unsqueeze(-1), qdtype).reshape(T, H)
    return q, s


def per_token_group_fp8_quant_literal(x: torch.Tensor, group_size: int,
                                      eps: float = 1e-10):
    """tests/test_fp8_quant.py:86-142 operation by operation, i.e. with amax,
    the scale and the division evaluated in x.dtype (bf16/f16 round-off and
    no saturation).  Only good to the reference's own tolerance (atol=rtol
    0.15 on values, 0.01 on scales, :338-345)."""
    T, H = x.shape
    G = group_size
    xg = x.reshape(T, H // G, G)
    amax = xg.abs().amax(dim=-1)                 # x.dtype
    s = torch.clamp(amax / 448.0, min=eps)       # x.dtype
    q = (xg / s.unsqueeze(-1)).to(E4M3).reshape(T, H)
    return q, s.to(F32)


# --------------------------------------------------------------------------
# a5  per-token-group MXFP4           tests/test_mxfp4_quant.py:59-110
# --------------------------------------------------------------------------
def per_token_group_quant_mxfp4(x: torch.Tensor, group_size: int = 32,
                                exact_exponent: bool = True):
    """to_mxfp(x, block_size=32, format="mxfp4") with fp16 upcast to fp32
    first (test_mxfp4_quant.py:77).  Returns (packed uint8 [T,H/2],
    scales fp32 [T,H/32] = 2^(e-127), scale bytes uint8)."""
    assert group_size == 32
    xin = x if x.dtype in (F32, torch.bfloat16) else x.to(F32)
    e, q = mx_quantize(xin.contiguous(), "mxfp4", block=32,
                       exact_exponent=exact_exponent)
    return q.view(torch.uint8), e8m0_to_f32(e), e.view(torch.uint8)


# --------------------------------------------------------------------------
# rms-norm prologue                   tests/test_fused_norm_quant.py:37-53
# --------------------------------------------------------------------------
# The one step of the rms-norm refs whose fp32 result depends on the HOST: torch's fp32
# x.pow(2).mean() is a vectorised cascade sum whose grouping follows the CPU's vector width, so its
# last bits differ between machines (and from any GPU reduction tree).  EXACT_VARIANCE = True
# evaluates the same formula with the mean carried out in fp64 and rounded once to fp32 -- the
# correctly rounded value every summation order approximates; the CUDA kernels are held to THAT
# bit for bit, and tests/test_quant_parity_gpu.py measures how far torch's own fp32 order strays
# from it (a few ulp of the variance at most).
EXACT_VARIANCE = False


def rms_norm_f32(x: torch.Tensor, weight: torch.Tensor, eps: float,
                 residual: Optional[torch.Tensor] = None,
                 exact_variance: Optional[bool] = None):
    """z = x_dtype(float(x) + float(res)) (written back as the new residual);
    var = mean(z^2); normed = z * rsqrt(var + eps) * float(w), all fp32 and no
    intermediate rounding (tests/test_fused_norm_quant.py:37-53).
    exact_variance (default: module flag EXACT_VARIANCE): mean in fp64, rounded once.
    Returns (normed fp32, new residual or None)."""
    if exact_variance is None:
        exact_variance = EXACT_VARIANCE
    xf = x.float()
    if residual is not None:
        z = (xf + residual.float()).to(x.dtype)
        residual = z.clone()
        xf = z.float()
    if exact_variance:
        var = xf.double().pow(2).mean(dim=-1, keepdim=True).float()
    else:
        var = xf.pow(2).mean(dim=-1, keepdim=True)
    inv_rms = torch.rsqrt(var + eps)
    return xf * inv_rms * weight.float(), residual


def _per_token_quant_f32(normed: torch.Tensor, qdtype, scale_ub=None):
    """tests/test_fused_norm_quant.py:56-80."""
    n2 = normed.reshape(-1, normed.shape[-1]).float()
    amax = n2.abs().amax(dim=1, keepdim=True)
    if qdtype == torch.int8:
        s = torch.where(amax > 0, amax / 127.0, torch.ones_like(amax))
        q = (n2 / s).round().clamp(-128, 127).to(torch.int8)
        return q, s.squeeze(1)
    qmax = FP8_MAX[qdtype]
    if scale_ub is not None:
        amax = torch.min(amax, scale_ub.float())
    s = torch.clamp(amax / qmax, min=1.0 / (qmax * 512.0))
    q = (n2 * (1.0 / s)).clamp(-qmax, qmax).to(qdtype)
    return q, s.squeeze(1)


def _per_group_quant_f32(normed: torch.Tensor, G: int, qdtype,
                         scale_ue8m0: bool = False):
    """tests/test_fused_norm_quant.py:83-125 (vectorised over groups)."""
    T, H = normed.reshape(-1, normed.shape[-1]).shape
    ng = normed.reshape(T, H // G, G).float()
    amax = ng.abs().amax(dim=-1, keepdim=True)
    if qdtype == torch.int8:
        s = torch.where(amax > 0, amax / 127.0, torch.ones_like(amax))
        q = (ng / s).round().clamp(-128, 127).to(torch.int8)
    else:
        qmax = FP8_MAX[qdtype]
        s = torch.clamp(amax / qmax, min=1.0 / (qmax * 512.0))
        if scale_ue8m0:
            s = exp2_ceil_log2(torch.clamp(s.abs(), min=1e-10))
        q = (ng * (1.0 / s)).clamp(-qmax, qmax).to(qdtype)
    return q.reshape(T, H), s.squeeze(-1)


def exp2_ceil_log2(s: torch.Tensor) -> torch.Tensor:
    """exp2(ceil(log2(s))) for positive normal fp32 s, evaluated on the bit
    pattern (exact).  The reference evaluates the same expression with fp32
    log2 (test_fused_norm_quant.py:113-116); the two agree except within a few
    ulp above a power of two -- see DESIGN.md 'ceil(log2)'."""
    bits = s.contiguous().view(torch.int32)
    up = ((bits & 0x7FFFFF) != 0).to(torch.int32)
    return (((bits >> 23) + up) << 23).view(F32)


# --------------------------------------------------------------------------
# a6  rms_norm + dynamic per-token     tests/test_fused_norm_quant.py:179-248
# --------------------------------------------------------------------------
def rms_norm_dynamic_per_token_quant(x, weight, eps, qdtype=E4M3, scale_ub=None,
                                     residual=None, exact_variance=None):
    """Returns (q [T,H], scales [T], new residual or None)."""
    normed, res = rms_norm_f32(x, weight, eps, residual, exact_variance)
    q, s = _per_token_quant_f32(normed, qdtype, scale_ub)
    return q, s, res


# --------------------------------------------------------------------------
# a7  rms_norm + per-block             tests/test_fused_norm_quant.py:251-317,565-636
# --------------------------------------------------------------------------
def rms_norm_per_block_quant(x, weight, eps, group_size, qdtype=E4M3,
                             residual=None, scale_ue8m0=False, exact_variance=None):
    """Returns (q [T,H], scales [T,H/G], new residual or None)."""
    normed, res = rms_norm_f32(x, weight, eps, residual, exact_variance)
    q, s = _per_group_quant_f32(normed, group_size, qdtype, scale_ue8m0)
    return q, s, res


# --------------------------------------------------------------------------
# a8  rms_norm + MXFP4                 tests/test_fused_norm_quant.py:666-737
# --------------------------------------------------------------------------
def rms_norm_mxfp4_quant(x, weight, eps, residual=None, exact_exponent=True,
                         exact_variance=None):
    """to_mxfp(normed.float(), 32, "mxfp4").  Returns (packed u8 [T,H/2],
    scales fp32 [T,H/32] = 2^(e-127), new residual or None)."""
    normed, res = rms_norm_f32(x, weight, eps, residual, exact_variance)
    e, q = mx_quantize(normed.float().contiguous(), "mxfp4",
                       exact_exponent=exact_exponent)
    return q.view(torch.uint8), e8m0_to_f32(e), res


# --------------------------------------------------------------------------
# a9  rms_norm (+add) + static FP8     tests/test_fused_norm_quant.py:448-456
# --------------------------------------------------------------------------
def rms_norm_static_fp8_quant(x, weight, scale, eps, residual=None, exact_variance=None):
    """fp32 normed (no rounding to the activation dtype) * (1/scale) -> e4m3.
    Returns (q, new residual or None)."""
    normed, res = rms_norm_f32(x, weight, eps, residual, exact_variance)
    inv = 1.0 / scale.float()
    q = (normed.float() * inv).clamp(-448.0, 448.0).to(E4M3)
    return q, res


# --------------------------------------------------------------------------
# a10 silu_and_mul + static FP8        tests/test_fused_quant_activation.py:28-40
# --------------------------------------------------------------------------
def silu_and_mul_quant(x: torch.Tensor, scale: torch.Tensor, qdtype=E4M3):
    """y = silu(gate) * up evaluated in x.dtype (two roundings), then
    fp8(clamp(float(y) * (1/scale)))."""
    d = x.shape[-1] // 2
    y = F.silu(x[..., :d]) * x[..., d:]
    inv = _f32(1.0 / float(scale.reshape(-1)[0].item()))
    m = FP8_MAX[qdtype]
    return (y.float() * inv).clamp(-m, m).to(qdtype)


# --------------------------------------------------------------------------
# a11 silu_and_mul + per-block         tests/test_fused_silu_mul_block_quant.py:49-79
# --------------------------------------------------------------------------
def silu_and_mul_per_block_quant(x: torch.Tensor, group_size: int, qdtype=E4M3,
                                 scale_ub: Optional[torch.Tensor] = None,
                                 scale_ue8m0: bool = False):
    """y = silu(gate) * up in x.dtype; per 1xG group:
       FP8 : s = max(gmax, float32.tiny) / 448 ; q = fp8(clamp(y / s))
       INT8: s = max(gmax, 1e-10) / 127       ; q = clamp(round(y / s))
    scale_ub / scale_ue8m0 are op features the reference test does not cover;
    they follow the kernel (fused_silu_mul_block_quant.cpp:93-110):
    s = min(s, ub) then s = 2^ceil(log2(max(s, 1e-10))).
    Returns (q [T,H], s [T,H/G])."""
    H = x.shape[-1] // 2
    G = group_size
    y = (F.silu(x[..., :H]) * x[..., H:]).contiguous()
    yg = y.float().reshape(*y.shape[:-1], H // G, G)
    gmax = yg.abs().amax(dim=-1)
    if qdtype == torch.int8:
        s = gmax.clamp(min=1e-10) / 127.0
        q = (yg / s.unsqueeze(-1)).round().clamp(-128, 127)
    else:
        m = FP8_MAX[qdtype]
        s = gmax.clamp(min=torch.finfo(F32).tiny) / m
        if scale_ub is not None:
            s = torch.min(s, scale_ub.float())
        if scale_ue8m0:
            s = exp2_ceil_log2(torch.clamp(s, min=1e-10))
        q = (yg / s.unsqueeze(-1)).clamp(-m, m)
    return q.reshape(y.shape).to(qdtype), s.to(F32)


# --------------------------------------------------------------------------
# a12 silu_and_mul + MXFP4             tests/test_fused_silu_mul_mxfp4_quant.py:22-41
# --------------------------------------------------------------------------
def silu_and_mul_mxfp4_quant(x: torch.Tensor, exact_exponent: bool = True):
    """y = gate_f32 * sigmoid(gate_f32) * up_f32 (fp32, no rounding) ->
    to_mxfp(y, 32, "mxfp4").  Returns (packed u8 [T,H/2], scales fp32 [T,H/32])."""
    H = x.shape[-1] // 2
    g = x[..., :H].float()
    u = x[..., H:].float()
    y = (g * torch.sigmoid(g) * u).contiguous()
    e, q = mx_quantize(y, "mxfp4", exact_exponent=exact_exponent)
    return q.view(torch.uint8), e8m0_to_f32(e)


# --------------------------------------------------------------------------
# "codes given the scale": the second half of every quantiser above, on its own.
# Used by the parity tests where a kernel scale legitimately differs from the
# oracle's by one step (a 1-ulp difference upstream): the codes must then be
# exactly what the reference formula gives FOR THAT SCALE.
# --------------------------------------------------------------------------
def fp8_codes_given_scale(v: torch.Tensor, s: torch.Tensor, qdtype, group: int = 0,
                          divide: bool = False) -> torch.Tensor:
    """v [T,H] fp32 values, s [T] / [T,1] (group == 0) or [T,H/group] fp32 scales.
    divide=False: sat(v * (1/s)) (a6, a7, a9, a10 -- the refs multiply by the reciprocal);
    divide=True:  sat(v / s) (a1, a4, a11 -- the refs divide).  int8: round-half-even of v / s."""
    T, H = v.shape
    v = v.float()
    se = s.reshape(T, -1).float()
    se = se.expand(T, H) if se.shape[1] == 1 else se.repeat_interleave(group, dim=1)
    if qdtype == torch.int8:
        return (v / se).round().clamp(-128, 127).to(torch.int8)
    m = FP8_MAX[qdtype]
    q = (v / se) if divide else (v * (1.0 / se))
    return q.clamp(-m, m).to(qdtype)


def mxfp4_codes_given_exponent(v: torch.Tensor, e: torch.Tensor) -> torch.Tensor:
    """to_mxfp's data path (mx_utils.py:327-353) for given E8M0 bytes e [T,H/32] (uint8):
    clamp(v * 2^(127-e), +-6) -> bf16 -> E2M1 RNE -> packed bytes [T,H/2]."""
    from .formats import e2m1_encode, pack_nibbles, E8M0_BIAS
    T, H = v.shape
    eb = e.view(torch.uint8).reshape(T, H // 32, 1)
    inv = torch.where(eb == 0, torch.tensor(1.0), torch.exp2(E8M0_BIAS - eb.to(F32)))
    q = torch.clamp(v.float().reshape(T, H // 32, 32) * inv, min=-6.0, max=6.0)
    codes = e2m1_encode(q.to(torch.bfloat16).to(F32))
    return pack_nibbles(codes.reshape(T, H))


def silu_mul_T(x: torch.Tensor) -> torch.Tensor:
    """y = silu(gate) * up evaluated in x.dtype (a10 / a11 prologue)."""
    H = x.shape[-1] // 2
    return (F.silu(x[..., :H]) * x[..., H:]).contiguous()


def silu_mul_f32(x: torch.Tensor) -> torch.Tensor:
    """y = gate_f32 * sigmoid(gate_f32) * up_f32 (a12 prologue)."""
    H = x.shape[-1] // 2
    g = x[..., :H].float()
    return (g * torch.sigmoid(g) * x[..., H:].float()).contiguous()
