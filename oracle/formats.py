"""Number formats of the hot path: FP8 (E4M3 / E5M2), FP4 E2M1, UE8M0 scales,
and OCP-MX block quantisation.  TEST INFRASTRUCTURE ONLY (see oracle/__init__).

The MX helpers restate ``/root/reference/tests/ops/mx_utils.py`` (``to_mxfp``
:277-353, ``_f32_to_floatx_unpacked`` :152-266, ``_floatx_unpacked_to_f32``
:21-110, ``pack_uint4``/``unpack_uint4`` :122-149, ``from_blocked_format``
:113-119) with a different formulation (comparison ladder / lookup table
instead of the bit-twiddling there); ``tests/test_oracle_golden.py`` proves
the two agree on every bf16 bit pattern.
"""
from __future__ import annotations

import torch

E4M3 = torch.float8_e4m3fn
E5M2 = torch.float8_e5m2
E8M0 = torch.float8_e8m0fnu
FP8_MAX = {E4M3: 448.0, E5M2: 57344.0}
E2M1_MAX = 6.0
E8M0_BIAS = 127
MX_BLOCK = 32

# positive E2M1 code points, index == 3-bit magnitude code
E2M1_VALUES = (0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0)


def fp8_max(dtype: torch.dtype) -> float:
    return FP8_MAX[dtype]


def min_scaling_factor(dtype: torch.dtype) -> torch.Tensor:
    """fp32 ``1 / (fp8_max * 512)`` -- csrc/quantization/fp8/quant_utils.h:35-37,
    tests/test_fp8_quant.py:76 (computed in fp32 there as well)."""
    one = torch.tensor(1.0, dtype=torch.float32)
    return one / (torch.tensor(FP8_MAX[dtype], dtype=torch.float32) *
                  torch.tensor(512.0, dtype=torch.float32))


# --------------------------------------------------------------------------
# E2M1
# --------------------------------------------------------------------------
def e2m1_encode(v: torch.Tensor) -> torch.Tensor:
    """fp32 -> 4-bit E2M1 code (uint8, low nibble), round-to-nearest-even,
    saturating at +-6, sign bit kept for -0 and for values that round to 0.

    Same function as mx_utils._f32_to_floatx_unpacked(x, 2, 1) (:152-266):
    the mid-points 0.25/0.75/1.25/1.75/2.5/3.5/5 go to the neighbour with an
    even mantissa, i.e. down/up/down/up/down/up/down.
    """
    assert v.dtype == torch.float32
    a = v.abs()
    code = ((a > 0.25).to(torch.uint8) + (a >= 0.75).to(torch.uint8) +
            (a > 1.25).to(torch.uint8) + (a >= 1.75).to(torch.uint8) +
            (a > 2.5).to(torch.uint8) + (a >= 3.5).to(torch.uint8) +
            (a > 5.0).to(torch.uint8))
    sign = (v.view(torch.int32) < 0).to(torch.uint8) << 3
    return code | sign


def e2m1_decode(code: torch.Tensor) -> torch.Tensor:
    """4-bit E2M1 code (uint8) -> fp32 (mx_utils._floatx_unpacked_to_f32 :21-110)."""
    assert code.dtype == torch.uint8
    table = torch.tensor(E2M1_VALUES + tuple(-x for x in E2M1_VALUES),
                         dtype=torch.float32)
    return table[(code & 0xF).long()]


def pack_nibbles(codes: torch.Tensor) -> torch.Tensor:
    """[..., 2n] uint8 codes -> [..., n] bytes, even index in the LOW nibble
    (mx_utils.pack_uint4 :144-149; csrc/quantization/fp4/mxfp4_quant.h:231-232)."""
    assert codes.dtype == torch.uint8 and codes.shape[-1] % 2 == 0
    c = codes.reshape(*codes.shape[:-1], codes.shape[-1] // 2, 2)
    return (c[..., 1] << 4) | (c[..., 0] & 0xF)


def unpack_nibbles(packed: torch.Tensor) -> torch.Tensor:
    """Inverse of pack_nibbles (mx_utils.unpack_uint4 :122-135)."""
    p = packed.view(torch.uint8)
    out = torch.stack((p & 0xF, p >> 4), dim=-1)
    return out.reshape(*p.shape[:-1], p.shape[-1] * 2)


# --------------------------------------------------------------------------
# UE8M0
# --------------------------------------------------------------------------
def e8m0_to_f32(byte: torch.Tensor) -> torch.Tensor:
    """Biased exponent byte -> fp32 2^(e-127).  Byte 0 is 2^-127 (an fp32
    subnormal) exactly as ``Tensor.to(float32)`` of float8_e8m0fnu gives."""
    b = byte.view(torch.uint8)
    return b.view(E8M0).to(torch.float32)


def rceil_exponent(descale: torch.Tensor) -> torch.Tensor:
    """``clamp(ceil(log2(descale)), -127, 127) + 127`` as uint8, NaN -> 0xFF.
    mx_utils.to_mxfp/_to_mx_rceil :313-325.  Evaluated in fp32 like the
    reference (so it inherits torch's fp32 log2; see DESIGN.md 'ceil(log2)')."""
    e = torch.clamp(torch.ceil(torch.log2(descale)), min=-E8M0_BIAS,
                    max=E8M0_BIAS) + E8M0_BIAS
    return torch.where(torch.isnan(descale),
                       torch.tensor(0xFF, dtype=torch.uint8),
                       e.to(torch.uint8))


def rceil_exponent_exact(descale: torch.Tensor) -> torch.Tensor:
    """Same quantity from the fp32 bit pattern (what the CUDA kernels do):
    exact for every positive input, including the few-ulp-above-2^k inputs
    where fp32 log2 rounds down to k."""
    assert descale.dtype == torch.float32
    bits = descale.contiguous().view(torch.int32)
    exp = (bits >> 23) & 0xFF
    man = bits & 0x7FFFFF
    ce = exp - 127 + (man != 0).to(torch.int32)              # normal numbers
    sub = torch.where(man > 0x400000, -126, -127)            # subnormals / zero
    ce = torch.where(exp == 0, sub, ce)
    ce = torch.clamp(ce, -127, 127)
    out = (ce + 127).to(torch.uint8)
    return torch.where(torch.isnan(descale),
                       torch.tensor(0xFF, dtype=torch.uint8), out)


def mx_quantize(x: torch.Tensor, fmt: str = "mxfp8", block: int = MX_BLOCK,
                exact_exponent: bool = False):
    """OCP-MX block quantisation with RCEIL scale selection.

    Restates mx_utils.to_mxfp (:277-353):
      amax per ``block`` -> e = clamp(ceil(log2(amax / max_pos)), +-127) + 127
      descale = 2^(127-e)  (1.0 when e == 0)
      q = clamp(x * descale, +-max_pos)
      mxfp8: q -> float8_e4m3fn ; mxfp4: q -> bf16 -> E2M1 (RNE) -> packed bytes
    Returns (scale_e8m0 [.., K/block], data) like the reference.
    x must be fp32 or bf16 (fp16 callers upcast first, tests/test_mxfp4_quant.py:77).
    """
    assert x.dtype in (torch.float32, torch.bfloat16)
    assert x.shape[-1] % block == 0 and x.is_contiguous()
    shape = x.shape
    xb = x.reshape(*shape[:-1], shape[-1] // block, block)
    amax = xb.abs().amax(dim=-1, keepdim=True).to(torch.float32)
    xb = xb.to(torch.float32)
    max_pos = FP8_MAX[E4M3] if fmt == "mxfp8" else E2M1_MAX
    descale = amax / max_pos
    e = rceil_exponent_exact(descale) if exact_exponent else rceil_exponent(descale)
    inv = torch.where(e == 0, torch.tensor(1.0),
                      torch.exp2(E8M0_BIAS - e.to(torch.float32)))
    q = torch.clamp(xb * inv, min=-max_pos, max=max_pos)
    if fmt == "mxfp8":
        data = q.to(E4M3).reshape(shape)
    elif fmt == "mxfp4":
        codes = e2m1_encode(q.to(torch.bfloat16).to(torch.float32))
        data = pack_nibbles(codes.reshape(shape)).view(torch.float4_e2m1fn_x2)
    else:
        raise ValueError(fmt)
    return e.squeeze(-1).view(E8M0), data


def mx_dequantize(data: torch.Tensor, scale_e8m0: torch.Tensor,
                  block: int = MX_BLOCK) -> torch.Tensor:
    """MX data + E8M0 scales -> bf16, as mx_utils.from_blocked_format :113-119
    (fp32 multiply, then one rounding to bf16)."""
    if data.dtype == torch.float4_e2m1fn_x2 or data.dtype == torch.uint8:
        vals = e2m1_decode(unpack_nibbles(data.view(torch.uint8)))
    else:
        vals = data.to(torch.float32)
    s = torch.repeat_interleave(scale_e8m0.to(torch.float32), block, dim=-1)
    return (vals * s).to(torch.bfloat16)
