"""Parity of the CUDA quantisation kernels (through torch.ops -> C ABI) against
the CPU oracle and the committed golden fixtures.  Shapes follow the
reference's own parametrisation (tests/test_fp8_quant.py, test_mxfp4_quant.py,
test_fused_norm_quant.py, test_fused_silu_mul_*_quant.py,
test_fused_quant_activation.py), trimmed so the oracle finishes in seconds.

Bars (DESIGN.md "Parity"):
  * a1-a5: codes, E8M0 bytes and fp32 scales BIT-EXACT;
  * a6-a9 (rms_norm + quant): BIT-EXACT against the oracle with the variance reduced exactly (fp64 mean
    rounded once -- the kernels accumulate the exact squares in fp64); against torch's own fp32 summation
    order on this host: scales within TORCH_ORDER_MAX_ULP, codes at most one step apart on a measured
    < 1e-5 of elements;
  * a10 / a11 (silu_and_mul + static / per-block): BIT-EXACT, both input types;
  * a12 (silu_and_mul + MXFP4, fp32 product): at most 1e-6 of codes one step apart, scale bytes equal.
"""
import os

import pytest
import torch

from oracle import formats, quant
from parity_util import code_stats, nibble_stats, scale_stats

pytestmark = pytest.mark.gpu

E4, E5 = torch.float8_e4m3fn, torch.float8_e5m2
GOLD = os.path.join(os.path.dirname(__file__), "golden")
DEV = "cuda"


@pytest.fixture(scope="module")
def ops():
    from vllm_xpu_kernels_b200 import ops as _ops
    return _ops


def load(name):
    return torch.load(os.path.join(GOLD, name + ".pt"), weights_only=False)


def u8(t):
    return t.contiguous().view(torch.uint8).cpu()


def ulps(a, b):
    a = a.detach().cpu().contiguous().float().view(torch.int32).long()
    b = b.detach().cpu().contiguous().float().view(torch.int32).long()
    return (a - b).abs().max().item() if a.numel() else 0


def ordered(code_u8):
    """sign-magnitude 8-bit float / int8 code -> monotone integer."""
    c = code_u8.to(torch.int16)
    return torch.where(c >= 128, 127 - c, c)


def assert_codes_close(got, ref, max_frac, what, is_int8=False):
    got, ref = u8(got), u8(ref)
    assert got.shape == ref.shape
    neq = got != ref
    frac = neq.float().mean().item() if got.numel() else 0.0
    if neq.any():
        if is_int8:
            d = (got.view(torch.int8).to(torch.int16) - ref.view(torch.int8).to(torch.int16)).abs()
        else:
            d = (ordered(got) - ordered(ref)).abs()
        assert d.max().item() <= 1, f"{what}: codes differ by more than one step"
    assert frac <= max_frac, f"{what}: {frac:.2e} of codes differ (allowed {max_frac:.1e})"


def seed(s=0):
    torch.manual_seed(s)


# ------------------------------------------------------------------ a1
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("qd", [E4, E5])
@pytest.mark.parametrize("T,H", [(1, 1), (7, 3), (7, 67), (83, 768), (83, 1024), (7, 1025), (7, 1031),
                                 (83, 5137), (7, 8192), (7, 8193), (1024, 4096), (3, 40960)])
@pytest.mark.parametrize("use_ub", [False, True])
def test_a1_dynamic_per_token(ops, dtype, qd, T, H, use_ub):
    seed()
    x = torch.rand(T, H, dtype=dtype) + 1e-6
    ub = torch.mean(x).to(torch.float32) if use_ub else None
    rq, rs = quant.dynamic_per_token_scaled_fp8_quant(x, qd, ub)
    q, s = ops.scaled_fp8_quant(x.to(DEV), scale_ub=None if ub is None else ub.to(DEV),
                                use_per_token_if_dynamic=True, fp8_dtype=qd)
    assert torch.equal(s.cpu(), rs)
    assert torch.equal(u8(q), u8(rq))


def test_a1_golden(ops):
    for c in load("a1_dynamic_per_token"):
        qd = E4 if "e4m3" in c["qdtype"] else E5
        ub = c["ub"]
        q, s = ops.scaled_fp8_quant(c["x"].to(DEV), scale_ub=None if ub is None else ub.to(DEV),
                                    use_per_token_if_dynamic=True, fp8_dtype=qd)
        assert torch.equal(s.cpu(), c["s"]) and torch.equal(u8(q), c["q"])


def test_a1_randn_full_config0(ops):
    """BASELINE config 0: 4096 x 4096 bf16; randn exercises the division rounding."""
    seed(1)
    x = torch.randn(4096, 4096, dtype=torch.bfloat16)
    rq, rs = quant.dynamic_per_token_scaled_fp8_quant(x, E4)
    q, s = ops.scaled_fp8_quant(x.to(DEV), use_per_token_if_dynamic=True, fp8_dtype=E4)
    assert torch.equal(s.cpu(), rs)
    assert torch.equal(u8(q), u8(rq))


# ------------------------------------------------------------------ a2 / a3
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("qd", [E4, E5])
@pytest.mark.parametrize("T,H", [(1, 1), (7, 16), (83, 67), (83, 2048), (7, 5137), (512, 1024)])
def test_a3_dynamic_per_tensor_and_a2_static(ops, dtype, qd, T, H):
    seed()
    x = torch.rand(T, H, dtype=dtype)
    rq, rs = quant.dynamic_scaled_fp8_quant(x, qd)
    q, s = ops.scaled_fp8_quant(x.to(DEV), fp8_dtype=qd)
    assert torch.equal(s.cpu(), rs)
    assert torch.equal(u8(q), u8(rq))
    q2, _ = ops.scaled_fp8_quant(x.to(DEV), scale=s, fp8_dtype=qd)
    assert torch.equal(u8(q2), u8(rq))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("gs", [(-1, -1), (-1, 1), (1, -1), (-1, 128), (1, 128), (128, 128), (1, 64),
                                (1, 16), (4, 256)])
def test_a2_static_group_shapes(ops, dtype, gs):
    seed()
    T, H = 128, 1024
    x = torch.rand(T, H, dtype=dtype)
    gm = T if gs[0] == -1 else gs[0]
    gn = H if gs[1] == -1 else gs[1]
    scale = (torch.rand(T // gm, H // gn) * 0.01 + 0.002)
    ref = quant.static_scaled_fp8_quant(x, scale, E4, gs)
    q, _ = ops.scaled_fp8_quant(x.to(DEV), scale=scale.to(DEV), fp8_dtype=E4, group_shape=gs)
    assert torch.equal(u8(q), u8(ref))
    if gs in ((1, -1), (-1, 1)):
        q1, _ = ops.scaled_fp8_quant(x.to(DEV), scale=scale.flatten().to(DEV), fp8_dtype=E4, group_shape=gs)
        assert torch.equal(u8(q1), u8(ref))


def test_a2_golden_reference_tolerance(ops):
    for c in load("a2_static_scaled_quantize"):
        q, _ = ops.scaled_fp8_quant(c["x"].to(DEV), scale=c["s"].to(DEV), fp8_dtype=E4,
                                    group_shape=tuple(c["group_shape"]))
        torch.testing.assert_close(q.float().cpu(), c["q"].view(E4).float(), rtol=0.12, atol=1e-5)


def test_a2_strided_rows_and_errors(ops):
    seed()
    big = torch.rand(16, 512, dtype=torch.bfloat16)
    x = big[:, :256]
    s = torch.tensor(0.01)
    out = torch.empty(16, 256, dtype=E4, device=DEV)
    ops.static_scaled_fp8_quant(out, x.to(DEV)[:, :], s.to(DEV))
    xs = big.to(DEV)[:, :256]
    assert not xs.is_contiguous()
    out2 = torch.empty(16, 256, dtype=E4, device=DEV)
    ops.static_scaled_fp8_quant(out2, xs, s.to(DEV))
    ref = quant.static_scaled_fp8_quant(x.contiguous(), s, E4)
    assert torch.equal(u8(out2), u8(ref)) and torch.equal(u8(out), u8(ref))
    with pytest.raises(RuntimeError, match="1-D scale needs group_shape"):
        ops.static_scaled_fp8_quant(out, xs, torch.ones(16, device=DEV))
    with pytest.raises(RuntimeError, match="must be divisible"):
        ops.static_scaled_fp8_quant(out, xs, torch.ones(3, 2, device=DEV))


# ------------------------------------------------------------------ a4
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("G", [32, 64, 128, 256, 16, 24])
@pytest.mark.parametrize("T", [1, 3, 31, 300])
@pytest.mark.parametrize("col_major,tma", [(False, False), (True, False), (True, True)])
def test_a4_per_token_group(ops, dtype, G, T, col_major, tma):
    seed()
    H = 768
    x = torch.rand(T, H, dtype=dtype) + 1e-6
    rq, rs = quant.per_token_group_fp8_quant(x, G)
    q, s = ops.per_token_group_quant_fp8(x.to(DEV), G, column_major_scales=col_major,
                                         tma_aligned_scales=tma, use_ue8m0=False)
    assert torch.equal(s.cpu(), rs)
    assert torch.equal(u8(q), u8(rq))
    if col_major:
        assert s.stride(-2) == 1
        if tma:
            assert s.stride(-1) == ops.get_tma_aligned_size(T, 4)
    # the reference's own (looser) bar against its literal bf16-arithmetic ref
    lq, ls = quant.per_token_group_fp8_quant_literal(x, G)
    assert torch.allclose(q.float().cpu(), lq.float(), atol=0.15, rtol=0.15)
    assert torch.allclose(s.cpu(), ls, atol=0.01, rtol=0.01)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("G", [32, 128])
@pytest.mark.parametrize("col_major", [False, True])
def test_a4_mxfp8(ops, dtype, G, col_major):
    seed()
    x = torch.rand(31, 512, dtype=dtype) + 1e-6
    rq, rs = quant.per_token_group_fp8_quant(x, G, scale_ue8m0=True)
    q, s = ops.per_token_group_quant_fp8(x.to(DEV), G, column_major_scales=col_major, use_ue8m0=True)
    assert s.dtype == torch.float8_e8m0fnu
    assert torch.equal(u8(s), u8(rs.to(torch.float8_e8m0fnu)))
    assert torch.equal(u8(q), u8(rq))


def test_a4_golden(ops):
    for c in load("a4_mxfp8"):
        q, s = ops.per_token_group_quant_fp8(c["x"].to(DEV), 32, use_ue8m0=True)
        assert torch.equal(u8(q), c["q"]) and torch.equal(u8(s), c["s_e8m0"])
    for c in load("a4_per_block"):
        q, s = ops.per_token_group_quant_fp8(c["x"].to(DEV), c["G"], use_ue8m0=False)
        assert torch.allclose(q.float().cpu(), c["q"].view(E4).float(), atol=0.15, rtol=0.15)
        assert torch.allclose(s.cpu(), c["s"], atol=0.01, rtol=0.01)
        if c["x"].dtype == torch.float32:
            assert torch.equal(u8(q), c["q"]) and torch.equal(s.cpu(), c["s"])


# ------------------------------------------------------------------ a5
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T", [1, 2, 4, 8, 257])
@pytest.mark.parametrize("H", [32, 64, 128, 256, 512, 4096])
@pytest.mark.parametrize("col_major", [False, True])
def test_a5_mxfp4(ops, dtype, T, H, col_major):
    seed()
    x = torch.rand(T, H, dtype=dtype) + 1e-6
    rq, rs, re = quant.per_token_group_quant_mxfp4(x)
    q, s = ops.per_token_group_quant_mxfp4(x.to(DEV), 32, column_major_scales=col_major)
    assert q.dtype == torch.float4_e2m1fn_x2 and s.dtype == torch.float8_e8m0fnu
    assert torch.equal(u8(s), re)
    assert torch.equal(u8(q), rq)
    if col_major:
        assert s.stride(0) == 1 and s.stride(1) == T


def test_a5_randn_and_golden(ops):
    seed(3)
    for dtype in (torch.float32, torch.bfloat16, torch.float16):
        x = torch.randn(512, 2048, dtype=dtype) * 3.0
        x[5, 64:96] = 0.0          # all-zero group -> exponent byte 0
        rq, rs, re = quant.per_token_group_quant_mxfp4(x)
        q, s = ops.per_token_group_quant_mxfp4(x.to(DEV))
        assert torch.equal(u8(s), re) and torch.equal(u8(q), rq)
    for c in load("a5_mxfp4"):
        q, s = ops.per_token_group_quant_mxfp4(c["x"].to(DEV))
        assert torch.equal(u8(s), c["s_e8m0"]) and torch.equal(u8(q), c["q"])


def test_e2m1_hardware_conversion_matches_reference_on_every_bf16(ops):
    """Every finite bf16 bit pattern through the MXFP4 kernel with the group scale pinned to 1
    (group max 6.0 -> e8m0 byte 127), against the reference's _f32_to_floatx_unpacked table."""
    g = load("e2m1_all_bf16")
    allbf16 = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)
    v = torch.where(g["finite"], allbf16.float(), torch.zeros(())).clamp(-6.0, 6.0).to(torch.bfloat16)
    x = torch.zeros(65536 // 16, 32, dtype=torch.bfloat16)
    x[:, :16] = v.reshape(-1, 16)
    x[:, 16] = 6.0
    q, s = ops.per_token_group_quant_mxfp4(x.to(DEV))
    assert (u8(s) == 127).all()
    codes = formats.unpack_nibbles(u8(q))[:, :16].reshape(-1)
    ref = torch.where(g["finite"], g["codes"], torch.zeros((), dtype=torch.uint8))
    vv = torch.where(g["finite"], allbf16.float(), torch.zeros(()))
    inr = vv.abs() <= 6.0
    assert torch.equal(codes[inr], ref[inr])


# ------------------------------------------------------------------ a6 .. a9
# Bar: BIT-EXACT scales, E8M0 bytes, codes and residual against the oracle with the variance reduced
# exactly (fp64 mean, rounded once -- oracle.quant.rms_norm_f32(exact_variance=True)); torch's own fp32
# summation order on this host is the second, measured check (its variance is a few ulp off the exact one).
def _norm_inputs(T, H, dtype, add_res, strided=False):
    seed()
    w = torch.empty(H, dtype=dtype).normal_(mean=1.0, std=0.1)
    last = 2 * H if strided else H
    x = (torch.randn(T, last, dtype=dtype) * (1.0 / H))[..., :H]
    res = torch.randn(T, H, dtype=dtype) * (1.0 / H) if add_res else None
    return x, w, res


NORM_SHAPES = [(1, 64), (7, 128), (83, 1024), (7, 5120), (64, 4096), (5, 8192), (3, 1000), (2, 16384)]
TORCH_ORDER_MAX_ULP = 8          # measured: <= 4 (profiles/parity_r02.txt)
TORCH_ORDER_MAX_CODE_FRAC = 1e-3  # measured: <= 6e-6 at T = 8192


def _check_vs_torch_order(q, s, rq, rs, what, is_int8=False):
    n, frac, d = code_stats(q, rq, is_int8)
    assert d <= 1 and frac <= TORCH_ORDER_MAX_CODE_FRAC, f"{what}: {n} codes differ, max {d} steps"
    if s is not None:
        assert scale_stats(s, rs)[2] <= TORCH_ORDER_MAX_ULP, f"{what}: scales {scale_stats(s, rs)}"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("qd", [E4, torch.int8])
@pytest.mark.parametrize("T,H", NORM_SHAPES)
@pytest.mark.parametrize("add_res", [False, True])
def test_a6_rms_norm_dynamic_per_token(ops, dtype, qd, T, H, add_res):
    x, w, res = _norm_inputs(T, H, dtype, add_res)
    dres = res.clone().to(DEV) if add_res else None
    q, s = ops.rms_norm_dynamic_per_token_quant(x.to(DEV), w.to(DEV), 1e-6, qd, None, dres)
    rq, rs, rres = quant.rms_norm_dynamic_per_token_quant(x, w, 1e-6, qd, None, res, exact_variance=True)
    assert torch.equal(s.cpu(), rs), f"scales: {scale_stats(s, rs)}"
    assert torch.equal(u8(q), u8(rq)), f"codes: {code_stats(q, rq, qd == torch.int8)}"
    if add_res:
        assert torch.equal(dres.cpu(), rres)       # residual update is exact
    tq, ts, _ = quant.rms_norm_dynamic_per_token_quant(x, w, 1e-6, qd, None, res, exact_variance=False)
    _check_vs_torch_order(q, s, tq, ts, "a6 vs torch order", qd == torch.int8)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("H", [1024, 4096])
@pytest.mark.parametrize("eps", [1e-6, 1e-30])
def test_a6_a9_variance_over_the_whole_input_range(ops, dtype, H, eps):
    """The 16-bit static and MXFP4 kernels sum exact fp32 squares widened to fp64 by integer operations and fall back to
    real conversions when that sum leaves its trusted window (quant_row_impl.cuh, widen_sq): one row per magnitude
    from squares that flush in fp32 to squares that overflow it, all-zero rows, rows that mix the extremes,
    and -0.0.  Same bar as every a6 / a7 test: bit-exact vs the exact-variance oracle."""
    seed()
    mags = ([1e-30, 1e-25, 1e-20, 1e-15, 1e-12, 1e-10, 1e-8, 1e-4, 1.0, 1e6, 1e15, 1e19, 3e19, 1e25]
            if dtype == torch.bfloat16 else [1e-7, 1e-6, 1e-5, 1e-3, 1.0, 30.0, 1e3, 1e4])
    rows = [torch.randn(H) * m for m in mags]
    rows.append(torch.zeros(H))
    z = torch.zeros(H); z[::2] = -0.0; z[5] = 1e-3 if dtype == torch.float16 else 1e-22
    rows.append(z)
    mix = torch.randn(H) * mags[0]; mix[H // 2:] = torch.randn(H - H // 2) * mags[-3]
    rows.append(mix)
    one = torch.zeros(H); one[7] = 1.0
    rows.append(one)
    x = torch.stack(rows).to(dtype)
    w = torch.empty(H, dtype=dtype).normal_(mean=1.0, std=0.1)
    q, s_ = ops.rms_norm_dynamic_per_token_quant(x.to(DEV), w.to(DEV), eps, E4, None, None)
    rq, rs, _ = quant.rms_norm_dynamic_per_token_quant(x, w, eps, E4, None, None, exact_variance=True)
    fin = torch.isfinite(rs)
    assert torch.equal(s_.cpu()[fin], rs[fin]) and torch.equal(torch.isnan(s_.cpu()), torch.isnan(rs)), \
        f"a6 scales differ in rows {(s_.cpu() != rs).nonzero().flatten().tolist()}"
    assert torch.equal(u8(q)[fin], u8(rq)[fin])
    q, s_ = ops.rms_norm_per_block_quant(x.to(DEV), w.to(DEV), eps, 128, E4, None)
    rq, rs, _ = quant.rms_norm_per_block_quant(x, w, eps, 128, E4, None, exact_variance=True)
    fin = torch.isfinite(rs).all(dim=1)
    assert torch.equal(s_.cpu()[fin], rs[fin]), f"a7 scales differ in rows {(s_.cpu() != rs).any(1).nonzero().flatten().tolist()}"
    assert torch.equal(u8(q)[fin], u8(rq)[fin])
    q, s_ = ops.rms_norm_mxfp4_quant(x.to(DEV), w.to(DEV), eps, None)
    rq, rs, _ = quant.rms_norm_mxfp4_quant(x, w, eps, None, exact_variance=True)
    bad = (s_.cpu() != rs).any(1) | (u8(q) != rq).any(1)
    assert not bad.any(), f"a8 differs in rows {bad.nonzero().flatten().tolist()}"
    scale = torch.tensor(0.02, dtype=torch.float32)
    q = ops.rms_norm_static_fp8_quant(x.to(DEV), w.to(DEV), scale.to(DEV), eps, None)
    rq, _ = quant.rms_norm_static_fp8_quant(x, w, scale, eps, None, exact_variance=True)
    bad = (u8(q) != u8(rq)).any(1)
    assert not bad.any(), f"a9 differs in rows {bad.nonzero().flatten().tolist()}"


def test_a6_scale_ub(ops):
    x, w, _ = _norm_inputs(83, 1024, torch.bfloat16, False)
    normed, _ = quant.rms_norm_f32(x, w, 1e-6, exact_variance=True)
    ub = torch.tensor(0.5 * float(normed.abs().max()))
    rq, rs, _ = quant.rms_norm_dynamic_per_token_quant(x, w, 1e-6, E4, ub, exact_variance=True)
    q, s = ops.rms_norm_dynamic_per_token_quant(x.to(DEV), w.to(DEV), 1e-6, E4, ub.to(DEV))
    assert torch.equal(s.cpu(), rs) and torch.equal(u8(q), u8(rq))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("qd", [E4, torch.int8])
@pytest.mark.parametrize("T,H", [(1, 128), (7, 1024), (83, 5120), (33, 7168), (4, 1536)])
@pytest.mark.parametrize("G", [64, 128])
@pytest.mark.parametrize("add_res", [False, True])
@pytest.mark.parametrize("transposed", [False, True])
def test_a7_rms_norm_per_block(ops, dtype, qd, T, H, G, add_res, transposed):
    x, w, res = _norm_inputs(T, H, dtype, add_res)
    rq, rs, rres = quant.rms_norm_per_block_quant(x, w, 1e-6, G, qd, res, exact_variance=True)
    dres = res.clone().to(DEV) if add_res else None
    q, s = ops.rms_norm_per_block_quant(x.to(DEV), w.to(DEV), 1e-6, G, qd, dres, transposed)
    assert s.shape == (T, H // G)
    if transposed:
        assert s.stride(0) == 1
    assert torch.equal(s.contiguous().cpu(), rs), f"scales: {scale_stats(s, rs)}"
    assert torch.equal(u8(q), u8(rq)), f"codes: {code_stats(q, rq, qd == torch.int8)}"
    if add_res:
        assert torch.equal(dres.cpu(), rres)
    tq, ts, _ = quant.rms_norm_per_block_quant(x, w, 1e-6, G, qd, res, exact_variance=False)
    _check_vs_torch_order(q, s, tq, ts, "a7 vs torch order", qd == torch.int8)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,H", [(7, 128), (83, 1024), (9, 5120)])
@pytest.mark.parametrize("add_res", [False, True])
def test_a7_mxfp8_ue8m0(ops, dtype, T, H, add_res):
    x, w, res = _norm_inputs(T, H, dtype, add_res)
    rq, rs, _ = quant.rms_norm_per_block_quant(x, w, 1e-6, 32, E4, res, scale_ue8m0=True, exact_variance=True)
    dres = res.clone().to(DEV) if add_res else None
    q, s = ops.rms_norm_per_block_quant(x.to(DEV), w.to(DEV), 1e-6, 32, E4, dres, False, True)
    log2 = torch.log2(s.cpu())
    assert torch.equal(log2, log2.round())                       # exact powers of two
    assert torch.equal(s.cpu(), rs) and torch.equal(u8(q), u8(rq))


def test_a7_noncontiguous_rows(ops):
    """DeepSeek-V3 MLA pattern: row stride hidden + 576 (test_fused_norm_quant.py:320-421)."""
    seed()
    T, H, extra = 83, 1536, 576
    w = torch.empty(H, dtype=torch.bfloat16).normal_(mean=1.0, std=0.1).to(DEV)
    fused = (torch.randn(T, H + extra, dtype=torch.bfloat16) / H).to(DEV)
    xnc, _ = fused.split([H, extra], dim=-1)
    assert not xnc.is_contiguous()
    q1, s1 = ops.rms_norm_per_block_quant(xnc, w, 1e-6, 128, E4)
    q2, s2 = ops.rms_norm_per_block_quant(xnc.contiguous(), w, 1e-6, 128, E4)
    assert torch.equal(s1, s2) and torch.equal(u8(q1), u8(q2))
    rq, rs, _ = quant.rms_norm_per_block_quant(xnc.contiguous().cpu(), w.cpu(), 1e-6, 128, E4, exact_variance=True)
    assert torch.equal(s1.cpu(), rs) and torch.equal(u8(q1), u8(rq))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,H", [(1, 128), (7, 1024), (83, 5120), (64, 4096)])
@pytest.mark.parametrize("add_res", [False, True])
def test_a8_rms_norm_mxfp4(ops, dtype, T, H, add_res):
    x, w, res = _norm_inputs(T, H, dtype, add_res)
    rq, rs, rres = quant.rms_norm_mxfp4_quant(x, w, 1e-6, res, exact_variance=True)
    dres = res.clone().to(DEV) if add_res else None
    q, s = ops.rms_norm_mxfp4_quant(x.to(DEV), w.to(DEV), 1e-6, dres)
    assert torch.equal(s.cpu(), rs), "E8M0 scales"
    assert torch.equal(u8(q), rq), f"codes: {nibble_stats(q, rq)}"
    if add_res:
        assert torch.equal(dres.cpu(), rres)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,H", [(1, 64), (7, 128), (83, 1024), (7, 5120), (5, 1000)])
@pytest.mark.parametrize("strided", [False, True])
@pytest.mark.parametrize("add_res", [False, True])
def test_a9_rms_norm_static(ops, dtype, T, H, strided, add_res):
    x, w, res = _norm_inputs(T, H, dtype, add_res, strided)
    for sv in (1.0, 0.02):
        scale = torch.tensor(sv, dtype=torch.float32)
        rq, rres = quant.rms_norm_static_fp8_quant(x, w, scale, 1e-6, res, exact_variance=True)
        dres = res.clone().to(DEV) if add_res else None
        xd = (torch.zeros(T, 2 * H, dtype=dtype, device=DEV) if strided else None)
        if strided:
            xd[:, :H] = x.to(DEV)
            xin = xd[:, :H]
            assert T == 1 or not xin.is_contiguous()
        else:
            xin = x.to(DEV)
        q = ops.rms_norm_static_fp8_quant(xin, w.to(DEV), scale.to(DEV), 1e-6, dres)
        assert torch.equal(u8(q), u8(rq)), f"codes: {code_stats(q, rq)}"
        if add_res:
            assert torch.equal(dres.cpu(), rres)


def test_a6_a9_golden(ops):
    """The committed outputs of the reference's own functions (torch fp32 summation order of the build
    container): the second, measured check."""
    for c in load("a6_a9_rms_norm"):
        x, w, res = c["x"], c["w"], c["res"]
        for qd in (E4, torch.int8):
            dres = res.clone().to(DEV) if res is not None else None
            q, s = ops.rms_norm_dynamic_per_token_quant(x.to(DEV), w.to(DEV), 1e-6, qd, None, dres)
            gq, gs = c[f"tok_{qd}"]
            _check_vs_torch_order(q, s, gq, gs, "golden a6", qd == torch.int8)
            if res is not None:
                assert torch.equal(dres.cpu(), c["res_out"])
            for G in (64, 128):
                dres = res.clone().to(DEV) if res is not None else None
                q, s = ops.rms_norm_per_block_quant(x.to(DEV), w.to(DEV), 1e-6, G, qd, dres)
                gq, gs = c[f"grp{G}_{qd}"]
                _check_vs_torch_order(q, s, gq, gs, "golden a7", qd == torch.int8)
        dres = res.clone().to(DEV) if res is not None else None
        q = ops.rms_norm_static_fp8_quant(x.to(DEV), w.to(DEV), torch.tensor(0.02, device=DEV), 1e-6, dres)
        _check_vs_torch_order(q, None, c["static_0p02"], None, "golden a9")
        dres = res.clone().to(DEV) if res is not None else None
        q, s = ops.rms_norm_mxfp4_quant(x.to(DEV), w.to(DEV), 1e-6, dres)
        gq, ge = c["mxfp4"]
        assert (u8(s.to(torch.float8_e8m0fnu)) != ge).float().mean().item() <= 1e-3
        assert nibble_stats(q, gq)[1] <= 1e-3


# ------------------------------------------------------------------ a10 .. a12
# Bar: a10 / a11 BIT-EXACT (silu(gate) is rounded to the 16-bit input type at once, and
# tests/test_silu_exhaustive_gpu.py proves the kernels' silu -> T equals torch's on all 65536 inputs of
# either type); a12 keeps the fp32 product and is held to its measured worst case (profiles/parity_r02.txt:
# at most 5 of 5.9e7 codes one step apart, no scale byte differs), stated here as <= 1e-6.
A12_MAX_FRAC = 1e-6


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("qd", [E4, E5])
@pytest.mark.parametrize("T,d", [(1, 16), (7, 128), (83, 512), (512, 4096), (7, 100), (5, 14336)])
@pytest.mark.parametrize("mag", [1.0, 8.0])
def test_a10_silu_and_mul_quant(ops, dtype, qd, T, d, mag):
    seed()
    x = torch.randn(T, 2 * d, dtype=dtype) * mag
    scale = torch.tensor([0.5], dtype=torch.float32)
    ref = quant.silu_and_mul_quant(x, scale, qd)
    out = torch.empty(T, d, dtype=qd, device=DEV)
    ops.silu_and_mul_quant(out, x.to(DEV), scale.to(DEV))
    assert torch.equal(u8(out), u8(ref)), f"codes: {code_stats(out, ref)}"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("qd", [E4, E5, torch.int8])
@pytest.mark.parametrize("T,H", [(1, 64), (1, 1024), (16, 5120), (128, 1024), (512, 5120), (3, 14336)])
@pytest.mark.parametrize("G", [32, 64, 128])
@pytest.mark.parametrize("transposed", [False, True])
def test_a11_silu_and_mul_per_block(ops, dtype, qd, T, H, G, transposed):
    if H % G:
        pytest.skip("hidden not divisible by group")
    seed()
    x = torch.randn(T, 2 * H, dtype=dtype) * (1.0 / H)
    rq, rs = quant.silu_and_mul_per_block_quant(x, G, qd)
    q, s = ops.silu_and_mul_per_block_quant(x.to(DEV), G, qd, None, transposed)
    assert s.shape == (T, H // G)
    assert torch.equal(s.contiguous().cpu(), rs), f"scales: {scale_stats(s, rs)}"
    assert torch.equal(u8(q), u8(rq)), f"codes: {code_stats(q, rq, qd == torch.int8)}"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("mag", [1.0, 1.0 / 64, 8.0])
def test_a11_full_width_magnitudes(ops, dtype, mag):
    """Llama-3-8B intermediate width, gate magnitudes from tiny to saturating."""
    seed()
    x = torch.randn(256, 2 * 14336, dtype=dtype) * mag
    for qd in (E4, torch.int8):
        rq, rs = quant.silu_and_mul_per_block_quant(x, 128, qd)
        q, s = ops.silu_and_mul_per_block_quant(x.to(DEV), 128, qd)
        assert torch.equal(s.cpu(), rs) and torch.equal(u8(q), u8(rq)), (scale_stats(s, rs), code_stats(q, rq))


def test_a11_scale_ub_and_ue8m0(ops):
    seed()
    T, H, G = 64, 1024, 128
    x = torch.randn(T, 2 * H, dtype=torch.bfloat16) * (1.0 / H)
    _, s0 = quant.silu_and_mul_per_block_quant(x, G, E4)
    ub = s0.median().reshape(1)
    rq, rs = quant.silu_and_mul_per_block_quant(x, G, E4, scale_ub=ub)
    q, s = ops.silu_and_mul_per_block_quant(x.to(DEV), G, E4, ub.to(DEV))
    assert torch.equal(s.cpu(), rs) and torch.equal(u8(q), u8(rq))
    rq, rs = quant.silu_and_mul_per_block_quant(x, G, E4, scale_ue8m0=True)
    q, s = ops.silu_and_mul_per_block_quant(x.to(DEV), G, E4, None, False, True)
    assert torch.equal(s.cpu(), rs) and torch.equal(u8(q), u8(rq))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,H", [(1, 128), (7, 1024), (83, 5120), (2048, 1024), (2, 28672)])
@pytest.mark.parametrize("mag", [None, 1.0])
def test_a12_silu_and_mul_mxfp4(ops, dtype, T, H, mag):
    seed()
    x = torch.randn(T, 2 * H, dtype=dtype) * (mag if mag else 1.0 / H)
    rq, rs = quant.silu_and_mul_mxfp4_quant(x)
    q, s = ops.silu_and_mul_mxfp4_quant(x.to(DEV))
    log2 = torch.log2(s.cpu())
    assert torch.equal(log2, log2.round())
    n_s = int((s.cpu() != rs).sum())
    n_c, frac_c, d_c = nibble_stats(q, rq)
    assert n_s <= max(1, A12_MAX_FRAC * s.numel()), f"{n_s} scale bytes differ"
    # a differing scale byte re-codes its whole group: compare against the formula with the kernel's scales
    given = quant.mxfp4_codes_given_exponent(quant.silu_mul_f32(x), u8(s.to(torch.float8_e8m0fnu)))
    n_g, frac_g, d_g = nibble_stats(q, given)
    assert d_g <= 1 and n_g <= max(1, A12_MAX_FRAC * rq.numel() * 2), f"{n_g} codes differ (max {d_g} steps)"
    if n_s == 0:
        assert d_c <= 1 and n_c <= max(1, A12_MAX_FRAC * rq.numel() * 2)


def test_a10_a12_golden(ops):
    for c in load("a10_silu_static"):
        qd = E4 if "e4m3" in c["qdtype"] else E5
        out = torch.empty(c["x"].shape[0], c["x"].shape[1] // 2, dtype=qd, device=DEV)
        ops.silu_and_mul_quant(out, c["x"].to(DEV), c["scale"].to(DEV))
        assert torch.equal(u8(out), c["q"])
    for c in load("a11_silu_block"):
        qd = E4 if "e4m3" in c["qdtype"] else torch.int8
        q, s = ops.silu_and_mul_per_block_quant(c["x"].to(DEV), c["G"], qd)
        assert torch.equal(s.cpu(), c["s"]) and torch.equal(u8(q), c["q"])
    for c in load("a12_silu_mxfp4"):
        q, s = ops.silu_and_mul_mxfp4_quant(c["x"].to(DEV))
        assert torch.equal(s.cpu(), c["s"])
        assert nibble_stats(q, c["q"])[0] <= 1


@pytest.mark.parametrize("fast", [0, 1])
@pytest.mark.parametrize("mag", [1.0, 1.0 / 64, 8.0])
def test_silu_modes_forced(ops, fast, mag):
    """vxk_set_silu_fast(0 / 1) overrides the auto choice; both forms stay within one code step on a
    vanishing fraction of elements (bf16: the two are identical, see the exhaustive test)."""
    from vllm_xpu_kernels_b200 import _loader
    lib = _loader.cdll()
    seed()
    x = torch.randn(256, 2 * 4096, dtype=torch.bfloat16) * mag
    rq, rs = quant.silu_and_mul_per_block_quant(x, 128, E4)
    r4, rs4 = quant.silu_and_mul_mxfp4_quant(x)
    prev = lib.vxk_set_silu_fast(fast)
    try:
        q, s = ops.silu_and_mul_per_block_quant(x.to(DEV), 128, E4)
        q4, s4 = ops.silu_and_mul_mxfp4_quant(x.to(DEV))
    finally:
        lib.vxk_set_silu_fast(prev)
    assert torch.equal(s.cpu(), rs) and torch.equal(u8(q), u8(rq))
    assert torch.equal(s4.cpu(), rs4)
    n, frac, d = nibble_stats(q4, r4)
    assert d <= 1 and n <= 2


def test_silu_extreme_gates(ops):
    """Gates far outside the sigmoid's useful range (exp overflow side) next to ordinary ones."""
    seed()
    x = torch.randn(64, 2 * 1024, dtype=torch.bfloat16)
    x[:, 0:1024:7] = -60.0
    x[:, 1:1024:7] = -90.0
    x[:, 2:1024:11] = 100.0
    x[3, :1024] = -80.0          # a whole row of vanishing silu values
    rq, rs = quant.silu_and_mul_per_block_quant(x, 128, E4)
    q, s = ops.silu_and_mul_per_block_quant(x.to(DEV), 128, E4)
    assert torch.isfinite(s).all()
    assert torch.equal(s.cpu(), rs) and torch.equal(u8(q), u8(rq))


# ------------------------------------------------------------------ boundary behaviour
def test_empty_inputs_and_errors(ops):
    x = torch.empty(0, 128, dtype=torch.bfloat16, device=DEV)
    q, s = ops.scaled_fp8_quant(x, use_per_token_if_dynamic=True, fp8_dtype=E4)
    assert q.shape == (0, 128)
    q, s = ops.per_token_group_quant_mxfp4(x)
    assert q.numel() == 0 and s.numel() == 0
    q, s = ops.silu_and_mul_per_block_quant(torch.empty(0, 256, dtype=torch.bfloat16, device=DEV), 64, E4)
    assert q.shape == (0, 128)
    y = torch.rand(4, 64, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError, match="group_size == 32"):
        torch.ops._C.per_token_group_quant_mxfp4(y, torch.empty(4, 32, dtype=torch.uint8, device=DEV),
                                                 torch.empty(4, 1, device=DEV), 64, 1e-10)
    with pytest.raises(RuntimeError, match="group_size must be 32, 64, or 128"):
        ops.silu_and_mul_per_block_quant(torch.rand(4, 512, dtype=torch.bfloat16, device=DEV), 256, E4)
    with pytest.raises(RuntimeError):
        ops.scaled_fp8_quant(y.t(), use_per_token_if_dynamic=True, fp8_dtype=E4)  # .contiguous() inside -> ok
        torch.ops._C.dynamic_per_token_scaled_fp8_quant(torch.empty(64, 4, dtype=E4, device=DEV), y.t(),
                                                        torch.empty(64, 1, device=DEV), None)
    with pytest.raises(NotImplementedError):
        torch.ops._C.silu_and_mul_quant(torch.empty(4, 32, dtype=E4), torch.rand(4, 64, dtype=torch.bfloat16),
                                        torch.ones(1))  # CPU tensors: no CPU fallback is registered


def test_large_row_count_int32_overflow_regression(ops):
    """test_fp8_quant.py:401-425 (1,024,000 x 1152 bf16: element index > 2^31 / 2)."""
    T, H = 1024000, 1152
    seed()
    x = torch.rand(T, H, dtype=torch.bfloat16, device=DEV)
    scale = (x.abs().max().float() / 448.0).reshape(1)
    out = torch.empty(T, H, dtype=E4, device=DEV)
    ops.static_scaled_fp8_quant(out, x, scale)
    # size-independent property: every code equals the per-element quantisation of the same scale
    idx = torch.tensor([0, 1, T // 2, T - 2, T - 1], device=DEV)
    ref = quant.static_scaled_fp8_quant(x[idx].cpu(), scale.cpu(), E4)
    assert torch.equal(u8(out[idx]), u8(ref))
    q, s = ops.scaled_fp8_quant(x, use_per_token_if_dynamic=True, fp8_dtype=E4)
    rq, rs = quant.dynamic_per_token_scaled_fp8_quant(x[idx].cpu(), E4)
    assert torch.equal(u8(q[idx]), u8(rq)) and torch.equal(s[idx].cpu(), rs)
