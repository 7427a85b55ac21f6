"""Parity of torch.ops._xpu_C.fp8_gemm_w8a16 (16-bit activations x FP8 weights -> C ABI ->
tcgen05 kind::f16 kernel with in-loop FP8 widening) against the CPU oracle.  Shapes and
operand construction follow the reference's tests/test_fp8_gemm_onednn.py:80-126,
224-267, 367-410.

Bar: |out - ref| <= 1e-2 + 1e-2 * |ref| against the exact-arithmetic reference (float64
dequantise + accumulate, one rounding to the activation dtype); the reference's own bar
(5e-2 / 6e-2 against a matmul in the activation dtype) is checked as well.
"""
import ctypes

import pytest
import torch

from oracle import gemm as og, quant as oq

pytestmark = pytest.mark.gpu
E4, E5 = torch.float8_e4m3fn, torch.float8_e5m2
DEV = "cuda"
MNK_FACTORS = [(1, 32, 1024), (1, 32, 2048), (1, 32, 5120), (8, 512, 2048), (512, 1024, 2048), (1024, 2048, 2048)]


@pytest.fixture(scope="module")
def ops():
    from vllm_xpu_kernels_b200 import ops as _ops
    return _ops


@pytest.fixture(scope="module")
def lib():
    from vllm_xpu_kernels_b200 import _loader
    return _loader.cdll()


def gemm_stats(lib):
    a, b = ctypes.c_uint64(), ctypes.c_uint64()
    lib.vxk_gemm_stats(ctypes.byref(a), ctypes.byref(b))
    return a.value, b.value


def close(out, ref, atol=1e-2, rtol=1e-2):
    torch.testing.assert_close(out.float().cpu(), ref.float(), atol=atol, rtol=rtol)


@pytest.mark.parametrize("fp8", [E4, E5])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("trans_wei", [True, False])
@pytest.mark.parametrize("is_mbk", [True, False])
@pytest.mark.parametrize("batch", [1, 2, 8])
@pytest.mark.parametrize("mnk", MNK_FACTORS)
def test_fp8_gemm_w8a16(ops, lib, fp8, dtype, trans_wei, is_mbk, batch, mnk):
    """tests/test_fp8_gemm_onednn.py:80-126: weight of ones quantised with a per-tensor scale of 4."""
    torch.manual_seed(1234)
    m, n, k = mnk
    if batch > 1 and m > 8:
        pytest.skip("batched variants only on the small-M shapes (keeps the float64 CPU reference fast)")
    if is_mbk and batch > 1 and m > 1:
        pytest.skip("[m,b,k] view of a [b,m,k] tensor has no single row stride (same restriction as fp8_gemm)")
    x = torch.randn(batch, m, k, dtype=dtype) / 10.0
    w = torch.ones(n, k, dtype=dtype) if trans_wei else torch.ones(k, n, dtype=dtype)
    s_wei = torch.tensor(4.0)
    wq = oq.static_scaled_fp8_quant(w, s_wei, fp8)
    xin = x.to(DEV)
    if is_mbk:
        xin = xin.transpose(0, 1)
    tc0, fb0 = gemm_stats(lib)
    out = ops.fp8_gemm_w8a16(xin, wq.to(DEV).transpose(0, 1) if trans_wei else wq.to(DEV), s_wei.to(DEV),
                             torch.Tensor())
    tc1, fb1 = gemm_stats(lib)
    if trans_wei:
        assert (tc1 - tc0, fb1 - fb0) == (1, 0), "K-major weights must run on the tensor-core kernel"
    out = out.transpose(0, 1) if is_mbk else out
    assert out.dtype == dtype and out.shape == (batch, m, n)
    b_kn = wq.t() if trans_wei else wq
    ref = og.fp8_gemm_w8a16_ref(x.reshape(-1, k), b_kn, s_wei)
    close(out.reshape(-1, n), ref)
    w_nk = wq if trans_wei else wq.t()
    close(out.reshape(-1, n), og.fp8_gemm_w8a16_literal(x.reshape(-1, k), w_nk, s_wei), 5e-2, 5e-2)


@pytest.mark.parametrize("fp8", [E4, E5])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("is_nt", [True, False])
@pytest.mark.parametrize("scale_layout", ["flat_1d", "row_2d"])
@pytest.mark.parametrize("mnk", MNK_FACTORS + [(100, 200, 2048), (257, 129, 1040), (128, 128, 16), (3, 40, 72),
                                               # M > 256: the swap-AB TMEM-operand kernel (256-token tiles, 64-deep stages):
                                               # token tail, N tail (200 = 128 + 72), K = 17 x 64
                                               (300, 200, 1088), (777, 264, 4160), (1025, 8, 64)])
@pytest.mark.parametrize("bias", [False, True])
def test_fp8_gemm_w8a16_per_channel(ops, fp8, dtype, is_nt, scale_layout, mnk, bias):
    """tests/test_fp8_gemm_onednn.py:224-267: per-channel weight scales as [N] or [1, N]."""
    torch.manual_seed(1234)
    m, n, k = mnk
    x = torch.randn(m, k, dtype=dtype) / 10.0
    w = torch.randn(n, k, dtype=dtype) / 10.0
    wq, ws = oq.dynamic_per_token_scaled_fp8_quant(w, fp8)          # [N,K], [N,1]
    scale = ws.flatten() if scale_layout == "flat_1d" else ws.t().contiguous()
    wt = wq.to(DEV).transpose(0, 1)
    if is_nt:
        wt = wt.contiguous()       # reference naming: "is_nt" makes the [K,N] view contiguous (N-major)
    b = (torch.randn(n, dtype=dtype) / 4) if bias else None
    out = ops.fp8_gemm_w8a16(x.to(DEV), wt, scale.to(DEV), b.to(DEV) if bias else torch.Tensor())
    assert out.dtype == dtype and out.shape == (m, n)
    close(out, og.fp8_gemm_w8a16_ref(x, wq.t(), ws.flatten(), b))
    lit = og.fp8_gemm_w8a16_literal(x, wq, ws)
    if bias:
        lit = lit + b
    close(out, lit, 5e-2, 5e-2)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("is_nt", [True, False])
@pytest.mark.parametrize("batch", [1, 2, 8])
@pytest.mark.parametrize("group_size", [64, 128, 256])
@pytest.mark.parametrize("mnk", [(4, 256, 512), (1, 256, 256), (2, 256, 1024), (300, 512, 768)])
def test_fp8_gemm_w8a16_block(ops, dtype, is_nt, batch, group_size, mnk):
    """tests/test_fp8_gemm_onednn.py:367-410: weight scales [K/gs, N/gs], activations not quantised."""
    torch.manual_seed(1234)
    m, n, k = mnk
    if k % group_size or n % group_size:
        pytest.skip("shape not divisible by the group")
    x = torch.randn(batch * m, k, dtype=dtype) / 10.0
    w = torch.randn(n, k, dtype=dtype) / 10.0
    wq, ws = og.fp8_block_quant_2d(w, group_size, group_size)      # [N,K] e4m3, [N/gs, K/gs] fp32
    wt = wq.to(DEV).transpose(0, 1)
    if is_nt:
        wt = wt.contiguous()
    out = ops.fp8_gemm_w8a16(x.to(DEV), wt, ws.t().contiguous().to(DEV), torch.Tensor())
    close(out, og.fp8_gemm_w8a16_ref(x, wq.t(), ws.t().contiguous()))
    w_deq = og.fp8_block_dequant_2d(wq, ws, group_size, group_size, dtype)
    close(out, torch.matmul(x, w_deq.t()), 6e-2, 6e-2)


@pytest.mark.parametrize("fp8", [E4, E5])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_w8a16_every_code_widens_exactly(ops, fp8, dtype):
    """All 256 FP8 codes (NaN / inf codes excluded) through the in-loop widening: with A = identity the
    GEMM returns the decoded weights themselves, which are exactly representable in bf16 / f16."""
    codes = torch.arange(256, dtype=torch.uint8)
    vals = codes.view(fp8).float()
    keep = torch.isfinite(vals)
    codes = codes[keep]
    n = codes.numel()
    k = 256
    w = torch.zeros(n, k, dtype=torch.uint8)
    w[torch.arange(n), torch.arange(n) % k] = codes       # one code per row, on the diagonal
    w[:, 255] = codes                                     # and the same code in the last K column
    a = torch.zeros(k, k, dtype=dtype)
    a[torch.arange(k), torch.arange(k)] = 1.0
    out = ops.fp8_gemm_w8a16(a.to(DEV), w.view(fp8).to(DEV).t(), None, None)   # D[kk, n] = W[n, kk]
    ref = w.view(fp8).float().t().to(dtype)
    assert torch.equal(out.float().cpu(), ref.float())    # value equality (a lone -0 accumulates to +0)


def test_w8a16_tc_matches_cuda_core_path(lib):
    """The tensor-core kernel and the CUDA-core kernel agree to output rounding on the same operands
    (C ABI called directly with force_fallback = 1 for the second run)."""
    torch.manual_seed(0)
    m, n, k = 200, 384, 1088
    x = (torch.randn(m, k, dtype=torch.bfloat16) / 10).to(DEV)
    wq = (torch.randn(n, k) / 10).to(E4).to(DEV)
    sc = (torch.rand(n) + 0.5).to(DEV)
    outs = []
    f = lib.vxk_fp8_gemm_w8a16
    f.restype = ctypes.c_int
    for force in (0, 1):
        d = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
        args = [ctypes.c_void_p(d.data_ptr()), ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(wq.data_ptr()),
                ctypes.c_void_p(sc.data_ptr()), ctypes.c_void_p(0)]
        args += [ctypes.c_int64(v) for v in (m, n, k, k, 1, k, n)]
        args += [ctypes.c_int(1)] + [ctypes.c_int64(v) for v in (0, 0, 0, 0)]
        args += [ctypes.c_int(2), ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(force),
                 ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)]
        assert f(*args) == 0, lib.vxk_last_error()
        torch.cuda.synchronize()
        outs.append(d.float().cpu())
    torch.testing.assert_close(outs[0], outs[1], atol=2e-3, rtol=8e-3)


def test_w8a16_errors(ops):
    x = torch.randn(4, 64, dtype=torch.float32, device=DEV)
    w = torch.zeros(32, 64, dtype=torch.uint8, device=DEV).view(E4).t()
    with pytest.raises(RuntimeError):
        ops.fp8_gemm_w8a16(x, w, None, None)                                   # fp32 activations
    with pytest.raises(RuntimeError):
        ops.fp8_gemm_w8a16(x.half(), w.view(torch.uint8), None, None)          # weight not fp8
    with pytest.raises(RuntimeError):
        ops.fp8_gemm_w8a16(x.half()[:, :32], w, None, None)                    # K mismatch
    out = ops.fp8_gemm_w8a16(x.half()[:0], w, None, None)
    assert out.shape == (0, 32)


@pytest.mark.parametrize("fp8", [E4, E5])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("mode", ["tensor", "channel", "block"])
@pytest.mark.parametrize("mnk", [(1, 4096, 4096), (17, 1000, 1536), (33, 384, 8192), (65, 136, 640),
                                 (128, 256, 2048), (7, 28672, 512), (129, 384, 1024), (256, 256, 512),
                                 (300, 512, 1024), (385, 256, 640), (512, 1024, 2048)])
def test_w8a16_small_m_kernel(ops, lib, fp8, dtype, mode, mnk):
    """The swapped-operand, cluster split-K kernel (M <= 128, K % 128 == 0): every K split count from 1
    to 8, partial 128-row weight tiles, bias, all three scale modes; 129 ... 512 tokens as 128-token tiles side by
    side in one launch while tiles x weight tiles fit the 148 SMs, else (up to 256 tokens) one pass per tile."""
    torch.manual_seed(7)
    m, n, k = mnk
    if mode == "block" and (n % 128 or k % 128):
        pytest.skip("shape not divisible by the 128 x 128 block")
    x = torch.randn(m, k, dtype=dtype) / 10.0
    w = torch.randn(n, k, dtype=torch.float32) / 10.0
    b = torch.randn(n, dtype=dtype) / 4
    if mode == "tensor":
        s = torch.tensor(0.02)
        wq = oq.static_scaled_fp8_quant(w, s, fp8)
        s_op = s_ref = s
    elif mode == "channel":
        wq, ws = oq.dynamic_per_token_scaled_fp8_quant(w, fp8)
        s_op, s_ref = ws.flatten(), ws.flatten()
    else:
        wq, ws = og.fp8_block_quant_2d(w, 128, 128)
        wq = wq if fp8 == E4 else wq.float().to(fp8)
        s_op = s_ref = ws.t().contiguous()
    tc0, fb0 = gemm_stats(lib)
    out = ops.fp8_gemm_w8a16(x.to(DEV), wq.to(DEV).t(), s_op.to(DEV), b.to(DEV))
    tc1, fb1 = gemm_stats(lib)
    tiles, n_tiles = (m + 127) // 128, (n + 127) // 128
    assert (tc1 - tc0, fb1 - fb0) == (1 if tiles * n_tiles <= 148 else tiles, 0)
    close(out, og.fp8_gemm_w8a16_ref(x, wq.t(), s_ref, b))


@pytest.mark.parametrize("fp8", [E4, E5])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("groups", [(128, 128), (64, 192), (256, 64), (192, 128)])      # (group_n, group_k)
@pytest.mark.parametrize("mnk", [(257, 384, 768), (700, 1536, 1152), (1024, 768, 4224)])
def test_w8a16_large_m_block_scales(ops, lib, fp8, dtype, groups, mnk):
    """M > 256 with 2-D block scales: the swapped-operand kernel with the widened weight tile in tensor memory
    folds the scale in while widening (one rounding to the activation type, like the small-M kernel).  Rectangular
    groups, group_k a multiple of 64 only, partial 128-row weight tiles and partial 256-token tiles, bias."""
    torch.manual_seed(11)
    m, n, k = mnk
    gn, gk = groups
    if n % gn or k % gk:
        pytest.skip("shape not divisible by the group")
    x = torch.randn(m, k, dtype=dtype) / 10.0
    w = torch.randn(n, k, dtype=torch.float32) / 10.0
    w[: n // 2] *= 7.0                       # scales that differ a lot between blocks
    b = torch.randn(n, dtype=dtype) / 4
    wq, ws = og.fp8_block_quant_2d(w, gn, gk)
    wq = wq if fp8 == E4 else wq.float().to(fp8)
    sc = ws.t().contiguous()                 # [K / gk, N / gn]
    tc0, fb0 = gemm_stats(lib)
    out = ops.fp8_gemm_w8a16(x.to(DEV), wq.to(DEV).t(), sc.to(DEV), b.to(DEV))
    tc1, fb1 = gemm_stats(lib)
    assert (tc1 - tc0, fb1 - fb0) == (1, 0)
    # against the exact oracle the rounding of the dequantised weights to bf16 / f16 shows: K terms of ~0.07 with a
    # relative error of up to 2^-9 each (std ~0.005 at K = 4224, and 786 k outputs to take the maximum over)
    close(out, og.fp8_gemm_w8a16_ref(x, wq.t(), sc, b), atol=6e-2, rtol=1e-2)
    # the kernel's own statement: weights dequantised with ONE rounding to the activation type
    # (tests/ops/fp8_quant_op.py:81-106), exact accumulation, one output rounding
    w_deq = og.fp8_block_dequant_2d(wq, ws, gn, gk, dtype)
    ref = (x.double() @ w_deq.double().t() + b.double()).to(dtype)
    close(out, ref, atol=2e-3, rtol=8e-3)
