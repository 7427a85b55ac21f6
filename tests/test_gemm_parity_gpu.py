"""Parity of the scaled GEMMs (torch.ops._xpu_C.fp8_gemm / fp4_gemm -> C ABI -> tcgen05
kernels) against the CPU oracle.  Shapes and operand construction follow the
reference's tests/test_fp8_gemm_onednn.py and tests/test_fp4_gemm_onednn.py.

Bar: |out - ref| <= 1e-2 + 1e-2 * |ref| against the exact-arithmetic reference
(float64 accumulate of the dequantised operands, one rounding to the output dtype);
the reference's own bar (6e-2 / 5e-2 against a matmul in the output dtype) is
checked as well.  Large shapes use the size-independent checks described in each test.
"""
import ctypes

import pytest
import torch

from oracle import formats, gemm as og, quant as oq

pytestmark = pytest.mark.gpu
E4, E5 = torch.float8_e4m3fn, torch.float8_e5m2
E8 = torch.float8_e8m0fnu
DEV = "cuda"
MNK_FACTORS = [(1, 32, 1024), (1, 32, 2048), (1, 32, 5120), (8, 512, 2048), (512, 1024, 2048), (1024, 2048, 2048)]
MNK_BLOCK_FACTORS = [(1, 256, 1024), (4, 256, 1024), (8, 256, 1024), (8, 512, 1024)]


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


# ------------------------------------------------------------------ per tensor / per channel
@pytest.mark.parametrize("out_dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("is_nt", [True, False])
@pytest.mark.parametrize("batch", [1, 2, 8])
@pytest.mark.parametrize("mnk", MNK_FACTORS)
def test_fp8_gemm_per_tensor(ops, out_dtype, is_nt, batch, mnk):
    """tests/test_fp8_gemm_onednn.py:129-171 (is_nt=False passes a contiguous [K,N] weight)."""
    torch.manual_seed(1234)
    m, n, k = mnk
    if batch > 1 and m > 8:
        pytest.skip("batched variants only on the small-M shapes (keeps the float64 CPU reference fast)")
    x = torch.randn(batch, m, k, dtype=out_dtype) / 10.0
    w = torch.randn(n, k, dtype=out_dtype) / 10.0
    s_src, s_wei = torch.tensor(4.0), torch.tensor(4.0)
    xq = oq.static_scaled_fp8_quant(x.reshape(-1, k), s_src, E4)
    wq = oq.static_scaled_fp8_quant(w, s_wei, E4)
    wt = wq.to(DEV).transpose(0, 1)
    if is_nt:
        wt = wt.contiguous()
    out = ops.fp8_gemm(xq.to(DEV), wt, out_dtype, s_src.to(DEV), s_wei.to(DEV), torch.Tensor())
    assert out.dtype == out_dtype and out.shape == (batch * m, n)
    close(out, og.fp8_gemm_ref(xq, wq.t(), out_dtype, s_src, s_wei))
    lit = og.fp8_gemm_literal(xq.to(out_dtype) * s_src.to(out_dtype), wq.to(out_dtype) * s_wei.to(out_dtype))
    close(out, lit, 6e-2, 6e-2)


@pytest.mark.parametrize("fp8", [E4, E5])
@pytest.mark.parametrize("out_dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("is_nt", [True, False])
@pytest.mark.parametrize("mnk", MNK_FACTORS + [(100, 200, 2048), (257, 129, 1040), (128, 128, 32)])
@pytest.mark.parametrize("bias", [False, True])
def test_fp8_gemm_per_channel(ops, fp8, out_dtype, is_nt, mnk, bias):
    """tests/test_fp8_gemm_onednn.py:174-214: dynamic per-token activations x per-channel weights."""
    torch.manual_seed(1234)
    m, n, k = mnk
    x = torch.randn(m, k, dtype=out_dtype) / 10.0
    w = torch.randn(n, k, dtype=out_dtype) / 10.0
    xq, xs = oq.dynamic_per_token_scaled_fp8_quant(x, fp8)
    wq, ws = oq.dynamic_per_token_scaled_fp8_quant(w, fp8)
    b = torch.randn(n, dtype=out_dtype) if bias else None
    wt = wq.to(DEV).transpose(0, 1)
    if is_nt:
        wt = wt.contiguous()
    out = ops.fp8_gemm(xq.to(DEV), wt, out_dtype, xs.to(DEV), ws.to(DEV), b.to(DEV) if bias else torch.Tensor())
    close(out, og.fp8_gemm_ref(xq, wq.t(), out_dtype, xs, ws, b))


def test_fp8_gemm_defaults_and_3d_input(ops):
    """Missing scales default to 1.0, missing out_dtype to fp16 (onednn_matmul.cpp:49,75-76); 3-D A."""
    torch.manual_seed(0)
    b, m, n, k = 2, 64, 256, 512
    xq = (torch.randn(b, m, k) / 4).to(E4)
    wq = (torch.randn(n, k) / 4).to(E4)
    out = ops.fp8_gemm(xq.to(DEV), wq.to(DEV).t(), None, None, None, None)
    assert out.dtype == torch.float16 and out.shape == (b, m, n)
    close(out.reshape(-1, n), og.fp8_gemm_ref(xq.reshape(-1, k), wq.t(), torch.float16))
    # [m, b, k]-strided activations (onednn_matmul.cpp:40-46)
    xt = xq.to(DEV).transpose(0, 1)
    with pytest.raises(RuntimeError):
        ops.fp8_gemm(xt, wq.to(DEV).t(), torch.bfloat16, None, None, None)  # rows not uniformly strided
    with pytest.raises(RuntimeError, match="f8_e5m2 or f8_e4m3fn"):
        ops.fp8_gemm(xq.to(DEV).view(torch.uint8), wq.to(DEV).t(), None, None, None, None)


# ------------------------------------------------------------------ MX / block scaled
@pytest.mark.parametrize("out_dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("mnk", MNK_FACTORS)
def test_mxfp8_gemm(ops, lib, out_dtype, mnk):
    """tests/test_fp8_gemm_onednn.py:281-308: to_mxfp both operands, B_scale = scale.t().contiguous()."""
    torch.manual_seed(0)
    m, n, k = mnk
    x = torch.randn(m, k, dtype=torch.bfloat16) / 10.0
    w = torch.randn(n, k, dtype=torch.bfloat16) / 10.0
    xs, xq = formats.mx_quantize(x, "mxfp8")
    ws, wq = formats.mx_quantize(w, "mxfp8")
    tc0, _ = gemm_stats(lib)
    out = ops.fp8_gemm(xq.to(DEV), wq.to(DEV).t(), out_dtype, xs.to(DEV), ws.t().contiguous().to(DEV), torch.Tensor())
    assert gemm_stats(lib)[0] == tc0 + 1, "MXFP8 must run on the block-scaled tcgen05 kernel"
    close(out, og.fp8_gemm_ref(xq, wq.t(), out_dtype, xs, ws.t().contiguous()))
    lit = torch.matmul(formats.mx_dequantize(xq, xs).to(out_dtype), formats.mx_dequantize(wq, ws).to(out_dtype).t())
    close(out, lit, 5e-2, 5e-2)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("is_nt", [True, False])
@pytest.mark.parametrize("batch", [1, 8])
@pytest.mark.parametrize("group", [128, 256])
@pytest.mark.parametrize("ue8m0", [False, True])
@pytest.mark.parametrize("mnk", MNK_BLOCK_FACTORS)
def test_fp8_gemm_per_block(ops, dtype, is_nt, batch, group, ue8m0, mnk):
    """tests/test_fp8_gemm_onednn.py:318-358: 1 x G activation groups (our own quant op) x G x G weight blocks."""
    torch.manual_seed(1234)
    m, n, k = mnk
    x = (torch.randn(batch * m, k, dtype=dtype) / 10.0).to(DEV)
    w = torch.randn(n, k, dtype=dtype) / 10.0
    xq, xs = ops.per_token_group_quant_fp8(x, group, dtype=E4, use_ue8m0=ue8m0)
    wq, ws = og.fp8_block_quant_2d(w, group, group, use_ue8m0=ue8m0)
    wt = wq.to(DEV).transpose(0, 1)
    if is_nt:
        wt = wt.contiguous()
    bs = ws.t().contiguous()
    out = ops.fp8_gemm(xq, wt, dtype, xs, bs.to(DEV), torch.Tensor())
    close(out, og.fp8_gemm_ref(xq.cpu(), wq.t(), dtype, xs.cpu(), bs))


# ------------------------------------------------------------------ MXFP4
@pytest.mark.parametrize("out_dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("mnk", MNK_FACTORS + [(300, 384, 1536), (128, 128, 64), (5, 40, 96)])
def test_mxfp4_gemm(ops, out_dtype, mnk):
    """tests/test_fp4_gemm_onednn.py:49-78: B_scale passed as [N, K/32]."""
    torch.manual_seed(0)
    m, n, k = mnk
    x = torch.randn(m, k, dtype=torch.bfloat16) * 0.01
    w = torch.randn(n, k, dtype=torch.bfloat16) * 0.01
    xs, xq = formats.mx_quantize(x, "mxfp4")
    ws, wq = formats.mx_quantize(w, "mxfp4")
    out = ops.fp4_gemm(xq.to(DEV), wq.to(DEV).transpose(0, 1), xs.to(DEV), ws.to(DEV), out_dtype)
    assert out.dtype == out_dtype and out.shape == (m, n)
    ref = og.fp4_gemm_ref(xq, wq.t(), xs, ws, out_dtype)
    close(out, ref, 1e-2, 1e-2)
    lit = torch.matmul(formats.mx_dequantize(xq, xs).to(out_dtype), formats.mx_dequantize(wq, ws).to(out_dtype).t())
    close(out, lit, 5e-2, 5e-2)


def test_mxfp4_gemm_errors(ops):
    a = torch.zeros(4, 32, dtype=torch.uint8, device=DEV).view(torch.float4_e2m1fn_x2)
    w = torch.zeros(8, 32, dtype=torch.uint8, device=DEV).view(torch.float4_e2m1fn_x2)
    sa = torch.zeros(4, 2, dtype=torch.uint8, device=DEV).view(E8)
    sw = torch.zeros(8, 2, dtype=torch.uint8, device=DEV).view(E8)
    with pytest.raises(RuntimeError, match="Unsupported scale type"):
        ops.fp4_gemm(a, w.t(), sa.view(torch.uint8).float(), sw, torch.bfloat16)
    with pytest.raises(RuntimeError, match="K-major"):
        ops.fp4_gemm(a, w.t().contiguous(), sa, sw, torch.bfloat16)
    out = ops.fp4_gemm(a, w.t(), sa, sw, torch.bfloat16)
    assert out.shape == (4, 8) and float(out.abs().max()) == 0.0


# ------------------------------------------------------------------ tensor-core path == fallback path
@pytest.mark.parametrize("kind", ["per_channel", "mxfp8", "mxfp4"])
def test_tensor_core_path_matches_cuda_core_fallback(ops, lib, kind):
    """The same call through the tcgen05 kernel and through the CUDA-core fallback (two independent
    implementations of the C-ABI contract) agree to bf16 rounding; each path is proven taken."""
    torch.manual_seed(7)
    m, n, k = 384, 640, 2048
    x = torch.randn(m, k, dtype=torch.bfloat16) / 10.0
    w = torch.randn(n, k, dtype=torch.bfloat16) / 10.0

    def run():
        if kind == "per_channel":
            xq, xs = oq.dynamic_per_token_scaled_fp8_quant(x, E4)
            wq, ws = oq.dynamic_per_token_scaled_fp8_quant(w, E4)
            return ops.fp8_gemm(xq.to(DEV), wq.to(DEV).t(), torch.bfloat16, xs.to(DEV), ws.to(DEV), None)
        if kind == "mxfp8":
            xs, xq = formats.mx_quantize(x, "mxfp8")
            ws, wq = formats.mx_quantize(w, "mxfp8")
            return ops.fp8_gemm(xq.to(DEV), wq.to(DEV).t(), torch.bfloat16, xs.to(DEV), ws.t().contiguous().to(DEV), None)
        xs, xq = formats.mx_quantize(x * 0.1, "mxfp4")
        ws, wq = formats.mx_quantize(w * 0.1, "mxfp4")
        return ops.fp4_gemm(xq.to(DEV), wq.to(DEV).t(), xs.to(DEV), ws.to(DEV), torch.bfloat16)

    tc0, fb0 = gemm_stats(lib)
    fast = run()
    assert gemm_stats(lib) == (tc0 + 1, fb0)
    prev = lib.vxk_set_gemm_force_fallback(1)
    try:
        slow = run()
    finally:
        lib.vxk_set_gemm_force_fallback(prev)
    assert gemm_stats(lib) == (tc0 + 1, fb0 + 1)
    torch.testing.assert_close(fast.float(), slow.float(), atol=2e-3, rtol=8e-3)


# ------------------------------------------------------------------ BASELINE shapes (size-independent checks)
LLAMA8B = [(4096, 6144), (4096, 4096), (4096, 28672), (14336, 4096)]        # (K, N), config 3
LLAMA70B = [(8192, 10240), (8192, 8192), (8192, 57344), (28672, 8192)]      # (K, N), config 4
DEEPSEEK = [(7168, 1536), (1536, 24576), (512, 32768), (16384, 7168), (7168, 36864), (18432, 7168)]  # config 5


@pytest.mark.parametrize("fp8", [E4, E5])
@pytest.mark.parametrize("per_channel", [True, False])
@pytest.mark.parametrize("bias", [False, True])
@pytest.mark.parametrize("nk", [(384, 512), (1000, 1040), (4096, 4096), (136, 7168), (6144, 640)])
@pytest.mark.parametrize("m", [3, 33, 64, 65, 128, 129, 200, 256, 257, 300, 384, 385, 511, 512])
def test_fp8_gemm_skinny(ops, lib, fp8, per_channel, bias, nk, m):
    """Small-M kernel (gemm_skinny.cuh: operands swapped, K split over a cluster, DSMEM reduction): token
    tiles 32 / 64 / 128 / 256, two to four token tiles of 128 side by side (129 ... 512 tokens, partial last tile),
    N and K tails, odd numbers of K slices per cluster rank, every scale mode.
    Same checker as tests/test_fp8_gemm_onednn.py:174-214 (dequantised matmul), atol = rtol = 1e-2."""
    torch.manual_seed(m * 7 + nk[0])
    n, k = nk
    out_dtype = torch.bfloat16 if (m + n) % 2 else torch.float16
    x = torch.randn(m, k, dtype=torch.float32) / 10.0
    w = torch.randn(n, k, dtype=torch.float32) / 10.0
    if per_channel:
        xq, xs = oq.dynamic_per_token_scaled_fp8_quant(x, fp8)
        wq, ws = oq.dynamic_per_token_scaled_fp8_quant(w, fp8)
    else:
        xs = x.abs().max().reshape(1) / torch.finfo(fp8).max
        ws = w.abs().max().reshape(1) / torch.finfo(fp8).max
        xq = (x / xs).clamp(-torch.finfo(fp8).max, torch.finfo(fp8).max).to(fp8)
        wq = (w / ws).clamp(-torch.finfo(fp8).max, torch.finfo(fp8).max).to(fp8)
    b = torch.randn(n, dtype=out_dtype) if bias else None
    tc0, fb0 = gemm_stats(lib)
    out = ops.fp8_gemm(xq.to(DEV), wq.to(DEV).t(), out_dtype, xs.to(DEV), ws.to(DEV),
                       b.to(DEV) if bias else torch.Tensor())
    assert gemm_stats(lib) == (tc0 + 1, fb0)
    close(out, og.fp8_gemm_ref(xq, wq.t(), out_dtype, xs, ws, b))
    # deterministic: the cluster reduction sums in rank order
    out2 = ops.fp8_gemm(xq.to(DEV), wq.to(DEV).t(), out_dtype, xs.to(DEV), ws.to(DEV),
                        b.to(DEV) if bias else torch.Tensor())
    assert torch.equal(out, out2)


def _gpu_ref(a_deq, w_deq, out_dtype):
    """fp32 matmul of dequantised operands on the GPU (TF32 off) as the full-size checker."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        return torch.matmul(a_deq.float(), w_deq.float().t()).to(out_dtype)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


@pytest.mark.parametrize("kn", LLAMA8B)
@pytest.mark.parametrize("m", [1, 16, 1000, 8192])
def test_llama8b_w8a8_shapes(ops, lib, kn, m):
    """Config 3: per-token x per-channel FP8 at the Llama-3-8B shapes; the activation comes from our own
    dynamic per-token quant op (the real hot path), the checker is an fp32 GPU matmul of the dequantised operands."""
    k, n = kn
    torch.manual_seed(1)
    x = (torch.randn(m, k, dtype=torch.bfloat16, device=DEV) / 10.0)
    w = (torch.randn(n, k, dtype=torch.bfloat16, device=DEV) / 10.0)
    xq, xs = ops.scaled_fp8_quant(x, use_per_token_if_dynamic=True, fp8_dtype=E4)
    wq, ws = ops.scaled_fp8_quant(w, use_per_token_if_dynamic=True, fp8_dtype=E4)
    tc0, _ = gemm_stats(lib)
    out = ops.fp8_gemm(xq, wq.t(), torch.bfloat16, xs, ws, None)
    assert gemm_stats(lib)[0] == tc0 + 1
    ref = _gpu_ref(xq.float() * xs, wq.float() * ws, torch.bfloat16)
    torch.testing.assert_close(out.float(), ref.float(), atol=1e-2, rtol=1e-2)


@pytest.mark.parametrize("kn", LLAMA70B)
@pytest.mark.parametrize("m", [1, 100, 2048])
def test_llama70b_mxfp4_shapes(ops, lib, kn, m):
    """Config 4: MXFP4 GEMM fed by silu_and_mul_mxfp4_quant (down_proj) / per_token_group_quant_mxfp4."""
    k, n = kn
    torch.manual_seed(2)
    if k == 28672:   # down_proj: activation = silu(gate) * up, quantised by the fused op
        gu = torch.randn(m, 2 * k, dtype=torch.bfloat16, device=DEV)
        q4, s4 = ops.silu_and_mul_mxfp4_quant(gu)
        xq, xs = q4.view(torch.float4_e2m1fn_x2), s4.to(E8)
    else:
        x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
        xq, xs = ops.per_token_group_quant_mxfp4(x)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV) * 0.05
    wq, ws = ops.per_token_group_quant_mxfp4(w)
    tc0, _ = gemm_stats(lib)
    out = ops.fp4_gemm(xq, wq.t(), xs, ws, torch.bfloat16)
    assert gemm_stats(lib)[0] == tc0 + 1
    lut = torch.tensor(list(formats.E2M1_VALUES) + [-v for v in formats.E2M1_VALUES], device=DEV)

    def deq(q, s):
        b = q.view(torch.uint8)
        vals = torch.stack((lut[(b & 0xF).long()], lut[(b >> 4).long()]), dim=-1).reshape(b.shape[0], -1)
        return vals * s.float().repeat_interleave(32, dim=1)

    ref = _gpu_ref(deq(xq, xs), deq(wq, ws), torch.bfloat16)
    torch.testing.assert_close(out.float(), ref.float(), atol=1e-2 * float(ref.float().abs().max()), rtol=1e-2)


@pytest.mark.parametrize("kn", DEEPSEEK)
@pytest.mark.parametrize("m", [4, 512])
@pytest.mark.parametrize("ue8m0", [True, False])
def test_deepseek_block_fp8_shapes(ops, kn, m, ue8m0):
    """Config 5: 1x128 activation groups x 128x128 weight blocks (fp32 or UE8M0 scales)."""
    k, n = kn
    torch.manual_seed(3)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV) / 10.0
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV) / 10.0
    xq, xs = ops.per_token_group_quant_fp8(x, 128, dtype=E4, use_ue8m0=ue8m0)
    wb = w.float().reshape(n // 128, 128, k // 128, 128)
    amax = wb.abs().amax(dim=(1, 3)).clamp(min=1e-6)
    s = amax / 448.0
    if ue8m0:
        s = torch.exp2(torch.ceil(torch.log2(s)))
    wq = (wb / s[:, None, :, None]).clamp(-448, 448).to(E4).reshape(n, k)
    ws = s.to(E8) if ue8m0 else s
    out = ops.fp8_gemm(xq, wq.t(), torch.bfloat16, xs, ws.t().contiguous(), None)
    a_deq = xq.float() * xs.float().repeat_interleave(128, dim=1)
    w_deq = wq.float() * ws.float().repeat_interleave(128, dim=0).repeat_interleave(128, dim=1)
    ref = _gpu_ref(a_deq, w_deq, torch.bfloat16)
    torch.testing.assert_close(out.float(), ref.float(), atol=1e-2 * float(ref.float().abs().max()), rtol=1e-2)


def test_gemm_linearity_full_size(ops):
    """Size-independent property at BASELINE's largest shape (M=8192, K=8192, N=57344 is 7.7 TFLOP):
    scaling every per-channel weight scale by 2 doubles the output exactly (power-of-two scaling is exact)."""
    torch.manual_seed(4)
    m, k, n = 8192, 8192, 57344
    xq = (torch.randn(m, k, device=DEV) / 4).to(E4)
    wq = (torch.randn(n, k, device=DEV) / 4).to(E4)
    xs = torch.rand(m, 1, device=DEV) * 0.01 + 0.001
    ws = torch.rand(n, 1, device=DEV) * 0.01 + 0.001
    out1 = ops.fp8_gemm(xq, wq.t(), torch.bfloat16, xs, ws, None)
    out2 = ops.fp8_gemm(xq, wq.t(), torch.bfloat16, xs, ws * 2.0, None)
    assert torch.equal(out2, out1 * 2.0)
    rows = torch.tensor([0, 1, 4095, 8191], device=DEV)
    ref = _gpu_ref(xq[rows].float() * xs[rows], wq.float() * ws, torch.bfloat16)
    torch.testing.assert_close(out1[rows].float(), ref.float(), atol=1e-2 * float(ref.float().abs().max()), rtol=1e-2)


# ------------------------------------------------------------------ round-2 additions
def _mx4_deq(q, s):
    lut = torch.tensor(list(formats.E2M1_VALUES) + [-v for v in formats.E2M1_VALUES], device=q.device)
    b = q.view(torch.uint8)
    vals = torch.stack((lut[(b & 0xF).long()], lut[(b >> 4).long()]), dim=-1).reshape(b.shape[0], -1)
    return vals * s.float().repeat_interleave(32, dim=1)


@pytest.mark.parametrize("k", [11008, 2880, 96, 544])
@pytest.mark.parametrize("m", [1, 64, 300])
def test_mxfp4_gemm_unaligned_scale_pitch(ops, lib, k, m):
    """K/32 not a multiple of 16 bytes (Llama-2 MLP K = 11008, K = 2880): the scale rows are not
    TMA-expressible as they stand, the call must still run on the tensor cores."""
    torch.manual_seed(5)
    n = 384
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV) * 0.05
    xq, xs = ops.per_token_group_quant_mxfp4(x)
    wq, ws = ops.per_token_group_quant_mxfp4(w)
    tc0, fb0 = gemm_stats(lib)
    out = ops.fp4_gemm(xq, wq.t(), xs, ws, torch.bfloat16)
    tc1, fb1 = gemm_stats(lib)
    assert (tc1, fb1) == (tc0 + 1, fb0), "must not fall back to the CUDA-core kernel"
    ref = _gpu_ref(_mx4_deq(xq, xs), _mx4_deq(wq, ws), torch.bfloat16)
    torch.testing.assert_close(out.float(), ref.float(), atol=1e-2 * float(ref.float().abs().max()), rtol=1e-2)


def _rows_sample(m):
    return torch.tensor(sorted({0, 1, 127, 128, 255, 256, m // 2 - 1, m // 2, m - 2, m - 1}), device=DEV)


@pytest.mark.parametrize("kn", LLAMA70B)
def test_llama70b_mxfp4_full_m(ops, lib, kn):
    """Config 4 at M = 8192 on a row sample against an fp32 matmul of the dequantised operands
    (the rows cover both CTAs of a pair, tile edges and the last tile)."""
    k, n = kn
    m = 8192
    torch.manual_seed(6)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV) * 0.05
    xq, xs = ops.per_token_group_quant_mxfp4(x)
    wq, ws = ops.per_token_group_quant_mxfp4(w)
    del x, w
    out = ops.fp4_gemm(xq, wq.t(), xs, ws, torch.bfloat16)
    rows = _rows_sample(m)
    ref = _gpu_ref(_mx4_deq(xq.view(torch.uint8)[rows], xs.view(torch.uint8)[rows].view(E8)), _mx4_deq(wq, ws),
                   torch.bfloat16)
    torch.testing.assert_close(out[rows].float(), ref.float(), atol=1e-2 * float(ref.float().abs().max()), rtol=1e-2)
    # linearity: doubling every weight scale (an exponent-byte increment) doubles the output exactly
    ws2 = (ws.view(torch.uint8) + 1).view(E8)
    out2 = ops.fp4_gemm(xq, wq.t(), xs, ws2, torch.bfloat16)
    assert torch.equal(out2, out * 2.0)


@pytest.mark.parametrize("kn", DEEPSEEK + [(7168, 576)])
@pytest.mark.parametrize("ue8m0", [True, False])
def test_deepseek_block_fp8_full_m(ops, kn, ue8m0):
    """Config 5 at M = 8192 on a row sample; (7168, 576) is the N % 128 != 0 edge the reference's integer
    n / sB.size(1) cannot express with 128-wide blocks (fp8_gemm_w8a8.h:40) -- run with gn = 64."""
    k, n = kn
    gn = 128 if n % 128 == 0 else 64
    m = 8192
    torch.manual_seed(7)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV) / 10.0
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV) / 10.0
    xq, xs = ops.per_token_group_quant_fp8(x, 128, dtype=E4, use_ue8m0=ue8m0)
    wb = w.float().reshape(n // gn, gn, k // 128, 128)
    s = wb.abs().amax(dim=(1, 3)).clamp(min=1e-6) / 448.0
    if ue8m0:
        s = torch.exp2(torch.ceil(torch.log2(s)))
    wq = (wb / s[:, None, :, None]).clamp(-448, 448).to(E4).reshape(n, k)
    ws = s.to(E8) if ue8m0 else s
    del x, w, wb
    out = ops.fp8_gemm(xq, wq.t(), torch.bfloat16, xs, ws.t().contiguous(), None)
    rows = _rows_sample(m)
    a_deq = xq[rows].float() * xs[rows].float().repeat_interleave(128, dim=1)
    w_deq = wq.float() * ws.float().repeat_interleave(gn, dim=0).repeat_interleave(128, dim=1)
    ref = _gpu_ref(a_deq, w_deq, torch.bfloat16)
    torch.testing.assert_close(out[rows].float(), ref.float(), atol=1e-2 * float(ref.float().abs().max()), rtol=1e-2)


@pytest.mark.parametrize("kind", ["fp8", "mxfp4", "block_fp8"])
def test_gemm_m_sweep(ops, kind):
    """M = 1, 2, 4, ... 4096 on one shape per kind: every kernel-selection boundary is crossed."""
    torch.manual_seed(8)
    k, n = 4096, 6144
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV) / 10.0
    xall = torch.randn(4096, k, dtype=torch.bfloat16, device=DEV) / 10.0
    if kind == "fp8":
        wq, ws = ops.scaled_fp8_quant(w, use_per_token_if_dynamic=True, fp8_dtype=E4)
        w_deq = wq.float() * ws
    elif kind == "mxfp4":
        wq, ws = ops.per_token_group_quant_mxfp4(w)
        w_deq = _mx4_deq(wq, ws)
    else:
        wb = w.float().reshape(n // 128, 128, k // 128, 128)
        s = wb.abs().amax(dim=(1, 3)).clamp(min=1e-6) / 448.0
        wq = (wb / s[:, None, :, None]).clamp(-448, 448).to(E4).reshape(n, k)
        ws = s
        w_deq = wq.float() * s.repeat_interleave(128, dim=0).repeat_interleave(128, dim=1)
    for e in range(13):
        m = 1 << e
        x = xall[:m]
        if kind == "fp8":
            xq, xs = ops.scaled_fp8_quant(x, use_per_token_if_dynamic=True, fp8_dtype=E4)
            out = ops.fp8_gemm(xq, wq.t(), torch.bfloat16, xs, ws, None)
            a_deq = xq.float() * xs
        elif kind == "mxfp4":
            xq, xs = ops.per_token_group_quant_mxfp4(x)
            out = ops.fp4_gemm(xq, wq.t(), xs, ws, torch.bfloat16)
            a_deq = _mx4_deq(xq, xs)
        else:
            xq, xs = ops.per_token_group_quant_fp8(x, 128, dtype=E4, use_ue8m0=False)
            out = ops.fp8_gemm(xq, wq.t(), torch.bfloat16, xs, ws.t().contiguous(), None)
            a_deq = xq.float() * xs.repeat_interleave(128, dim=1)
        ref = _gpu_ref(a_deq, w_deq, torch.bfloat16)
        torch.testing.assert_close(out.float(), ref.float(), atol=1e-2 * float(ref.float().abs().max()), rtol=1e-2,
                                   msg=lambda s, m=m: f"{kind} M={m}: {s}")


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two CUDA devices in one process")
def test_ops_on_second_device_after_first(ops):
    """One process, cuda:0 then cuda:1: per-device kernel attributes (dynamic shared memory above 48 KB)
    must be set on both."""
    for dev in ("cuda:0", "cuda:1"):
        torch.manual_seed(9)
        x = torch.randn(300, 4096, dtype=torch.bfloat16, device=dev) / 10.0
        w = torch.randn(512, 4096, dtype=torch.bfloat16, device=dev) / 10.0
        xq, xs = ops.scaled_fp8_quant(x, use_per_token_if_dynamic=True, fp8_dtype=E4)
        wq, ws = ops.scaled_fp8_quant(w, use_per_token_if_dynamic=True, fp8_dtype=E4)
        out = ops.fp8_gemm(xq, wq.t(), torch.bfloat16, xs, ws, None)
        ref = _gpu_ref(xq.float() * xs, wq.float() * ws, torch.bfloat16)
        torch.testing.assert_close(out.float(), ref.float(), atol=1e-2, rtol=1e-2)
        x4, s4 = ops.per_token_group_quant_mxfp4(x)
        w4, t4 = ops.per_token_group_quant_mxfp4(w)
        out4 = ops.fp4_gemm(x4, w4.t(), s4, t4, torch.bfloat16)
        ref4 = _gpu_ref(_mx4_deq(x4, s4), _mx4_deq(w4, t4), torch.bfloat16)
        torch.testing.assert_close(out4.float(), ref4.float(), atol=1e-2 * float(ref4.float().abs().max()), rtol=1e-2)
        # rms weight staged in > 48 KB of dynamic shared memory (H = 14336)
        xr = torch.randn(64, 14336, dtype=torch.bfloat16, device=dev) / 14336
        wr = torch.empty(14336, dtype=torch.bfloat16, device=dev).normal_(1.0, 0.1)
        q, s = ops.rms_norm_dynamic_per_token_quant(xr, wr, 1e-6, E4)
        rq, rs, _ = oq.rms_norm_dynamic_per_token_quant(xr.cpu(), wr.cpu(), 1e-6, E4, exact_variance=True)
        assert torch.equal(s.cpu(), rs)


@pytest.mark.parametrize("k", [1024, 1280, 4096, 8192])
@pytest.mark.parametrize("n", [128, 1000, 4096])
@pytest.mark.parametrize("m", [1, 7, 33, 64, 100, 128, 129, 200, 256, 257, 385, 512])
def test_mxfp4_skinny_decode_kernel(ops, lib, m, n, k):
    """M <= 128, and 129 ... 512 tokens as two to four 128-token tiles side by side: the swap-AB split-K MXFP4 kernel
    (gemm_skinny_mx4.cuh) straight from the natural scale tensors;
    ragged N, K slices that do not fill the last ring stage (1280 = 5 x 256), bias, both output types."""
    torch.manual_seed(m * 131 + n + k)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV) * 0.05
    xq, xs = ops.per_token_group_quant_mxfp4(x)
    wq, ws = ops.per_token_group_quant_mxfp4(w)
    bias = torch.randn(n, dtype=torch.bfloat16, device=DEV)
    tc0, fb0 = gemm_stats(lib)
    out = ops.fp4_gemm(xq, wq.t(), xs, ws, torch.bfloat16, bias)
    out16 = ops.fp4_gemm(xq, wq.t(), xs, ws, torch.float16)
    tc1, fb1 = gemm_stats(lib)
    assert (tc1, fb1) == (tc0 + 2, fb0)
    ref = _mx4_deq(xq, xs).double() @ _mx4_deq(wq, ws).double().t()
    tol = 1e-2 * float(ref.abs().max())
    torch.testing.assert_close(out.double(), (ref + bias.double()).to(torch.bfloat16).double(), atol=tol, rtol=1e-2)
    torch.testing.assert_close(out16.double(), ref.to(torch.float16).double(), atol=tol, rtol=1e-2)


def test_weight_written_right_before_small_m_gemm(ops, lib):
    """the small-M kernels request weight tiles before griddepcontrol.wait, so a weight written by the
    kernel just before them on the same stream must be fenced off by the caller -- or the early loads switched off
    (vxk_set_early_weight_loads(0), include/vxk_b200.h).  With the switch off, 200 back-to-back
    (rewrite weight -> GEMM) pairs on one stream must all see the new weight."""
    torch.manual_seed(3)
    m, n, k = 8, 1024, 4096
    x = (torch.randn(m, k, device=DEV) / 4).to(E4)
    xs = torch.ones(m, 1, device=DEV)
    ws = torch.ones(n, 1, device=DEV)
    wa = (torch.randn(n, k, device=DEV) / 4).to(E4)
    wb = (torch.randn(n, k, device=DEV) / 4).to(E4)
    ref = [ops.fp8_gemm(x, w.t(), torch.bfloat16, xs, ws, None).clone() for w in (wa, wb)]
    torch.cuda.synchronize()
    assert not torch.equal(ref[0], ref[1])
    w = torch.empty_like(wa)
    prev = lib.vxk_set_early_weight_loads(0)
    try:
        outs = []
        for i in range(200):
            w.copy_(wb if i % 2 else wa)                         # the producer of the weight, same stream
            outs.append(ops.fp8_gemm(x, w.t(), torch.bfloat16, xs, ws, None))
        torch.cuda.synchronize()
        for i, o in enumerate(outs):
            assert torch.equal(o, ref[i % 2]), i
    finally:
        lib.vxk_set_early_weight_loads(prev)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("m", [1, 3, 17, 33, 64, 65, 128, 130, 200, 256])
@pytest.mark.parametrize("nk,gk,gn", [((1536, 7168), 128, 128), ((7168, 2048), 128, 128), ((576, 1024), 128, 64),
                                      ((1000, 1536), 256, 8), ((384, 1024), 512, 128), ((24576, 1536), 128, 128)])
@pytest.mark.parametrize("bias,col_major_sa", [(False, False), (True, True)])
def test_block_fp8_small_m_kernel(ops, lib, dtype, m, nk, gk, gn, bias, col_major_sa):
    """fp32 block scales at M <= 64 (BASELINE config 5 in the decode regime), and 65 ... 256 tokens as 64-token tiles
    side by side where they fit one wave (else the general kernels): gemm_skinny_kernel<MT, BLK> -- swap-AB,
    cluster split-K, per-slice promotion.  DeepSeek shapes, N tails (576, 1000), group_k 128 / 256 / 512, group_n
    8 ... 128, row- and column-major activation scales, bias; float64 oracle, atol = rtol = 1e-2; deterministic."""
    n, k = nk
    g = torch.Generator().manual_seed(m * 31 + n)
    xq = (torch.randn(m, k, generator=g) / 4).to(E4)
    wq = (torch.randn(n, k, generator=g) / 4).to(E4)
    xs = torch.rand(m, k // gk, generator=g) * 0.02 + 0.005
    ws = torch.rand(k // gk, n // gn, generator=g) * 0.02 + 0.005
    b = torch.randn(n, generator=g).to(dtype) if bias else None
    xs_d = xs.to(DEV)
    if col_major_sa:
        xs_d = xs.t().contiguous().to(DEV).t()          # logical [m, k/gk], column-major storage
    tc0, fb0 = gemm_stats(lib)
    out = ops.fp8_gemm(xq.to(DEV), wq.to(DEV).t(), dtype, xs_d, ws.to(DEV), b.to(DEV) if bias else torch.Tensor())
    assert gemm_stats(lib) == (tc0 + 1, fb0), "tensor-core path"
    close(out, og.fp8_gemm_ref(xq, wq.t(), dtype, xs, ws, b))
    out2 = ops.fp8_gemm(xq.to(DEV), wq.to(DEV).t(), dtype, xs_d, ws.to(DEV), b.to(DEV) if bias else torch.Tensor())
    assert torch.equal(out, out2)
