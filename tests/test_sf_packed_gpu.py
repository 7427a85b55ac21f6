"""SURVEY 8f #2, first half: MX scale factors leave the quantisation kernels (and a one-off weight pre-pack)
already in the layout tcgen05.cp consumes, so the MXFP4 GEMM issues no re-packing launch.

The layout is this repository's own (DESIGN.md "Scale-factor layouts"; the reference re-orders inside oneDNN,
fp4_gemm_w4a4.h:159-167), so the checker is a numpy restatement of the layout definition, and the functional
bar is: the packed path produces the SAME bits as the reference-API path.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
E8 = torch.float8_e8m0fnu


def pack_ref(sc: np.ndarray, nblk: int, sf_words: int) -> np.ndarray:
    """sc [rows, K/32] uint8 -> bytes in chunk order [kb][blk][w][lane][c][b] (include/vxk_b200.h)."""
    rows, groups = sc.shape
    gp = 4 * sf_words
    num_kb = (groups + gp - 1) // gp
    full = np.zeros((nblk * 128, num_kb * gp), dtype=np.uint8)
    full[:rows, :groups] = sc
    # row = blk*128 + c*32 + lane ; group = kb*gp + w*4 + b
    v = full.reshape(nblk, 4, 32, num_kb, sf_words, 4)          # [blk, c, lane, kb, w, b]
    return np.ascontiguousarray(v.transpose(3, 0, 4, 2, 1, 5)).reshape(-1)


@pytest.fixture(scope="module")
def ops():
    from vllm_xpu_kernels_b200 import ops as _ops
    return _ops


@pytest.fixture(scope="module")
def lib():
    from vllm_xpu_kernels_b200 import _loader
    lib = _loader.cdll()
    lib.vxk_sf_pack_launches.restype = ctypes.c_uint64
    return lib


@pytest.mark.parametrize("rows,K", [(1, 32), (127, 256), (128, 512), (300, 2880), (1000, 4096), (8192, 8192)])
@pytest.mark.parametrize("is_weight", [False, True])
def test_mx_sf_pack_layout(ops, rows, K, is_weight):
    torch.manual_seed(0)
    sc = torch.randint(100, 140, (rows, K // 32), dtype=torch.uint8)
    got = ops.mx_sf_pack(sc.to(DEV).view(E8), True, is_weight).cpu().numpy()
    nblk = (rows + 127) // 128 + (2 if is_weight else 1)
    ref = pack_ref(sc.numpy(), nblk, 2)
    assert got.shape == ref.shape
    assert np.array_equal(got, ref)
    # strided source (column-major scales, the reference's column_major_scales=True convention)
    sct = sc.to(DEV).t().contiguous().t()
    assert np.array_equal(ops.mx_sf_pack(sct.view(E8), True, is_weight).cpu().numpy(), ref)


@pytest.mark.parametrize("T,H", [(1, 64), (5, 256), (300, 2880), (1024, 4096), (333, 14336)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_quant_kernels_emit_packed_scales(ops, T, H, dtype):
    """a5 / a8 / a12 *_sf: same codes and scales as the plain op, plus the packed bytes = pack(scales)."""
    torch.manual_seed(1)
    x = torch.randn(T, H, dtype=dtype, device=DEV)
    nblk = (T + 127) // 128 + 1

    def check(sf, s_e8):
        ref = pack_ref(s_e8.view(torch.uint8).cpu().numpy(), nblk, 2)
        got = sf.cpu().numpy()
        if H % 256 == 0:
            # only real rows are written: compare those (padding rows of the last block are don't-care)
            mask = pack_ref(np.full((T, H // 32), 1, dtype=np.uint8), nblk, 2).astype(bool)
            assert np.array_equal(got[mask], ref[mask])
        else:
            assert np.array_equal(got, ref)     # zero-filled buffer: identical everywhere

    q0, s0 = ops.per_token_group_quant_mxfp4(x)
    q1, s1, sf = ops.per_token_group_quant_mxfp4_sf(x)
    assert torch.equal(q0.view(torch.uint8), q1.view(torch.uint8)) and torch.equal(s0.view(torch.uint8), s1.view(torch.uint8))
    check(sf, s1)

    w = torch.empty(H, dtype=dtype, device=DEV).normal_(1.0, 0.1)
    q0, s0 = ops.rms_norm_mxfp4_quant(x, w, 1e-6)
    q1, s1, sf = ops.rms_norm_mxfp4_quant_sf(x, w, 1e-6)
    assert torch.equal(q0, q1) and torch.equal(s0, s1)
    check(sf, s1.to(E8))
    res0 = torch.randn_like(x)
    res1 = res0.clone()
    q0, s0 = ops.rms_norm_mxfp4_quant(x, w, 1e-6, res0)
    q1, s1, sf = ops.rms_norm_mxfp4_quant_sf(x, w, 1e-6, res1)
    assert torch.equal(q0, q1) and torch.equal(s0, s1) and torch.equal(res0, res1)
    check(sf, s1.to(E8))

    if H % 64 == 0:
        g = torch.randn(T, H, dtype=dtype, device=DEV)        # gate || up, hidden = H / 2
        q0, s0 = ops.silu_and_mul_mxfp4_quant(g)
        q1, s1, sf = ops.silu_and_mul_mxfp4_quant_sf(g)
        assert torch.equal(q0, q1) and torch.equal(s0, s1)
        nb = (T + 127) // 128 + 1
        ref = pack_ref(s1.to(E8).view(torch.uint8).cpu().numpy(), nb, 2)
        got = sf.cpu().numpy()
        mask = pack_ref(np.full((T, H // 64), 1, dtype=np.uint8), nb, 2).astype(bool)
        assert np.array_equal(got[mask], ref[mask])


@pytest.mark.parametrize("m,n,k", [(129, 384, 512), (300, 1024, 2880), (1024, 2048, 4096), (2048, 8192, 8192)])
def test_fp4_gemm_with_packed_scales_is_bit_identical_and_packs_nothing(ops, lib, m, n, k):
    torch.manual_seed(2)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV) * 0.05
    xq, xs, xsf = ops.per_token_group_quant_mxfp4_sf(x)
    wq, ws = ops.per_token_group_quant_mxfp4(w)
    wsf = ops.mx_sf_pack(ws, True, True)                        # once per weight
    ref = ops.fp4_gemm(xq, wq.t(), xs, ws, torch.bfloat16)
    torch.cuda.synchronize()
    n0 = lib.vxk_sf_pack_launches()
    out = ops.fp4_gemm_sf(xq, wq.t(), xs, ws, torch.bfloat16, None, xsf, wsf)
    torch.cuda.synchronize()
    assert lib.vxk_sf_pack_launches() == n0, "the packed path must not launch the re-packing kernel"
    assert torch.equal(out, ref)
    # one side only
    out_b = ops.fp4_gemm_sf(xq, wq.t(), xs, ws, torch.bfloat16, None, None, wsf)
    torch.cuda.synchronize()
    # (small problems run on the single-CTA kernel, which reads the natural layout and never re-packs)
    d = lib.vxk_sf_pack_launches() - n0
    assert d <= 1 and (d == 1 or m < 2048)
    assert torch.equal(out_b, ref)
    bias = torch.randn(n, dtype=torch.bfloat16, device=DEV)
    assert torch.equal(ops.fp4_gemm_sf(xq, wq.t(), xs, ws, torch.bfloat16, bias, xsf, wsf),
                       ops.fp4_gemm(xq, wq.t(), xs, ws, torch.bfloat16, bias))


def test_small_m_ignores_packed_scales(ops):
    """M <= 128 runs on the single-CTA kernel, which reads the natural-layout scales: packed copies are optional extras."""
    torch.manual_seed(3)
    x = torch.randn(16, 1024, dtype=torch.bfloat16, device=DEV)
    w = torch.randn(512, 1024, dtype=torch.bfloat16, device=DEV) * 0.05
    xq, xs, xsf = ops.per_token_group_quant_mxfp4_sf(x)
    wq, ws = ops.per_token_group_quant_mxfp4(w)
    wsf = ops.mx_sf_pack(ws, True, True)
    assert torch.equal(ops.fp4_gemm_sf(xq, wq.t(), xs, ws, torch.bfloat16, None, xsf, wsf),
                       ops.fp4_gemm(xq, wq.t(), xs, ws, torch.bfloat16))


def test_packed_buffer_checks(ops):
    x = torch.randn(256, 512, dtype=torch.bfloat16, device=DEV)
    out_q = torch.empty(256, 256, dtype=torch.uint8, device=DEV)
    out_s = torch.empty(256, 16, dtype=torch.float32, device=DEV)
    with pytest.raises(RuntimeError, match="too small"):
        torch.ops._vxk.per_token_group_quant_mxfp4_sf(x, out_q, out_s, torch.empty(64, dtype=torch.uint8, device=DEV), 32, 1e-10)
    with pytest.raises(RuntimeError, match="uint8"):
        torch.ops._vxk.per_token_group_quant_mxfp4_sf(x, out_q, out_s, torch.empty(1 << 16, dtype=torch.float32, device=DEV), 32, 1e-10)
