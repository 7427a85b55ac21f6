"""vxk_fp4_gemm_static_scratch (include/vxk_b200.h): the MXFP4 GEMM for callers that may not allocate.  No
counterpart in the reference.

Called through the C ABI (ctypes) and compared with vxk_fp4_gemm given a caller workspace: same kernels, same
arithmetic, so the outputs are expected to be identical; the assertion uses the GEMM suites' tolerance
(atol = rtol = 1e-2, tests/test_gemm_parity_gpu.py) and additionally reports exact equality.  Two calls in a row
share the library's static scratch array: the second (a different shape) must not see the first one's contents.

This file is collected last on purpose.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda"
BF16 = 2          # VXK_BF16
i64, vp = ctypes.c_int64, ctypes.c_void_p


@pytest.fixture(scope="module")
def lib():
    from vllm_xpu_kernels_b200 import _loader
    lib = _loader.cdll()
    lib.vxk_fp4_gemm_workspace_bytes.restype = ctypes.c_int64
    lib.vxk_fp4_gemm_static_scratch_bytes.restype = ctypes.c_int64
    lib.vxk_last_error.restype = ctypes.c_char_p
    return lib


def operands(m, n, k, seed):
    g = torch.Generator().manual_seed(seed)
    a = torch.randint(0, 256, (m, k // 2), dtype=torch.uint8, generator=g).to(DEV)
    b = torch.randint(0, 256, (n, k // 2), dtype=torch.uint8, generator=g).to(DEV)
    sa = torch.randint(122, 130, (m, k // 32), dtype=torch.uint8, generator=g).to(DEV)   # e8m0: 2^-5 .. 2^2
    sb = torch.randint(122, 130, (n, k // 32), dtype=torch.uint8, generator=g).to(DEV)
    bias = (torch.randn(n, generator=g) / 2).to(torch.bfloat16).to(DEV)
    return a, b, sa, sb, bias


def common_args(d, a, b, sa, sb, bias, m, n, k):
    return [vp(d.data_ptr()), vp(a.data_ptr()), vp(b.data_ptr()), vp(sa.data_ptr()), vp(sb.data_ptr()),
            vp(bias.data_ptr()), i64(m), i64(n), i64(k), i64(k // 2), i64(k // 2), i64(n), i64(k // 32), i64(1),
            i64(k // 32), i64(1), ctypes.c_int(BF16), ctypes.c_int(BF16)]


def run_existing(lib, ops, ws):
    a, b, sa, sb, bias, m, n, k = ops
    d = torch.full((m, n), float("nan"), dtype=torch.bfloat16, device=DEV)
    st = vp(torch.cuda.current_stream().cuda_stream)
    rc = lib.vxk_fp4_gemm(*common_args(d, a, b, sa, sb, bias, m, n, k), vp(ws.data_ptr()), i64(ws.numel()), st)
    assert rc == 0, lib.vxk_last_error()
    return d


def run_static(lib, ops, d):
    a, b, sa, sb, bias, m, n, k = ops
    st = vp(torch.cuda.current_stream().cuda_stream)
    rc = lib.vxk_fp4_gemm_static_scratch(*common_args(d, a, b, sa, sb, bias, m, n, k), st)
    assert rc == 0, lib.vxk_last_error()


# (first call, second call): the second is smaller than the first in one case (stale bytes of the first call lie
# beyond and inside its region) and larger in the other; 130 rows = a partial 128-row block; K = 2880 takes the
# re-pitched scale rows of the single-CTA kernel (K / 32 not a multiple of 16)
@pytest.mark.parametrize("shapes", [((512, 1024, 2048), (130, 384, 1024)), ((130, 256, 512), (700, 1536, 4096)),
                                    ((300, 256, 2880), (256, 512, 1024))])
def test_static_scratch_equals_caller_workspace_on_two_calls_in_a_row(lib, shapes):
    cases = []
    for i, (m, n, k) in enumerate(shapes):
        a, b, sa, sb, bias = operands(m, n, k, seed=17 * m + i)
        cases.append((a, b, sa, sb, bias, m, n, k))
    # everything the test itself allocates exists before the calls under test
    refs, outs = [], []
    for c in cases:
        need = lib.vxk_fp4_gemm_workspace_bytes(i64(c[5]), i64(c[6]), i64(c[7]))
        assert 0 < need + 256 <= lib.vxk_fp4_gemm_static_scratch_bytes(), "the case must use scratch and fit the array"
        ws = torch.empty(need, dtype=torch.uint8, device=DEV)
        refs.append(run_existing(lib, c, ws))
        outs.append(torch.full((c[5], c[6]), float("nan"), dtype=torch.bfloat16, device=DEV))
    run_static(lib, cases[0], outs[0])           # warm-up: first use loads the kernels and resolves the symbol
    torch.cuda.synchronize()
    for o in outs:
        o.fill_(float("nan"))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    alloc0 = torch.cuda.memory_allocated()
    run_static(lib, cases[0], outs[0])           # two calls in a row on one stream, sharing the static array
    run_static(lib, cases[1], outs[1])
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == alloc0, "the entry point must not go through torch's allocator"
    assert free1 >= free0, f"device memory was allocated during the calls ({free0 - free1} bytes)"
    for d, ref in zip(outs, refs):
        assert torch.isfinite(d.float()).all()
        torch.testing.assert_close(d.float(), ref.float(), atol=1e-2, rtol=1e-2)   # the GEMM suites' tolerance
        assert torch.equal(d, ref), "same kernels, same inputs: expected identical bits"


def test_static_scratch_rejects_sizes_beyond_the_array(lib):
    m, n, k = 256, 28672, 8192                   # weight scale factors alone: 7.3 MB packed
    assert lib.vxk_fp4_gemm_workspace_bytes(i64(m), i64(n), i64(k)) > lib.vxk_fp4_gemm_static_scratch_bytes()
    d = torch.zeros(1, dtype=torch.bfloat16, device=DEV)        # never touched: the call is refused up front
    args = [vp(d.data_ptr())] * 6 + [i64(m), i64(n), i64(k), i64(k // 2), i64(k // 2), i64(n), i64(k // 32), i64(1),
                                      i64(k // 32), i64(1), ctypes.c_int(BF16), ctypes.c_int(BF16)]
    rc = lib.vxk_fp4_gemm_static_scratch(*args, vp(torch.cuda.current_stream().cuda_stream))
    assert rc == -1                               # VXK_ERR_INVALID
    msg = lib.vxk_last_error().decode()
    assert "static array holds" in msg and "28672" in msg
    torch.cuda.synchronize()
    assert float(d.float().item()) == 0.0
