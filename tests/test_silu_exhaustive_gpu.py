"""silu(gate) rounded to the 16-bit input type has only 65536 inputs per type: check ALL of them.

The reference's a10 / a11 refs evaluate ``F.silu(gate)`` on the bf16 / f16 tensor
(tests/test_fused_quant_activation.py:28-40, tests/test_fused_silu_mul_block_quant.py:49-79), i.e. the fp32
value x / (1 + exp(-x)) rounded once to the type; tests/golden/silu_all_16bit.pt holds that table (generated
in the build container, tests/golden/make_golden.py) and tests/test_oracle_golden.py pins it to the
closed-form definition.  Here:
  * the kernels' silu -> T (vxk_debug_silu = the device functions the fused kernels call) equals the table on
    every finite input, for the form the default (auto) mode selects per type and for the accurate form;
  * the production a10 / a11 kernels themselves reproduce the oracle's codes and scales on every gate value
    combined with several `up` multipliers.
"""
import ctypes
import os

import pytest
import torch
import torch.nn.functional as F

from oracle import quant
from parity_util import code_stats, scale_stats, u8

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
DEV = "cuda"
E4 = torch.float8_e4m3fn
DT = {torch.float16: 1, torch.bfloat16: 2}


@pytest.fixture(scope="module")
def lib():
    from vllm_xpu_kernels_b200 import _loader
    lib = _loader.cdll()
    lib.vxk_debug_silu.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int,
                                   ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_void_p]
    return lib


def all_values(dtype):
    return torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(dtype)


def kernel_silu(lib, x_dev, mode):
    out = torch.empty_like(x_dev)
    rc = lib.vxk_debug_silu(out.data_ptr(), None, x_dev.data_ptr(), x_dev.numel(), DT[x_dev.dtype], mode, 0.0, 1.0, None)
    assert rc == 0, lib.vxk_last_error()
    torch.cuda.synchronize()
    return out.cpu()


# the form the auto mode takes for a10 / a11 per input type (quant_flat.cu launch_flat)
AUTO_FORM = {torch.bfloat16: 1, torch.float16: 0}


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("which", ["auto", "accurate"])
def test_silu_to_T_matches_torch_on_every_input(lib, dtype, which):
    allv = all_values(dtype)
    fin = torch.isfinite(allv.float())
    table = torch.load(os.path.join(GOLD, "silu_all_16bit.pt"), weights_only=False)[str(dtype)]
    live = F.silu(allv).view(torch.int16)                      # torch on THIS host
    assert torch.equal(live[fin], table[fin]), "torch's table differs between the build container and this host"
    mode = AUTO_FORM[dtype] if which == "auto" else 0
    got = kernel_silu(lib, allv.to(DEV), mode).view(torch.int16)
    bad = ((got != table) & fin).nonzero().flatten().tolist()
    assert not bad, f"{len(bad)} inputs differ, first: {[hex(i) for i in bad[:8]]}"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_a10_a11_on_every_gate_value(dtype):
    """The production kernels on all 65536 gate values x 6 `up` multipliers: bit-exact codes and scales."""
    from vllm_xpu_kernels_b200 import ops
    allv = all_values(dtype)
    gate = torch.where(torch.isfinite(allv.float()), allv, torch.zeros((), dtype=dtype))
    ups = [1.0, -1.0, 0.37109375, -2.5, 96.0, 1.0 / 512]
    x = torch.empty(len(ups), 2 * 65536, dtype=dtype)
    for i, u in enumerate(ups):
        x[i, :65536] = gate
        x[i, 65536:] = u
    # overflowing products (inf) are outside the reference's domain: keep |silu * up| finite
    big = 600.0 if dtype == torch.float16 else 1e30
    x[:, :65536] = x[:, :65536].clamp(-big, big)
    xd = x.to(DEV)
    for sv in (1.0, 0.0078125):
        sc = torch.tensor([sv])
        out = torch.empty(len(ups), 65536, dtype=E4, device=DEV)
        ops.silu_and_mul_quant(out, xd, sc.to(DEV))
        ref = quant.silu_and_mul_quant(x, sc, E4)
        assert torch.equal(u8(out), u8(ref)), f"a10: {code_stats(out, ref)}"
    for qd in (E4, torch.int8):
        q, s = ops.silu_and_mul_per_block_quant(xd, 128, qd)
        rq, rs = quant.silu_and_mul_per_block_quant(x, 128, qd)
        assert torch.equal(s.cpu(), rs), f"a11 scales: {scale_stats(s, rs)}"
        assert torch.equal(u8(q), u8(rq)), f"a11 codes: {code_stats(q, rq, qd == torch.int8)}"
