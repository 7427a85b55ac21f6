"""GPU probe: T(silu(x)) of the fused kernels on every 16-bit input against torch's CPU table
(F.silu on the 16-bit tensor = fp32 evaluation rounded once, tests/test_fused_quant_activation.py:28-40,
tests/test_fused_silu_mul_block_quant.py:49-79).  Prints the mismatching inputs per mode and searches
the approximate form's two constants for a variant with zero mismatches.  Run on the GPU box:
    python tools/silu_probe.py [out.pt]
"""
import ctypes
import os
import struct
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.nn.functional as F

from vllm_xpu_kernels_b200 import _loader

lib = _loader.cdll()
lib.vxk_debug_silu.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int,
                               ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_void_p]
DT = {torch.float16: 1, torch.bfloat16: 2}


def f32_from_bits(b):
    return struct.unpack("f", struct.pack("I", b & 0xFFFFFFFF))[0]


def run(allv_dev, dt, mode, c=0.0, beta=1.0, want_f32=False):
    out = torch.empty_like(allv_dev)
    f32 = torch.empty(allv_dev.numel(), dtype=torch.float32, device="cuda") if want_f32 else None
    rc = lib.vxk_debug_silu(out.data_ptr(), f32.data_ptr() if want_f32 else None, allv_dev.data_ptr(),
                            allv_dev.numel(), DT[dt], mode, c, beta, None)
    assert rc == 0, lib.vxk_last_error()
    return out, f32


def main():
    dump = {}
    for dt in (torch.bfloat16, torch.float16):
        allv = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(dt)
        fin = torch.isfinite(allv.float())
        table = F.silu(allv)
        dev = allv.cuda()
        for mode in (0, 1):
            out, f32 = run(dev, dt, mode, want_f32=True)
            neq = (out.cpu().view(torch.int16) != table.view(torch.int16)) & fin
            idx = neq.nonzero().flatten().tolist()
            print(f"{dt} mode {mode}: {len(idx)} mismatching inputs", [hex(i) for i in idx[:40]])
            for i in idx[:12]:
                print(f"    x={float(allv[i]):+.6g} kernel={float(out[i]):+.9g} torch={float(table[i]):+.9g} "
                      f"f32={float(f32[i]):+.9g}")
            dump[f"{dt}_mode{mode}_t"] = out.cpu().view(torch.int16)
            dump[f"{dt}_mode{mode}_f32"] = f32.cpu()
        dump[f"{dt}_table"] = table.view(torch.int16)
        # search: c = -log2(e) +- k ulp, beta = 1 +- j ulp
        c0 = struct.unpack("I", struct.pack("f", -1.4426950408889634))[0]
        b0 = struct.unpack("I", struct.pack("f", 1.0))[0]
        best = []
        for kc in range(-12, 13):
            for kb in range(-6, 7):
                c = f32_from_bits(c0 + kc)
                beta = f32_from_bits(b0 + kb)
                out, _ = run(dev, dt, 2, c, beta)
                n = int(((out.cpu().view(torch.int16) != table.view(torch.int16)) & fin).sum())
                best.append((n, kc, kb))
        best.sort()
        print(f"{dt} search (mismatches, c ulp offset, beta ulp offset):", best[:12])
        zero = [b for b in best if b[0] == 0]
        print(f"{dt} variants with zero mismatches: {len(zero)}", zero[:20])
    if len(sys.argv) > 1:
        torch.save(dump, sys.argv[1])


if __name__ == "__main__":
    main()
