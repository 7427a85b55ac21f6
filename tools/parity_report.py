"""Observed distance between the CUDA quantisation kernels and the CPU oracle at BASELINE shapes.

For every fused op (a6-a12) it prints, per shape / dtype:
  scales : how many differ, max distance in fp32 ulps (or differing E8M0 bytes)
  codes  : how many differ from the oracle's codes, max distance in code steps -- over ALL elements
  given  : how many differ from the reference formula evaluated WITH THE KERNEL'S scale
           (a scale that legitimately differs by one step changes the codes of its whole group)
rms ops are compared twice: against the oracle with the variance reduced exactly (fp64 mean, rounded
once -- the value the kernels compute) and against torch's own fp32 summation order on this host.

    python tools/parity_report.py [out.txt] [--quick]
Needs a CUDA device; the oracle runs on the host cores (a few minutes at full size).
"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import formats, quant  # noqa: E402
from parity_util import code_stats, nibble_stats, scale_stats, u8  # noqa: E402
from vllm_xpu_kernels_b200 import ops  # noqa: E402

E4 = torch.float8_e4m3fn
DEV = "cuda"
LINES = []


def emit(s=""):
    print(s, flush=True)
    LINES.append(s)


def fmt_codes(tag, st):
    n, frac, d = st
    return f"{tag}: {n} differ ({frac:.2e}), max {d} step" + ("s" if d != 1 else "")


def fmt_scales(st, unit="ulp"):
    n, frac, d = st
    return f"scales: {n} differ ({frac:.2e}), max {d} {unit}"


def norm_inputs(T, H, dtype, add_res, seed=0):
    torch.manual_seed(seed)
    w = torch.empty(H, dtype=dtype).normal_(mean=1.0, std=0.1)
    x = torch.randn(T, H, dtype=dtype) * (1.0 / H)
    res = torch.randn(T, H, dtype=dtype) * (1.0 / H) if add_res else None
    return x, w, res


def report_rms(T, H, dtype, add_res):
    x, w, res = norm_inputs(T, H, dtype, add_res)
    tag = f"T={T} H={H} {str(dtype)[6:]}{' +residual' if add_res else ''}"
    xd, wd = x.to(DEV), w.to(DEV)

    def dres():
        return res.clone().to(DEV) if add_res else None

    for exact in (True, False):
        ref_name = "exact-variance oracle" if exact else "torch-fp32-order oracle"
        normed, rres = quant.rms_norm_f32(x, w, 1e-6, res, exact_variance=exact)
        # a6
        for qd in (E4, torch.int8):
            r = dres()
            q, s = ops.rms_norm_dynamic_per_token_quant(xd, wd, 1e-6, qd, None, r)
            rq, rs = quant._per_token_quant_f32(normed, qd)
            given = quant.fp8_codes_given_scale(normed, s.cpu(), qd, divide=(qd == torch.int8))
            i8 = qd == torch.int8
            emit(f"a6  rms+per-token {str(qd)[6:]:12s} {tag} vs {ref_name}: {fmt_scales(scale_stats(s, rs))}; "
                 f"{fmt_codes('codes', code_stats(q, rq, i8))}; {fmt_codes('given kernel scale', code_stats(q, given, i8))}"
                 + (f"; residual exact: {torch.equal(r.cpu(), rres)}" if add_res else ""))
        # a7
        for G, ue in ((128, False), (64, False), (32, True)):
            q, s = ops.rms_norm_per_block_quant(xd, wd, 1e-6, G, E4, dres(), False, ue)
            rq, rs = quant._per_group_quant_f32(normed, G, E4, ue)
            given = quant.fp8_codes_given_scale(normed, s.cpu(), E4, group=G)
            emit(f"a7  rms+1x{G}{' ue8m0' if ue else ''} e4m3 {tag} vs {ref_name}: {fmt_scales(scale_stats(s, rs))}; "
                 f"{fmt_codes('codes', code_stats(q, rq))}; {fmt_codes('given kernel scale', code_stats(q, given))}")
        # a8
        q, s = ops.rms_norm_mxfp4_quant(xd, wd, 1e-6, dres())
        e, rq = formats.mx_quantize(normed.float().contiguous(), "mxfp4", exact_exponent=True)
        se = u8(s.to(torch.float8_e8m0fnu))
        neb = int((se != u8(e)).sum())
        given = quant.mxfp4_codes_given_exponent(normed, se)
        emit(f"a8  rms+mxfp4 {tag} vs {ref_name}: E8M0 bytes: {neb} differ ({neb / se.numel():.2e}); "
             f"{fmt_codes('codes', nibble_stats(q, rq))}; {fmt_codes('given kernel scale', nibble_stats(q, given))}")
        # a9
        sc = torch.tensor(0.02)
        q = ops.rms_norm_static_fp8_quant(xd, wd, sc.to(DEV), 1e-6, dres())
        rq = quant.fp8_codes_given_scale(normed, sc.reshape(1, 1).expand(T, 1), E4)
        emit(f"a9  rms+static e4m3 {tag} vs {ref_name}: {fmt_codes('codes', code_stats(q, rq))}")


def report_silu(T, H, dtype, mag, lib):
    torch.manual_seed(0)
    x = torch.randn(T, 2 * H, dtype=dtype) * mag
    xd = x.to(DEV)
    tag = f"T={T} H={H} {str(dtype)[6:]} gate~N(0,{mag:g})"
    yT = quant.silu_mul_T(x)
    for fast in (-1, 0, 1):
        prev = lib.vxk_set_silu_fast(fast)
        try:
            # a10
            sc = torch.tensor([0.5])
            out = torch.empty(T, H, dtype=E4, device=DEV)
            ops.silu_and_mul_quant(out, xd, sc.to(DEV))
            rq = quant.silu_and_mul_quant(x, sc, E4)
            emit(f"a10 silu_mul+static e4m3 {tag} silu_fast={fast}: {fmt_codes('codes', code_stats(out, rq))}")
            # a11
            for G in (128,):
                for qd in (E4, torch.int8):
                    q, s = ops.silu_and_mul_per_block_quant(xd, G, qd)
                    rq, rs = quant.silu_and_mul_per_block_quant(x, G, qd)
                    given = quant.fp8_codes_given_scale(yT.float(), s.cpu(), qd, group=G, divide=True)
                    i8 = qd == torch.int8
                    emit(f"a11 silu_mul+1x{G} {str(qd)[6:]:12s} {tag} silu_fast={fast}: {fmt_scales(scale_stats(s, rs))}; "
                         f"{fmt_codes('codes', code_stats(q, rq, i8))}; "
                         f"{fmt_codes('given kernel scale', code_stats(q, given, i8))}")
            # a12
            q, s = ops.silu_and_mul_mxfp4_quant(xd)
            rq, rs = quant.silu_and_mul_mxfp4_quant(x)
            se = u8(s.to(torch.float8_e8m0fnu))
            re = u8(rs.to(torch.float8_e8m0fnu))
            neb = int((se != re).sum())
            given = quant.mxfp4_codes_given_exponent(quant.silu_mul_f32(x), se)
            emit(f"a12 silu_mul+mxfp4 {tag} silu_fast={fast}: E8M0 bytes: {neb} differ ({neb / se.numel():.2e}); "
                 f"{fmt_codes('codes', nibble_stats(q, rq))}; {fmt_codes('given kernel scale', nibble_stats(q, given))}")
        finally:
            lib.vxk_set_silu_fast(prev)


def main():
    quick = "--quick" in sys.argv
    out = next((a for a in sys.argv[1:] if not a.startswith("--")), None)
    from vllm_xpu_kernels_b200 import _loader
    lib = _loader.cdll()
    t0 = time.time()
    emit(f"# parity report: CUDA kernels vs CPU oracle, {torch.cuda.get_device_name(0)}, torch {torch.__version__}, "
         f"host threads {torch.get_num_threads()}, cpu capability {torch.backends.cpu.get_cpu_capability()}")
    emit("# rms_norm ops (a6-a9)")
    shapes = [(1024, 4096)] if quick else [(8192, 4096), (2048, 7168), (512, 16384)]
    for T, H in shapes:
        for dtype in (torch.bfloat16, torch.float16):
            for add_res in (False, True):
                report_rms(T, H, dtype, add_res)
    emit("# silu_and_mul ops (a10-a12)")
    shapes = [(512, 14336)] if quick else [(8192, 14336), (2048, 28672)]
    for T, H in shapes:
        for dtype in (torch.bfloat16, torch.float16):
            for mag in ((1.0,) if (quick or T > 4096) else (1.0, 1.0 / 64, 8.0)):
                report_silu(T, H, dtype, mag, lib)
    emit(f"# done in {time.time() - t0:.0f} s")
    if out:
        with open(out, "w") as f:
            f.write("\n".join(LINES) + "\n")


if __name__ == "__main__":
    main()
