"""GEMM section of bench.py: FP8 W8A8 (per-token x per-channel) and MXFP4 scaled GEMMs on the
Llama-3-70B shapes of the reference's benchmark table (benchmark/benchmark_gemm_onednn.py:88-116),
column-parallel over the ranks (each rank owns W[n0:n1, :]), timed with CUDA events on the device.

Reported per shape: whole-job TFLOP/s (2*M*N*K over the max-over-ranks time), the fraction of the
dense tensor roofline (FP8 = 2x, FP4 = 4x the measured bf16 cuBLAS peak of MEASURED_PEAKS.json),
and for one shape the time including the NCCL all-gather that reassembles the [M, N] output.
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

LLAMA70B_KN = [(8192, 10240), (8192, 8192), (8192, 57344), (28672, 8192)]
E4 = torch.float8_e4m3fn
E8 = torch.float8_e8m0fnu


def _time(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / steps


def bench_gemms(dev, rank, world, pk, steps=10, m_tokens=8192):
    from vllm_xpu_kernels_b200 import _loader, ops, shard
    lib = _loader.cdll()

    def stats():
        a, b = ctypes.c_uint64(), ctypes.c_uint64()
        lib.vxk_gemm_stats(ctypes.byref(a), ctypes.byref(b))
        return a.value, b.value

    peak_fp8 = 2.0 * pk["bf16"]
    peak_fp4 = 4.0 * pk["bf16"]
    rows = []
    torch.manual_seed(1234)
    M = m_tokens
    for (K, N) in LLAMA70B_KN:
        n0, n1 = shard.column_shard(N, rank, world)
        nloc = n1 - n0
        for kind in ("fp8_w8a8", "mxfp4", "fp8_w8a16"):
            if nloc == 0:
                ms = 0.0
            elif kind == "fp8_w8a8":
                a = (torch.randn(M, K, device=dev) / 4).to(E4)
                w = (torch.randn(nloc, K, device=dev) / 4).to(E4)
                sa = torch.rand(M, 1, device=dev) * 0.01 + 0.001
                sw = torch.rand(nloc, 1, device=dev) * 0.01 + 0.001
                wt = w.t()
                fn = lambda: torch.ops._xpu_C.fp8_gemm(a, wt, torch.bfloat16, sa, sw, None)  # noqa: E731
                ms = _time(fn, steps)
            elif kind == "fp8_w8a16":      # bf16 activations x per-channel FP8 weights
                a = torch.randn(M, K, device=dev, dtype=torch.bfloat16) / 4
                w = (torch.randn(nloc, K, device=dev) / 4).to(E4)
                sa = None
                sw = torch.rand(nloc, device=dev) * 0.01 + 0.001
                wt = w.t()
                fn = lambda: torch.ops._xpu_C.fp8_gemm_w8a16(a, wt, sw, None)  # noqa: E731
                ms = _time(fn, steps)
            else:
                a = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
                w = torch.randint(0, 256, (nloc, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
                sa = torch.randint(118, 130, (M, K // 32), dtype=torch.uint8, device=dev).view(E8)
                sw = torch.randint(118, 130, (nloc, K // 32), dtype=torch.uint8, device=dev).view(E8)
                wt = w.t()
                fn = lambda: torch.ops._xpu_C.fp4_gemm(a, wt, sa, sw, torch.bfloat16, None)  # noqa: E731
                ms = _time(fn, steps)
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_max = float(t.item())
            tflops = 2.0 * M * N * K / (ms_max * 1e-3) / 1e12
            peak = peak_fp8 if kind == "fp8_w8a8" else pk["bf16"] if kind == "fp8_w8a16" else peak_fp4
            rows.append({"kind": kind, "M": M, "N": N, "K": K, "n_per_gpu": nloc if world > 1 else N,
                         "ms": ms_max, "tflops": tflops, "peak_tflops": peak * world,
                         "frac_of_dense_peak": tflops / (peak * world)})
            del a, w, sa, sw, wt
            torch.cuda.empty_cache()

    # reassembled output for one shape (gate_up of Llama-3-70B): GEMM + NCCL all-gather of the bf16 slabs
    gather = None
    if world > 1:
        K, N = 8192, 57344
        n0, n1 = shard.column_shard(N, rank, world)
        a = (torch.randn(M, K, device=dev) / 4).to(E4)
        w = (torch.randn(n1 - n0, K, device=dev) / 4).to(E4)
        sa = torch.rand(M, 1, device=dev) * 0.01
        sw = torch.rand(n1 - n0, 1, device=dev) * 0.01
        wt = w.t()

        def fn():
            out = torch.ops._xpu_C.fp8_gemm(a, wt, torch.bfloat16, sa, sw, None)
            return shard.all_gather_columns(out, N)

        ms = _time(fn, max(2, steps // 2))
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        gather = {"shape": [M, N, K], "kind": "fp8_w8a8", "ms_gemm_plus_allgather": float(t.item()),
                  "tflops_reassembled": 2.0 * M * N * K / (float(t.item()) * 1e-3) / 1e12,
                  "how": "fp8_gemm, then NCCL all-gather of the [M, N/p] slabs"}
        # the same product through ONE kernel: the GEMM epilogue stores each tile into every rank's
        # [M, N] output over NVLink peer mappings (vxk_fp8_gemm_allgather), then one barrier
        try:
            ref_full = fn()
            so = shard.SymmetricOutput(M, N, torch.bfloat16, dev)
            fused = lambda: shard.fp8_gemm_allgather(a, wt, sa, sw, so, n0, multicast=False)  # noqa: E731
            got = fused()
            torch.cuda.synchronize()
            same = bool(torch.equal(got, ref_full))
            ms_f = _time(fused, max(2, steps // 2))
            tf = torch.tensor([ms_f], dtype=torch.float64, device=dev)
            dist.all_reduce(tf, op=dist.ReduceOp.MAX)
            ok = torch.tensor([1.0 if same else 0.0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            gather["fused"] = {"ms_gemm_with_fused_allgather": float(tf.item()),
                               "tflops_reassembled": 2.0 * M * N * K / (float(tf.item()) * 1e-3) / 1e12,
                               "bit_identical_to_nccl_path_on_all_ranks": bool(ok.item() == 1.0),
                               "how": "vxk_fp8_gemm_allgather: epilogue peer stores over NVLink + one barrier"}
            if so.mc_ptr:
                so.buf.zero_()
                so.barrier()
                fmc = lambda: shard.fp8_gemm_allgather(a, wt, sa, sw, so, n0, multicast=True)  # noqa: E731
                got = fmc()
                torch.cuda.synchronize()
                same = torch.tensor([1.0 if torch.equal(got, ref_full) else 0.0], device=dev)
                dist.all_reduce(same, op=dist.ReduceOp.MIN)
                ms_m = _time(fmc, max(2, steps // 2))
                tm = torch.tensor([ms_m], dtype=torch.float64, device=dev)
                dist.all_reduce(tm, op=dist.ReduceOp.MAX)
                gather["fused_multicast"] = {"ms_gemm_with_fused_allgather": float(tm.item()),
                                             "tflops_reassembled": 2.0 * M * N * K / (float(tm.item()) * 1e-3) / 1e12,
                                             "bit_identical_to_nccl_path_on_all_ranks": bool(same.item() == 1.0),
                                             "how": "same kernel, one multimem.st per row segment through the NVSwitch (NVLS)"}
            else:
                gather["fused_multicast"] = {"unavailable": "symmetric-memory handle has no multicast mapping"}
            del ref_full, got, so
        except Exception as exc:  # symmetric memory unavailable on this box / build
            gather["fused"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
        # the same for the MXFP4 GEMM (BASELINE config 4)
        try:
            del a, w, sa, sw, wt
            torch.cuda.empty_cache()
            a4 = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
            w4 = torch.randint(0, 256, (n1 - n0, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
            sa4 = torch.randint(118, 130, (M, K // 32), dtype=torch.uint8, device=dev).view(E8)
            sw4 = torch.randint(118, 130, (n1 - n0, K // 32), dtype=torch.uint8, device=dev).view(E8)
            wt4 = w4.t()

            def fn4():
                return shard.all_gather_columns(torch.ops._xpu_C.fp4_gemm(a4, wt4, sa4, sw4, torch.bfloat16, None), N)

            ref4 = fn4()
            ms_u = _time(fn4, max(2, steps // 2))
            so4 = shard.SymmetricOutput(M, N, torch.bfloat16, dev)
            fused4 = lambda: shard.fp4_gemm_allgather(a4, wt4, sa4, sw4, so4, n0)  # noqa: E731
            got4 = fused4()
            torch.cuda.synchronize()
            same4 = torch.tensor([1.0 if torch.equal(got4, ref4) else 0.0], device=dev)
            ms_f4 = _time(fused4, max(2, steps // 2))
            tt = torch.tensor([ms_u, ms_f4], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(same4, op=dist.ReduceOp.MIN)
            gather["mxfp4"] = {"ms_gemm_plus_allgather": float(tt[0].item()),
                               "ms_gemm_with_fused_allgather": float(tt[1].item()),
                               "tflops_reassembled_fused": 2.0 * M * N * K / (float(tt[1].item()) * 1e-3) / 1e12,
                               "bit_identical_to_nccl_path_on_all_ranks": bool(same4.item() == 1.0)}
            del ref4, got4, so4
        except Exception as exc:
            gather["mxfp4"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
        # and for the DeepSeek-V3 block-scaled FP8 GEMM (BASELINE config 5: 1x128 / 128x128 fp32 scales)
        try:
            torch.cuda.empty_cache()
            Kd, Nd = 7168, 36864
            d0, d1 = shard.column_shard(Nd, rank, world)
            ab = (torch.randn(M, Kd, device=dev) / 4).to(E4)
            wb = (torch.randn(d1 - d0, Kd, device=dev) / 4).to(E4)
            sab = torch.rand(M, Kd // 128, device=dev) + 0.5
            swb = torch.rand(Kd // 128, (d1 - d0) // 128, device=dev) + 0.5
            wtb = wb.t()

            def fnb():
                return shard.all_gather_columns(torch.ops._xpu_C.fp8_gemm(ab, wtb, torch.bfloat16, sab, swb, None), Nd)

            refb = fnb()
            ms_u = _time(fnb, max(2, steps // 2))
            sob = shard.SymmetricOutput(M, Nd, torch.bfloat16, dev)
            fusedb = lambda: shard.fp8_gemm_allgather(ab, wtb, sab, swb, sob, d0)  # noqa: E731
            gotb = fusedb()
            torch.cuda.synchronize()
            sameb = torch.tensor([1.0 if torch.equal(gotb, refb) else 0.0], device=dev)
            ms_fb = _time(fusedb, max(2, steps // 2))
            tt = torch.tensor([ms_u, ms_fb], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(sameb, op=dist.ReduceOp.MIN)
            gather["block_fp8_dsv3"] = {"shape": [M, Nd, Kd], "ms_gemm_plus_allgather": float(tt[0].item()),
                                        "ms_gemm_with_fused_allgather": float(tt[1].item()),
                                        "tflops_reassembled_fused": 2.0 * M * Nd * Kd / (float(tt[1].item()) * 1e-3) / 1e12,
                                        "bit_identical_to_nccl_path_on_all_ranks": bool(sameb.item() == 1.0)}
            del refb, gotb, sob
        except Exception as exc:
            gather["block_fp8_dsv3"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
    tc, fb = stats()
    best = max((r for r in rows if r["kind"] != "fp8_w8a16"), key=lambda r: r["frac_of_dense_peak"]) if rows else None
    return {"shapes": "Llama-3-70B (K,N) x M=%d, N sharded over %d GPU(s)" % (M, world), "rows": rows,
            "best": best, "reassembled": gather, "tensor_core_calls": tc, "fallback_calls": fb,
            "peak_source": pk["src"] + " bf16 cuBLAS x2 (fp8) / x4 (fp4) / x1 (fp8_w8a16: bf16 MMA)"}
