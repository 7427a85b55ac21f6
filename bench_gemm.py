"""GEMM half of bench.py (BASELINE.json metric "FP8/MXFP4 GEMM TFLOPS (Llama-3 shapes) + ..."):

  config 4 / north star   Llama-3-70B (K,N) at M = 8192: FP8 W8A8 (per-token x per-channel), MXFP4, fp8_w8a16
  config 5                DeepSeek-V3 1x128 / 128x128 block-scaled FP8 (fp32 scale grids) at M = 8192
  config 3                Llama-3-8B W8A8 at M in {1, 256, 8192}
                          (shape tables: benchmark/benchmark_gemm_onednn.py:78-116,156-188 of the reference)

Column-parallel over the ranks (rank r owns W[n0:n1, :], shard.column_shard), timed on the device with CUDA
events: >= 10 warm-up + >= 100 timed launches per case (the reference harness uses 5 + 195,
benchmark_gemm_onednn.py:596-617), max over ranks.  Weights rotate over enough copies that the working set
exceeds 2 x the 126 MB L2 (or a single set is already larger), so small-M numbers are HBM numbers.  Cases whose
kernel is shorter than the eager launch path (M <= 256) are timed as CUDA-graph replays of the same rotation,
which is how a serving engine issues its decode step; the row says so ("timed").

Per row: whole-job TFLOP/s (2 M N K / t), the bounding roofline time max(flops / tensor peak, algorithmic bytes /
HBM peak) over t ("frac"), tensor peak = measured bf16 cuBLAS of MEASURED_PEAKS.json x2 (FP8) / x4 (FP4) / x1.
"""
from __future__ import annotations

import ctypes
import os
import time

import torch
import torch.distributed as dist

LLAMA70B_KN = [(8192, 10240), (8192, 8192), (8192, 57344), (28672, 8192)]
LLAMA8B_KN = [(4096, 6144), (4096, 4096), (4096, 28672), (14336, 4096)]
DSV3_KN = [(7168, 1536), (1536, 24576), (512, 32768), (16384, 7168), (7168, 36864), (18432, 7168)]
E4 = torch.float8_e4m3fn
E8 = torch.float8_e8m0fnu
L2_BYTES = 126e6

KERNEL_OF = {   # dominant kernel of each kind at large M (vllm_xpu_kernels_b200/csrc)
    "fp8_w8a8": "gemm_tc2_kernel<KIND_F8,256,*> (tcgen05 kind::f8f6f4, cta_group::2)",
    "mxfp4": "gemm_tc2_kernel<KIND_MXF4,192,*> (tcgen05 kind::mxf4.block_scale, cta_group::2)",
    "fp8_w8a16": "gemm_w8a16_kernel (tcgen05 kind::f16, FP8 widened in the main loop)",
    "block_fp8": "gemm_tc2_blk_kernel<192> (tcgen05 kind::f8f6f4 per K slice + fp32 promotion)",
    "mxfp4_sf": "gemm_tc2_kernel<KIND_MXF4,192,*> via torch.ops._vxk.fp4_gemm_sf (pre-packed scale factors, no sf_pack launches)",
    "block_fp8 (M <= 64)": "gemm_skinny_kernel<MT, BLK> (swap-AB, split-K over a cluster, per-slice promotion in registers)",
    "mxfp4 (M <= 128)": "gemm_skinny_mx4_kernel (swap-AB, split-K over a cluster, natural-layout scales re-tiled in the kernel)",
    "moe_bf16 / moe_fp8 / moe_mxfp4": "gemm_grouped_kernel<MT,kind,*> (tcgen05 kind::f16, swap-AB, device-side tile list; "
                                      "fp8 widened into shared memory, mxfp4 into tensor memory)",
}


def make_case(kind, M, K, nloc, dev, copies):
    """(closures over `copies` independent weight sets, algorithmic bytes of one launch)"""
    fns = []
    if kind == "fp8_w8a8":
        a = (torch.randn(M, K, device=dev) / 4).to(E4)
        sa = torch.rand(M, 1, device=dev) * 0.01 + 0.001
        for _ in range(copies):
            w = (torch.randn(nloc, K, device=dev) / 4).to(E4)
            sw = torch.rand(nloc, 1, device=dev) * 0.01 + 0.001
            fns.append(lambda w=w, sw=sw: torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None))
        nbytes = M * K + K * nloc + 2 * M * nloc + 4 * (M + nloc)
    elif kind == "fp8_w8a16":
        a = torch.randn(M, K, device=dev, dtype=torch.bfloat16) / 4
        for _ in range(copies):
            w = (torch.randn(nloc, K, device=dev) / 4).to(E4)
            sw = torch.rand(nloc, device=dev) * 0.01 + 0.001
            fns.append(lambda w=w, sw=sw: torch.ops._xpu_C.fp8_gemm_w8a16(a, w.t(), sw, None))
        nbytes = 2 * M * K + K * nloc + 2 * M * nloc + 4 * nloc
    elif kind == "mxfp4":
        a = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
        sa = torch.randint(118, 130, (M, K // 32), dtype=torch.uint8, device=dev).view(E8)
        for _ in range(copies):
            w = torch.randint(0, 256, (nloc, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
            sw = torch.randint(118, 130, (nloc, K // 32), dtype=torch.uint8, device=dev).view(E8)
            fns.append(lambda w=w, sw=sw: torch.ops._xpu_C.fp4_gemm(a, w.t(), sa, sw, torch.bfloat16, None))
        nbytes = M * K // 2 + K * nloc // 2 + 2 * M * nloc + (M + nloc) * K // 32
    elif kind == "mxfp4_sf":
        # the extension path: weight scales pre-packed once, activation scales packed by the quant kernel
        a = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
        sa = torch.randint(118, 130, (M, K // 32), dtype=torch.uint8, device=dev).view(E8)
        sa_p = torch.ops._vxk.mx_sf_pack(sa, True, False)
        for _ in range(copies):
            w = torch.randint(0, 256, (nloc, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
            sw = torch.randint(118, 130, (nloc, K // 32), dtype=torch.uint8, device=dev).view(E8)
            sw_p = torch.ops._vxk.mx_sf_pack(sw, True, True)
            fns.append(lambda w=w, sw=sw, sw_p=sw_p: torch.ops._vxk.fp4_gemm_sf(a, w.t(), sa, sw, torch.bfloat16, None, sa_p, sw_p))
        nbytes = M * K // 2 + K * nloc // 2 + 2 * M * nloc + (M + nloc) * K // 32
    elif kind == "block_fp8":
        a = (torch.randn(M, K, device=dev) / 4).to(E4)
        sa = torch.rand(M, K // 128, device=dev) + 0.5
        for _ in range(copies):
            w = (torch.randn(nloc, K, device=dev) / 4).to(E4)
            sw = torch.rand(K // 128, (nloc + 127) // 128, device=dev) + 0.5
            fns.append(lambda w=w, sw=sw: torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None))
        nbytes = M * K + K * nloc + 2 * M * nloc + 4 * (M * K // 128 + (K // 128) * (nloc // 128))
    else:
        raise KeyError(kind)
    return fns, nbytes


def _time_eager(fns, launches, warmup):
    n = len(fns)
    for i in range(max(warmup, n)):
        fns[i % n]()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(launches):
        fns[i % n]()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / launches


def _time_graph(fns, launches, warmup):
    n = len(fns)
    for i in range(max(warmup, n)):
        fns[i % n]()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for i in range(launches):
                fns[i % n]()
    torch.cuda.current_stream().wait_stream(side)
    g.replay()          # warm replay
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    g.replay()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / launches


def time_case(kind, M, K, N, dev, rank, world, pk, launches, warmup):
    from vllm_xpu_kernels_b200 import shard
    n0, n1 = shard.column_shard(N, rank, world)
    nloc = n1 - n0
    wbytes = K * max(nloc, 1) // (2 if kind.startswith("mxfp4") else 1)
    setbytes = wbytes + M * K + 2 * M * max(nloc, 1)
    copies = 1 if setbytes > 2 * L2_BYTES else max(2, min(16, int(2 * L2_BYTES // setbytes) + 1))
    graph = M <= 256
    if nloc == 0:
        ms, nbytes = 0.0, 0
    else:
        fns, nbytes = make_case(kind, M, K, nloc, dev, copies)
        ms = (_time_graph if graph else _time_eager)(fns, launches, warmup)
        del fns
    t = torch.tensor([ms, float(nbytes)], dtype=torch.float64, device=dev)
    if world > 1:
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ts = t.clone()
        dist.all_reduce(ts, op=dist.ReduceOp.SUM)
        ms, nbytes = float(tm[0].item()), float(ts[1].item())
    torch.cuda.empty_cache()
    peak_tf = {"fp8_w8a8": 2.0, "block_fp8": 2.0, "mxfp4": 4.0, "mxfp4_sf": 4.0, "fp8_w8a16": 1.0}[kind] * pk["bf16"] * world
    flops = 2.0 * M * N * K
    t_tensor = flops / (peak_tf * 1e12)
    t_hbm = nbytes / (pk["hbm"] * world * 1e9)
    bound = "tensor" if t_tensor >= t_hbm else "hbm"
    return {"kind": kind, "M": M, "N": N, "K": K, "n_per_gpu": nloc if world > 1 else N, "ms": ms,
            "tflops": flops / (ms * 1e-3) / 1e12 if ms > 0 else 0.0,
            "gbs": nbytes / (ms * 1e-3) / 1e9 if ms > 0 else 0.0,
            "bound": bound, "peak_tflops": peak_tf, "frac": max(t_tensor, t_hbm) * 1e3 / ms if ms > 0 else 0.0,
            "frac_of_dense_peak": flops / (ms * 1e-3) / 1e12 / peak_tf if ms > 0 else 0.0,
            "launches": launches, "warmup": warmup, "weight_copies": copies,
            "timed": "cuda-graph replay" if graph else "eager torch.ops launches"}


def bench_gemms(dev, rank, world, pk, launches=100, warmup=10, m_tokens=8192, sets=("70b", "dsv3", "8b")):
    from vllm_xpu_kernels_b200 import _loader
    lib = _loader.cdll()

    def stats():
        a, b = ctypes.c_uint64(), ctypes.c_uint64()
        lib.vxk_gemm_stats(ctypes.byref(a), ctypes.byref(b))
        return a.value, b.value

    torch.manual_seed(1234)
    M = m_tokens
    rows70, rows5, rows3, rows4 = [], [], [], []
    if "70b" in sets:
        for (K, N) in LLAMA70B_KN:
            for kind in ("fp8_w8a8", "mxfp4", "fp8_w8a16"):
                rows70.append(time_case(kind, M, K, N, dev, rank, world, pk, launches, warmup))
        # the same MXFP4 GEMM with both scale grids already in the tensor core's layout (no re-packing launches)
        rows70.append(time_case("mxfp4_sf", M, LLAMA70B_KN[0][0], LLAMA70B_KN[0][1], dev, rank, world, pk, launches, warmup))
        # config 4 in the decode regime (swap-AB split-K MXFP4 kernel)
        for (K, N) in LLAMA70B_KN:
            for m in (1, 64):
                rows4.append(time_case("mxfp4", m, K, N, dev, rank, world, pk, launches, warmup))
    rows5d = []
    if "dsv3" in sets:
        for (K, N) in DSV3_KN:
            rows5.append(time_case("block_fp8", M, K, N, dev, rank, world, pk, launches, warmup))
        # config 5 in the decode regime (block-scaled variant of the swap-AB split-K kernel)
        for (K, N) in DSV3_KN:
            for m in (1, 64):
                rows5d.append(time_case("block_fp8", m, K, N, dev, rank, world, pk, launches, warmup))
    if "8b" in sets:
        for (K, N) in LLAMA8B_KN:
            for m in (1, 256, M):
                rows3.append(time_case("fp8_w8a8", m, K, N, dev, rank, world, pk, launches, warmup))
    tc, fb = stats()
    return {"llama70b_m8192": rows70, "llama70b_mxfp4_decode": rows4, "dsv3_block_fp8_m8192": rows5,
            "dsv3_block_fp8_decode": rows5d,
            "llama8b_w8a8_m_sweep": rows3,
            "kernels": KERNEL_OF, "tensor_core_calls": tc, "fallback_calls": fb,
            "peak_source": pk["src"] + " bf16 cuBLAS x2 (fp8) / x4 (fp4) / x1 (fp8_w8a16: bf16 MMA); "
                           "hbm = " + pk["src"] + " copy bandwidth",
            "method": f">= {warmup} warm-up + {launches} timed launches per case, CUDA events, max over ranks, "
                      "weights rotated through > 2 x L2 unless one set already exceeds it"}


# ---------------------------------------------------------------------------------------------
# MoE grouped GEMM: the reference test's shape, decode and prefill token counts
# ---------------------------------------------------------------------------------------------
def bench_moe_grouped(dev, pk, launches=100, warmup=10):
    """torch.ops._xpu_C.cutlass_grouped_gemm_interface on tests/fused_moe/test_grouped_gemm.py:14-21 (N = 5120,
    K = 8192, 16 experts, top-1), bf16 activations, the three recipes.  rows_per_expert is redrawn per launch so
    that consecutive launches read different experts (HBM numbers, not L2 numbers).  Roofline time =
    max(2MNK / bf16 dense peak, (weights of the experts that own rows + activations + output) / HBM peak)."""
    from vllm_xpu_kernels_b200.fused_moe_interface import cutlass_grouped_gemm_xe2
    E, N, K = 16, 5120, 8192
    rows_out = []
    for kind in ("bf16", "fp8", "fp8_nk", "mxfp4"):
        if kind == "bf16":
            b, sc, wb = torch.randn(E, K, N, device=dev, dtype=torch.bfloat16) / 8, None, 2.0
        elif kind == "fp8_nk":    # extension: the fp8 recipe with the weights stored [E, N, K] (transposed view handed over)
            b = torch.empty(E, N, K, device=dev, dtype=E4)
            for e in range(E):
                b[e] = (torch.randn(N, K, device=dev) / 4).to(E4)
            b, sc, wb = b.transpose(1, 2), torch.rand(E, device=dev) + 0.5, 1.0
        elif kind == "fp8":
            b = torch.empty(E, K, N, device=dev, dtype=E4)
            for e in range(E):
                b[e] = (torch.randn(K, N, device=dev) / 4).to(E4)
            sc, wb = torch.rand(E, device=dev) + 0.5, 1.0
        else:
            b = torch.randint(0, 256, (E, N, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
            sc, wb = torch.randint(118, 130, (E, N, K // 32), dtype=torch.uint8, device=dev), 0.5 + 1.0 / 32
        for M in (16, 256, 8192):
            a = torch.randn(M, K, device=dev, dtype=torch.bfloat16) / 8
            d = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            copies = 8 if M < 1024 else 2
            fns, act = [], 0.0
            for i in range(copies):
                g = torch.Generator().manual_seed(100 + i)
                rows = torch.bincount(torch.rand(M, E, generator=g).argmax(dim=1), minlength=E).to(torch.int32)
                act += float((rows > 0).sum()) / copies
                r = rows.to(dev)
                fns.append(lambda r=r, a=a, d=d: cutlass_grouped_gemm_xe2(a, b, sc, None, d, r, N, K, E))
            try:
                ms, graph = _time_graph(fns, launches, warmup), True
            except Exception:
                ms, graph = _time_eager(fns, launches, warmup), False
            flops = 2.0 * M * N * K
            nbytes = act * N * K * wb + 2.0 * M * K + 2.0 * M * N
            t_tensor, t_hbm = flops / (pk["bf16"] * 1e12), nbytes / (pk["hbm"] * 1e9)
            rows_out.append({"kind": "moe_" + kind, "M": M, "N": N, "K": K, "experts": E, "active_experts": act,
                             "ms": ms, "tflops": flops / (ms * 1e-3) / 1e12, "gbs": nbytes / (ms * 1e-3) / 1e9,
                             "bound": "tensor" if t_tensor > t_hbm else "hbm",
                             "frac": max(t_tensor, t_hbm) * 1e3 / ms,
                             "timed": "cuda-graph replay" if graph else "eager torch.ops launches"})
        del b, sc
        torch.cuda.empty_cache()
    return rows_out


# ---------------------------------------------------------------------------------------------
# fused gate_up projection: GEMM -> silu_and_mul -> 1x128 block FP8 quant in one kernel
# ---------------------------------------------------------------------------------------------
def bench_fused_gate_up(dev, pk, launches=100, warmup=10, m_tokens=8192):
    """torch.ops._vxk.fp8_gemm_silu_mul_quant against the two reference-surface ops it replaces
    (torch.ops._xpu_C.fp8_gemm + torch.ops._C.silu_and_mul_per_block_quant), Llama-3-8B and -70B MLP shapes."""
    from vllm_xpu_kernels_b200 import ops
    rows = []
    for name, K, inter in (("llama3-8b", 4096, 14336), ("llama3-70b", 8192, 28672)):
        M = m_tokens
        a = (torch.randn(M, K, device=dev) / 4).to(E4)
        sa = torch.rand(M, 1, device=dev) * 0.01 + 0.001
        ws = [(torch.randn(2 * inter, K, device=dev) / 4).to(E4) for _ in range(2)]
        sw = torch.rand(2 * inter, 1, device=dev) * 0.01 + 0.001
        two = [lambda w=w: ops.silu_and_mul_per_block_quant(
            torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None), 128, E4) for w in ws]
        fused = [lambda w=w: ops.fp8_gemm_silu_mul_quant(a, w.t(), sa, sw, None, torch.bfloat16) for w in ws]
        q0, s0 = two[0]()
        q1, s1 = fused[0]()
        same = bool(torch.equal(q0.view(torch.uint8), q1.view(torch.uint8)) and torch.equal(s0, s1))
        t2, tf = _time_graph(two, launches, warmup), _time_graph(fused, launches, warmup)
        flops = 2.0 * M * 2 * inter * K
        rows.append({"kind": "fp8_gemm_silu_mul_quant", "model": name, "M": M, "K": K, "inter": inter,
                     "ms_two_ops": t2, "ms_fused": tf, "speedup": t2 / tf, "tflops_fused": flops / (tf * 1e-3) / 1e12,
                     "frac_of_fp8_peak": flops / (tf * 1e-3) / 1e12 / (2.0 * pk["bf16"]),
                     "bit_identical_to_two_ops": same,
                     "hbm_bytes_saved": 2 * 2 * M * 2 * inter})      # the bf16 intermediate: written once, read once
        del ws
        torch.cuda.empty_cache()
    return rows


# ---------------------------------------------------------------------------------------------
# e2e: one GEMM call through torch.ops._xpu_C with HOST activations and a HOST result
# ---------------------------------------------------------------------------------------------
def e2e_gemm(dev, rank, world, steps=5):
    """fp8_gemm at the first Llama-3-70B shape: every step copies the quantised activations + scales from
    pinned host memory, runs the op on the resident weight shard, and reads the bf16 result back."""
    from vllm_xpu_kernels_b200 import shard
    M, (K, N) = 8192, LLAMA70B_KN[0]
    n0, n1 = shard.column_shard(N, rank, world)
    nloc = n1 - n0
    ha = torch.randint(0, 120, (M, K), dtype=torch.uint8).pin_memory()
    hsa = (torch.rand(M, 1) * 0.01 + 0.001).pin_memory()
    w = (torch.randn(nloc, K, device=dev) / 4).to(E4)
    sw = torch.rand(nloc, 1, device=dev) * 0.01 + 0.001
    da = torch.empty(M, K, dtype=torch.uint8, device=dev)
    dsa = torch.empty(M, 1, device=dev)
    hout = torch.empty(M, nloc, dtype=torch.bfloat16).pin_memory()

    def step():
        da.copy_(ha, non_blocking=True)
        dsa.copy_(hsa, non_blocking=True)
        out = torch.ops._xpu_C.fp8_gemm(da.view(E4), w.t(), torch.bfloat16, dsa, sw, None)
        hout.copy_(out, non_blocking=True)

    step()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        step()
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / steps
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    return {"op": "torch.ops._xpu_C.fp8_gemm", "M": M, "N": N, "K": K, "ms_per_step": ms,
            "value": 2.0 * M * N * K / (ms * 1e-3) / 1e12, "unit": "TFLOP/s",
            "h2d_bytes_per_step": ha.numel() + hsa.numel() * 4, "d2h_bytes_per_step": hout.numel() * 2,
            "note": "activations + scales H2D and the bf16 result D2H every step; the weight shard is resident"}


# ---------------------------------------------------------------------------------------------
# CPU baseline of the GEMM: the reference test's formula on the host cores, reduced M
# ---------------------------------------------------------------------------------------------
def cpu_gemm_baseline(threads, m=512, reps=2):
    """dequantise (oracle helpers) -> torch.matmul in fp32 -> bf16, per-token x per-channel FP8 at the first
    Llama-3-70B shape with M reduced to `m`; returns TFLOP/s on `threads` cores."""
    from oracle import quant as oq
    torch.set_num_threads(threads)
    K, N = LLAMA70B_KN[0]
    torch.manual_seed(1234)
    a = torch.randn(m, K, dtype=torch.bfloat16) / 10.0
    w = torch.randn(N, K, dtype=torch.bfloat16) / 10.0
    aq, sa = oq.dynamic_per_token_scaled_fp8_quant(a, E4)
    wq, sw = oq.dynamic_per_token_scaled_fp8_quant(w, E4)

    def run():
        return torch.matmul(aq.float() * sa, (wq.float() * sw).t()).to(torch.bfloat16)

    run()
    t0 = time.perf_counter()
    for _ in range(reps):
        run()
    dt = (time.perf_counter() - t0) / reps
    return {"value": 2.0 * m * N * K / dt / 1e12, "unit": "TFLOP/s", "cores": threads, "kind": "port",
            "sample": f"M = {m} of 8192 rows, N = {N}, K = {K}: oracle dequantise -> fp32 torch.matmul -> bf16 "
                      f"(tests/test_fp8_gemm_onednn.py:174-214), {reps} timed passes", "ms": dt * 1e3}


# ---------------------------------------------------------------------------------------------
# reassembled output (N > 1): GEMM + NCCL all-gather vs the fused epilogue-peer-store kernels
# ---------------------------------------------------------------------------------------------
def bench_reassembled(dev, rank, world, steps=10):
    from vllm_xpu_kernels_b200 import shard
    M = 8192
    out = {}

    def timeit(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(n):
            fn()
        e.record()
        torch.cuda.synchronize()
        t = torch.tensor([s.elapsed_time(e) / n], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def all_same(a, b):
        ok = torch.tensor([1.0 if torch.equal(a, b) else 0.0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        return bool(ok.item() == 1.0)

    cases = [("fp8_w8a8", 8192, 57344), ("mxfp4", 8192, 57344), ("block_fp8", 7168, 36864)]
    for kind, K, N in cases:
        try:
            torch.cuda.empty_cache()
            n0, n1 = shard.column_shard(N, rank, world)
            nloc = n1 - n0
            if kind == "fp8_w8a8":
                a = (torch.randn(M, K, device=dev) / 4).to(E4)
                w = (torch.randn(nloc, K, device=dev) / 4).to(E4)
                sa = torch.rand(M, 1, device=dev) * 0.01
                sw = torch.rand(nloc, 1, device=dev) * 0.01
                plain = lambda: torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None)   # noqa: E731
                fused = lambda so: shard.fp8_gemm_allgather(a, w.t(), sa, sw, so, n0, multicast=False)  # noqa: E731
            elif kind == "mxfp4":
                a = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
                w = torch.randint(0, 256, (nloc, K // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
                sa = torch.randint(118, 130, (M, K // 32), dtype=torch.uint8, device=dev).view(E8)
                sw = torch.randint(118, 130, (nloc, K // 32), dtype=torch.uint8, device=dev).view(E8)
                plain = lambda: torch.ops._xpu_C.fp4_gemm(a, w.t(), sa, sw, torch.bfloat16, None)   # noqa: E731
                fused = lambda so: shard.fp4_gemm_allgather(a, w.t(), sa, sw, so, n0)               # noqa: E731
            else:
                a = (torch.randn(M, K, device=dev) / 4).to(E4)
                w = (torch.randn(nloc, K, device=dev) / 4).to(E4)
                sa = torch.rand(M, K // 128, device=dev) + 0.5
                sw = torch.rand(K // 128, nloc // 128, device=dev) + 0.5
                plain = lambda: torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, None)   # noqa: E731
                fused = lambda so: shard.fp8_gemm_allgather(a, w.t(), sa, sw, so, n0)               # noqa: E731
            nccl = lambda: shard.all_gather_columns(plain(), N)                                     # noqa: E731
            ref_full = nccl()
            ms_nccl = timeit(nccl, steps)
            so = shard.SymmetricOutput(M, N, torch.bfloat16, dev)
            got = fused(so)
            torch.cuda.synchronize()
            same = all_same(got, ref_full)
            ms_fused = timeit(lambda: fused(so), steps)
            ms_gemm = timeit(plain, steps)
            flops = 2.0 * M * N * K
            out[kind] = {"shape": [M, N, K], "ms_gemm_only": ms_gemm, "ms_gemm_plus_nccl_allgather": ms_nccl,
                         "ms_fused_gemm_allgather": ms_fused,
                         "tflops_reassembled_fused": flops / (ms_fused * 1e-3) / 1e12,
                         "bit_identical_to_nccl_path_on_all_ranks": same}
            del ref_full, got, so, a, w, sa, sw
        except Exception as exc:   # symmetric memory unavailable on this box / build
            out[kind] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
    return out
