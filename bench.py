#!/usr/bin/env python
"""Benchmark of the quantized-linear hot path (BASELINE.json metric:
"FP8/MXFP4 GEMM TFLOPS (Llama-3 shapes) + quant-kernel HBM GB/s vs peak").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    torchrun --nproc-per-node N bench.py --gpus N ...        (N > 1)

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path over one
batch of synthetic Llama-3-8B activations (BASELINE.json configs[1]):
    rms_norm_dynamic_per_token_quant   T x 4096  bf16 -> e4m3 + [T] scales
    silu_and_mul_per_block_quant       T x 2*14336 bf16 -> e4m3 + [T,112] scales (G = 128)
with T = 8192 tokens per GPU (weak scaling: every rank quantises its own T rows,
no collective -- token rows are independent).  `value` is the algorithmic GB/s of
that step (BASELINE.md section 3 bytes / CUDA-event time, inputs resident in HBM;
per-step working set 692 MB > 126 MB L2, so every step streams from HBM).
`e2e` is the same step through torch.ops with pinned HOST buffers (H2D of the
activations and D2H of codes + scales inside the timed region).
`gemm` reports the FP8 / MXFP4 scaled GEMMs on Llama-3-70B shapes (N-sharded over
the ranks when N > 1) as TFLOP/s, each with its own tensor roofline.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

T_TOKENS = 8192
HIDDEN = 4096
INTER = 14336
GROUP = 128
EPS = 1e-6

# algorithmic bytes per output element (bf16 in)
BYTES_RMS = lambda T: T * HIDDEN * 3 + T * 4                          # noqa: E731
BYTES_SILU = lambda T: T * INTER * 5 + T * (INTER // GROUP) * 4       # noqa: E731


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sustained=d.get("bf16_tflops_sustained"),
                    src="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, src="fallback")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 7:
                    self.samples.append(f)
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no-samples"]}
        sm = sorted(int(float(s[0])) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[3 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": int(float(self.samples[0][1])),
                "power_w_max": max(float(s[2]) for s in self.samples), "samples": len(sm), "reasons": reasons}


# --------------------------------------------------------------------------- reference arm
def cpu_reference_step(T: int, threads: int):
    """The reference's torch CPU implementation of the same step (oracle/, a restatement of
    /root/reference/tests/test_fused_norm_quant.py:37-80 and
    tests/test_fused_silu_mul_block_quant.py:49-79) on T rows.  Returns (seconds, bytes)."""
    import torch
    from oracle import quant as oq
    torch.set_num_threads(threads)
    torch.manual_seed(1234)
    x = torch.randn(T, HIDDEN, dtype=torch.bfloat16) / HIDDEN
    w = torch.empty(HIDDEN, dtype=torch.bfloat16).normal_(1.0, 0.1)
    gu = torch.randn(T, 2 * INTER, dtype=torch.bfloat16) / INTER
    E4 = torch.float8_e4m3fn

    def step():
        oq.rms_norm_dynamic_per_token_quant(x, w, EPS, E4)
        oq.silu_and_mul_per_block_quant(gu, GROUP, E4)

    return step, BYTES_RMS(T) + BYTES_SILU(T)


def run_reference(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    T = 512   # bounded sample of the 8192-row workload
    step, nbytes = cpu_reference_step(T, threads)
    for _ in range(max(1, min(args.warmup, 2))):
        step()
    steps = max(1, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    gbs = nbytes / dt / 1e9
    sample = f"{T} of {T_TOKENS} token rows of the same step, torch {torch.__version__} CPU ops"
    line = {
        "impl": "reference", "metric": "quant-kernel HBM GB/s (Llama-3-8B fused rms_norm+FP8 / silu_mul+block-FP8 step)",
        "value": gbs, "unit": "GB/s", "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16->e4m3", "data": "synthetic",
        "config": {"workload": "llama3-8b fused quant step: rms_norm_dynamic_per_token_quant(Tx4096) + "
                               "silu_and_mul_per_block_quant(Tx2x14336, G=128)", "tokens_per_gpu": T_TOKENS},
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    from vllm_xpu_kernels_b200 import _loader, ops  # raises if the CUDA extension is missing
    E4 = torch.float8_e4m3fn
    pk = peaks()
    T = T_TOKENS

    torch.manual_seed(1234 + rank)
    x = (torch.randn(T, HIDDEN, dtype=torch.bfloat16, device=dev) / HIDDEN)
    w = torch.empty(HIDDEN, dtype=torch.bfloat16, device=dev).normal_(1.0, 0.1)
    gu = (torch.randn(T, 2 * INTER, dtype=torch.bfloat16, device=dev) / INTER)
    q1 = torch.empty(T, HIDDEN, dtype=E4, device=dev)
    s1 = torch.empty(T, dtype=torch.float32, device=dev)
    q2 = torch.empty(T, INTER, dtype=E4, device=dev)
    s2 = torch.empty(T, INTER // GROUP, dtype=torch.float32, device=dev)

    def step():
        torch.ops._C.rms_norm_dynamic_per_token_quant(q1, x, w, s1, EPS, None, None)
        torch.ops._C.silu_and_mul_per_block_quant(q2, gu, s2, GROUP, None, False, False)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # W untimed warm-up steps (at least 3), extended until the device has been busy for ~150 ms so
    # that the timed region does not start on idle clocks (10 steps of this workload are only 2 ms)
    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    t_w = time.perf_counter()
    extra_warm = 0
    while time.perf_counter() - t_w < 0.15:
        for _ in range(20):
            step()
        torch.cuda.synchronize()
        extra_warm += 20
    barrier()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()

    # ---- timed region: K steps, CUDA events on the launching (current) stream ----
    K = args.steps
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
    n_launch0 = _loader.launch_count()
    barrier()
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for i in range(K):
        ev[i][0].record()
        torch.ops._C.rms_norm_dynamic_per_token_quant(q1, x, w, s1, EPS, None, None)
        ev[i][1].record()
        torch.ops._C.silu_and_mul_per_block_quant(q2, gu, s2, GROUP, None, False, False)
        ev[i][2].record()
    t_end.record()
    barrier()
    launches = _loader.launch_count() - n_launch0
    total_ms = t_start.elapsed_time(t_end)
    rms_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / K
    silu_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / K

    tmax = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    total_ms_max = float(tmax.item())
    ms_per_step = total_ms_max / K
    step_bytes = BYTES_RMS(T) + BYTES_SILU(T)
    value = world * step_bytes / (ms_per_step * 1e-3) / 1e9

    # ---- e2e: host buffers, H2D + ops + D2H inside the timed region ----
    hx = torch.empty(T, HIDDEN, dtype=torch.bfloat16).pin_memory()
    hgu = torch.empty(T, 2 * INTER, dtype=torch.bfloat16).pin_memory()
    hx.copy_(x.cpu())
    hgu.copy_(gu.cpu())
    hq1 = torch.empty(T, HIDDEN, dtype=torch.uint8).pin_memory()
    hs1 = torch.empty(T, dtype=torch.float32).pin_memory()
    hq2 = torch.empty(T, INTER, dtype=torch.uint8).pin_memory()
    hs2 = torch.empty(T, INTER // GROUP, dtype=torch.float32).pin_memory()
    dx = torch.empty_like(x)
    dgu = torch.empty_like(gu)

    def e2e_step():
        dx.copy_(hx, non_blocking=True)
        dgu.copy_(hgu, non_blocking=True)
        torch.ops._C.rms_norm_dynamic_per_token_quant(q1, dx, w, s1, EPS, None, None)
        torch.ops._C.silu_and_mul_per_block_quant(q2, dgu, s2, GROUP, None, False, False)
        hq1.copy_(q1.view(torch.uint8), non_blocking=True)
        hs1.copy_(s1, non_blocking=True)
        hq2.copy_(q2.view(torch.uint8), non_blocking=True)
        hs2.copy_(s2, non_blocking=True)

    e2e_steps = max(1, min(K, 5))
    e2e_step()
    barrier()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(e2e_steps):
        e2e_step()
    e1.record()
    barrier()
    e2e_ms = torch.tensor([e0.elapsed_time(e1) / e2e_steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_val = world * step_bytes / (float(e2e_ms.item()) * 1e-3) / 1e9
    h2d = hx.numel() * 2 + hgu.numel() * 2
    d2h = hq1.numel() + hs1.numel() * 4 + hq2.numel() + hs2.numel() * 4

    # ---- GEMMs (Llama-3-70B shapes, N-sharded over the ranks) ----
    gemm = None
    try:
        from bench_gemm import bench_gemms
        gemm = bench_gemms(dev, rank, world, pk, steps=max(3, min(K, 20)))
    except Exception as exc:  # GEMM section must never hide the quant numbers
        gemm = {"error": f"{type(exc).__name__}: {exc}"[:300]}

    if sampler:
        sampler.stop_flag.set()
        sampler.join(timeout=2)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- CPU baseline: the reference's torch CPU path on a bounded sample, rank 0 only at N=1 ----
    cpu_baseline = None
    if world == 1:
        threads = os.cpu_count() or 1
        Tc = 256
        cstep, cbytes = cpu_reference_step(Tc, threads)
        cstep()
        t0 = time.perf_counter()
        reps = 3
        for _ in range(reps):
            cstep()
        cdt = (time.perf_counter() - t0) / reps
        cpu_baseline = {"value": cbytes / cdt / 1e9, "unit": "GB/s", "cores": threads, "kind": "port",
                        "sample": f"{Tc} of {T} token rows of the same step (oracle/, torch CPU ops), "
                                  f"{reps} timed passes"}

    achieved = BYTES_SILU(T) / (silu_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("silu_and_mul_per_block_quant_T8192")
        except Exception:
            traffic = None
    line = {
        "metric": "quant-kernel HBM GB/s (Llama-3-8B fused rms_norm+FP8 / silu_mul+block-FP8 step)",
        "value": value, "unit": "GB/s", "n_gpus": world, "steps": K, "warmup": max(args.warmup, 3) + extra_warm,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16->e4m3", "data": "synthetic",
        "config": {"workload": "llama3-8b fused quant step: rms_norm_dynamic_per_token_quant(Tx4096) + "
                               "silu_and_mul_per_block_quant(Tx2x14336, G=128)",
                   "tokens_per_gpu": T, "hidden": HIDDEN, "intermediate": INTER, "group": GROUP,
                   "l2": "per-step working set 692 MB > 126 MB L2 (inputs larger than L2)",
                   "parallelism": f"token rows sharded over {world} GPU(s), no collective"},
        "roofline": {"bound": "hbm", "kernel": "flat_quant_kernel<bf16,e4m3,silu,div> (silu_and_mul_per_block_quant)",
                     "achieved": achieved, "peak": pk["hbm"], "unit": "GB/s", "frac": achieved / pk["hbm"],
                     "peak_source": pk["src"] + " copy bandwidth", "traffic": traffic,
                     "ms_per_launch": silu_ms, "algorithmic_bytes_per_launch": BYTES_SILU(T),
                     "also": {"rms_norm_dynamic_per_token_quant": {
                         "achieved": BYTES_RMS(T) / (rms_ms * 1e-3) / 1e9, "ms_per_launch": rms_ms,
                         "frac": BYTES_RMS(T) / (rms_ms * 1e-3) / 1e9 / pk["hbm"]}}},
        "cpu_baseline": cpu_baseline,
        "e2e": {"value": e2e_val, "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": float(e2e_ms.item())},
        "gpu_launches": int(launches),
        "clocks": sampler.summary() if sampler else None,
        "gemm": gemm,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
