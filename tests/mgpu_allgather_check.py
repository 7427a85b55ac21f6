"""Run under torchrun with >= 2 GPUs (tests/test_multigpu_gpu.py launches it): the fused column-parallel
FP8 GEMM + all-gather (vxk_fp8_gemm_allgather, epilogue peer stores over NVLink) against the unfused path
(fp8_gemm on the local shard + NCCL all-gather) and against the single-GPU product, on every rank."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

# The fused kernels accumulate K in ONE chain per output tile.  Up to 512 tokens the unfused ops would by default run
# their split-K small-M kernels (token tiles side by side), whose fp32 partial sums are added in another order: equal
# to output rounding, not bit for bit.  The bit-for-bit comparison below is against the same-chain kernels, so the
# token-tile dispatch is switched off for this process (read once, at the first call).
os.environ.setdefault("VXK_SKINNY_MTILES", "0")
os.environ.setdefault("VXK_SKINNY_MX4_MAX_M", "128")
os.environ.setdefault("VXK_SKINNY_BLK_MAX_M", "64")


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    import vllm_xpu_kernels_b200._xpu_C  # noqa: F401
    from vllm_xpu_kernels_b200 import shard
    E4 = torch.float8_e4m3fn
    for (M, N, K, per_token, with_bias) in [(300, 1024, 512, True, False), (1000, 4096 + 256, 2048, True, True),
                                            (512, 2048, 1024, False, False), (130, 384, 4096, True, True),
                                            (700, 4096, 1024, True, True)]:
        torch.manual_seed(5)     # same operands on every rank
        a = (torch.randn(M, K, device=dev) / 4).to(E4)
        w = (torch.randn(N, K, device=dev) / 4).to(E4)
        sa = (torch.rand(M, 1, device=dev) * 0.01 + 0.001) if per_token else torch.full((1,), 0.01, device=dev)
        sw = torch.rand(N, 1, device=dev) * 0.01 + 0.001
        bias = (torch.randn(N, device=dev, dtype=torch.bfloat16) / 8) if with_bias else None
        n0, n1 = shard.column_shard(N, rank, world)
        wl, swl = w[n0:n1], sw[n0:n1].contiguous()
        bl = bias[n0:n1].contiguous() if with_bias else None
        full = torch.ops._xpu_C.fp8_gemm(a, w.t(), torch.bfloat16, sa, sw, bias)          # single-GPU product
        so = shard.SymmetricOutput(M, N, torch.bfloat16, dev)
        so.buf.fill_(float("nan"))
        so.barrier()
        got = shard.fp8_gemm_allgather(a, wl.t(), sa, swl, so, n0, bl)
        torch.cuda.synchronize()
        assert not torch.isnan(got.float()).any(), (rank, "unwritten columns", M, N, K)
        # tile shapes differ between the sharded and the unsharded call (accumulation order inside the tensor
        # core does not): the results agree to output rounding
        torch.testing.assert_close(got.float(), full.float(), atol=1e-2, rtol=1e-2)
        local = torch.ops._xpu_C.fp8_gemm(a, wl.t(), torch.bfloat16, sa, swl, bl)
        unfused = shard.all_gather_columns(local, N)
        assert torch.equal(got, unfused), (rank, "fused != fp8_gemm + NCCL all-gather", M, N, K)
        so.barrier()
        if so.mc_ptr and M > 256 and (n1 - n0) % 256 == 0 and n0 % 8 == 0:      # NVLS multicast variant
            so.buf.fill_(float("nan"))
            so.barrier()
            got = shard.fp8_gemm_allgather(a, wl.t(), sa, swl, so, n0, bl, multicast=True)
            torch.cuda.synchronize()
            assert torch.equal(got, unfused), (rank, "multicast fused != fp8_gemm + NCCL all-gather", M, N, K)
            so.barrier()
        del so
    # ---- DeepSeek-V3 block scales (BASELINE config 5): A 1x128 fp32 / UE8M0, W 128x128
    for (M, N, K, ue8m0) in [(300, 1024, 512, False), (1000, 3072, 2048, False), (700, 1536, 1024, True)]:
        torch.manual_seed(11)
        a = (torch.randn(M, K, device=dev) / 4).to(E4)
        w = (torch.randn(N, K, device=dev) / 4).to(E4)
        if ue8m0:
            sa = torch.randint(120, 128, (M, K // 128), dtype=torch.uint8, device=dev).view(torch.float8_e8m0fnu)
            sw = torch.randint(120, 128, (K // 128, N // 128), dtype=torch.uint8, device=dev).view(torch.float8_e8m0fnu)
        else:
            sa = torch.rand(M, K // 128, device=dev) + 0.5
            sw = torch.rand(K // 128, N // 128, device=dev) + 0.5
        n0, n1 = shard.column_shard(N, rank, world)
        wl = w[n0:n1]
        swl = sw.view(torch.uint8)[:, n0 // 128:n1 // 128].contiguous().view(sw.dtype) if ue8m0 \
            else sw[:, n0 // 128:n1 // 128].contiguous()
        so = shard.SymmetricOutput(M, N, torch.bfloat16, dev)
        so.buf.fill_(float("nan"))
        so.barrier()
        got = shard.fp8_gemm_allgather(a, wl.t(), sa, swl, so, n0)
        torch.cuda.synchronize()
        local = torch.ops._xpu_C.fp8_gemm(a, wl.t(), torch.bfloat16, sa, swl, None)
        unfused = shard.all_gather_columns(local, N)
        assert torch.equal(got, unfused), (rank, "block-scaled fused != fp8_gemm + NCCL all-gather", M, N, K, ue8m0)
        so.barrier()
        del so
    # ---- MXFP4 (BASELINE config 4): packed E2M1 operands, E8M0 scales per 32
    E8 = torch.float8_e8m0fnu
    FP4 = torch.float4_e2m1fn_x2
    for (M, N, K) in [(300, 1024, 512), (1000, 3072, 2048), (257, 384, 4096)]:
        torch.manual_seed(9)
        a = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, device=dev).view(FP4)
        w = torch.randint(0, 256, (N, K // 2), dtype=torch.uint8, device=dev).view(FP4)
        sa = torch.randint(118, 130, (M, K // 32), dtype=torch.uint8, device=dev).view(E8)
        sw = torch.randint(118, 130, (N, K // 32), dtype=torch.uint8, device=dev).view(E8)
        n0, n1 = shard.column_shard(N, rank, world)
        wl = w.view(torch.uint8)[n0:n1].view(FP4)
        swl = sw.view(torch.uint8)[n0:n1].contiguous().view(E8)
        so = shard.SymmetricOutput(M, N, torch.bfloat16, dev)
        so.buf.fill_(float("nan"))
        so.barrier()
        got = shard.fp4_gemm_allgather(a, wl.t(), sa, swl, so, n0)
        torch.cuda.synchronize()
        local = torch.ops._xpu_C.fp4_gemm(a, wl.t(), sa, swl, torch.bfloat16, None)
        unfused = shard.all_gather_columns(local, N)
        assert torch.equal(got, unfused), (rank, "mxfp4 fused != fp4_gemm + NCCL all-gather", M, N, K)
        so.barrier()
        del so
    dist.barrier()
    if rank == 0:
        print("mgpu allgather ok", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
