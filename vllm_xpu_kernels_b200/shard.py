"""Multi-GPU partitioning of the hot path.  The reference has no
parallelism of its own (sharding is the caller's job in vLLM); this module states the
two partitionings the benchmark and a tensor-parallel caller use:

  * activation quantisation shards by TOKEN ROWS: rows are independent, no collective;
  * the scaled GEMMs shard by weight N (column parallel): rank r owns W[n0:n1, :], its
    per-channel / block / MX scales and bias[n0:n1]; activations and their scales are
    replicated.  An all-gather of the [M, N/p] bf16 slabs is needed only when a full
    [M, N] output must be reassembled.

Pure index arithmetic + one torch.distributed collective; works on any backend
(NCCL on the GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist


def row_shard(num_rows: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous [r0, r1) token-row range of `rank` (sizes differ by at most one row)."""
    base, rem = divmod(num_rows, world)
    r0 = rank * base + min(rank, rem)
    return r0, r0 + base + (1 if rank < rem else 0)


def column_shard(n: int, rank: int, world: int, align: int = 128) -> tuple[int, int]:
    """[n0, n1) output-column range of `rank`, boundaries on multiples of `align` (the weight
    block / MMA tile width) so block scales [K/128, N/128] split on whole blocks; the last rank
    takes the remainder."""
    blocks = (n + align - 1) // align
    b0, b1 = row_shard(blocks, rank, world)
    return min(b0 * align, n), min(b1 * align, n)


def shard_weight(w_nk: torch.Tensor, scale: Optional[torch.Tensor], rank: int, world: int,
                 scale_kind: str = "per_channel", align: int = 128):
    """Slice W[N, K] (and its scales) for `rank`.  scale_kind: 'per_channel' ([N] or [N,1]),
    'mx' ([N, K/32]), 'block' ([K/gk, N/gn], N on the last axis), 'per_tensor'."""
    n0, n1 = column_shard(w_nk.shape[0], rank, world, align)
    w = w_nk[n0:n1]
    if scale is None or scale_kind == "per_tensor":
        return w, scale, (n0, n1)
    if scale_kind in ("per_channel", "mx"):
        return w, scale[n0:n1], (n0, n1)
    if scale_kind == "block":
        gn = w_nk.shape[0] // scale.shape[1]
        assert n0 % gn == 0 and (n1 % gn == 0 or n1 == w_nk.shape[0])
        return w, scale[:, n0 // gn:(n1 + gn - 1) // gn], (n0, n1)
    raise ValueError(scale_kind)


def all_gather_columns(local: torch.Tensor, n_total: int, align: int = 128, group=None) -> torch.Tensor:
    """Reassemble the full [M, N] output from per-rank [M, n1-n0] slabs (column_shard layout).
    Slabs may differ in width, so each rank contributes through all_gather of padded slabs."""
    world = dist.get_world_size(group)
    widths = [column_shard(n_total, r, world, align) for r in range(world)]
    wmax = max(b - a for a, b in widths)
    pad = local.new_zeros(local.shape[0], wmax)
    pad[:, :local.shape[1]] = local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[:, :b - a] for o, (a, b) in zip(out, widths)], dim=1)


# ---------------------------------------------------------------------------------------------
# Fused column-parallel FP8 GEMM + all-gather over NVLink peer memory (C ABI vxk_fp8_gemm_allgather)
# ---------------------------------------------------------------------------------------------
class SymmetricOutput:
    """An [M, N_total] output allocated in torch symmetric memory on every rank and mapped into every
    peer: `ptrs[r]` is rank r's buffer as seen from this process.  Create once, reuse across calls."""

    def __init__(self, m: int, n_total: int, dtype: torch.dtype, device, group=None):
        import torch.distributed._symmetric_memory as symm
        self.buf = symm.empty((m, n_total), dtype=dtype, device=device)
        self.hdl = symm.rendezvous(self.buf, group=group if group is not None else dist.group.WORLD)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.world = len(self.ptrs)
        self.rank = int(self.hdl.rank)
        try:      # NVLS multicast mapping of the same buffers (0 / missing when the fabric has no multicast)
            self.mc_ptr = int(self.hdl.multicast_ptr or 0)
        except Exception:
            self.mc_ptr = 0

    def barrier(self):
        """Device-side barrier of all ranks on the current stream: after it every rank's columns have landed."""
        self.hdl.barrier()


def fp8_gemm_allgather(a: torch.Tensor, wt_local: torch.Tensor, a_scale: torch.Tensor, w_scale_local: torch.Tensor,
                       out: SymmetricOutput, col_offset: int, bias_local: Optional[torch.Tensor] = None,
                       sync: bool = True, multicast: bool = False) -> torch.Tensor:
    """D[:, col_offset : col_offset + N_local] = fp8_gemm(a, wt_local) written by the GEMM epilogue into EVERY
    rank's `out.buf` (NVLink peer stores), then one barrier: on return (stream order) `out.buf` holds the full
    [M, N_total] product on all ranks -- GEMM and all-gather in one kernel, no NCCL call.
    a [M, K] e4m3/e5m2; wt_local [K, N_local] K-major view of this rank's W[n0:n1, :]; scales fp32
    (per tensor, per token [M] / [M,1], per channel [N_local] / [N_local,1]), or block scales a_scale [M, K/gk] with
    w_scale_local [K/gk, N_local/gn] in fp32 or float8_e8m0fnu (DeepSeek-V3 recipe; needs M > 128)."""
    import ctypes

    from . import _loader
    lib = _loader.cdll()
    assert a.dim() == 2 and wt_local.dim() == 2 and a.stride(1) == 1 and a.shape[1] == wt_local.shape[0]
    M, K = a.shape
    n_local = wt_local.shape[1]
    dt = {torch.bfloat16: 2, torch.float16: 1}[out.buf.dtype]
    f8 = {torch.float8_e4m3fn: 0, torch.float8_e5m2: 1}[a.dtype]
    block = a_scale.dim() == 2 and a_scale.shape[1] > 1          # [M, K/gk] with w_scale_local [K/gk, N_local/gn]
    i64 = ctypes.c_int64
    gk = gn = sa0 = sa1 = sb0 = sb1 = 0
    scale_dt = 0
    ws, ws_bytes = None, 0
    if block:
        assert w_scale_local.dim() == 2 and w_scale_local.shape[0] == a_scale.shape[1]
        gk, gn = K // w_scale_local.shape[0], n_local // w_scale_local.shape[1]
        sa0, sa1, sb0, sb1 = a_scale.stride(0), a_scale.stride(1), w_scale_local.stride(0), w_scale_local.stride(1)
        a_mode = b_mode = 2
        if a_scale.dtype == torch.float8_e8m0fnu:
            scale_dt = 1
            lib.vxk_fp8_gemm_workspace_bytes.restype = ctypes.c_int64
            ws_bytes = int(lib.vxk_fp8_gemm_workspace_bytes(i64(M), i64(n_local), i64(K), ctypes.c_int(2), ctypes.c_int(1)))
            ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=a.device)
    else:
        a_scale = a_scale.contiguous()
        w_scale_local = w_scale_local.contiguous()
        a_mode = 0 if a_scale.numel() == 1 else 1
        b_mode = 0 if w_scale_local.numel() == 1 else 1
    bias_dt = 0
    if bias_local is not None:
        bias_dt = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}[bias_local.dtype]
    ptrs = (ctypes.c_void_p * out.world)(*out.ptrs)
    # multicast=True sends each row segment once through the NVSwitch (multimem.st) instead of once per peer.
    # An all-gather is bound by every rank's NVLink INGRESS (it must receive (p-1)/p of the output either way),
    # so this measured no faster than unicast at 2 and 8 GPUs (profiles/r01e_bench_n8.json) and is off by default.
    mc = out.mc_ptr if multicast else 0
    if multicast and not mc:
        raise RuntimeError("no multicast mapping on this symmetric-memory handle")
    rc = lib.vxk_fp8_gemm_allgather(
        ptrs, ctypes.c_int(out.world), ctypes.c_int(out.rank), ctypes.c_void_p(mc), i64(col_offset),
        i64(out.buf.stride(0)),
        ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(wt_local.data_ptr()),
        ctypes.c_void_p(a_scale.data_ptr()), ctypes.c_void_p(w_scale_local.data_ptr()),
        ctypes.c_void_p(bias_local.data_ptr() if bias_local is not None else 0),
        i64(M), i64(n_local), i64(K), i64(a.stride(0)), i64(wt_local.stride(0)), i64(wt_local.stride(1)),
        ctypes.c_int(a_mode), ctypes.c_int(b_mode), i64(gk), i64(gn), i64(sa0), i64(sa1), i64(sb0), i64(sb1),
        ctypes.c_int(scale_dt), ctypes.c_int(f8), ctypes.c_int(dt), ctypes.c_int(bias_dt),
        ctypes.c_void_p(ws.data_ptr() if ws is not None else 0), i64(ws_bytes),
        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        raise RuntimeError("vxk_fp8_gemm_allgather: " + lib.vxk_last_error().decode())
    if sync:
        out.barrier()
    return out.buf


def fp4_gemm_allgather(a: torch.Tensor, wt_local: torch.Tensor, a_scale: torch.Tensor, w_scale_local: torch.Tensor,
                       out: SymmetricOutput, col_offset: int, bias_local: Optional[torch.Tensor] = None,
                       sync: bool = True) -> torch.Tensor:
    """MXFP4 counterpart of fp8_gemm_allgather (C ABI vxk_fp4_gemm_allgather): a [M, K/2] float4_e2m1fn_x2,
    wt_local [K/2, N_local] = transposed view of this rank's packed W[n0:n1, :], a_scale [M, K/32] and
    w_scale_local [N_local, K/32] float8_e8m0fnu.  Needs M > 128 (CTA-pair kernel)."""
    import ctypes

    from . import _loader
    lib = _loader.cdll()
    assert a.dim() == 2 and wt_local.dim() == 2 and a.stride(1) == 1 and wt_local.stride(0) == 1
    M, K = a.shape[0], 2 * a.shape[1]
    n_local = wt_local.shape[1]
    dt = {torch.bfloat16: 2, torch.float16: 1}[out.buf.dtype]
    bias_dt = 0
    if bias_local is not None:
        bias_dt = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}[bias_local.dtype]
    i64 = ctypes.c_int64
    lib.vxk_fp4_gemm_workspace_bytes.restype = ctypes.c_int64
    ws_bytes = int(lib.vxk_fp4_gemm_workspace_bytes(i64(M), i64(n_local), i64(K)))
    ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=a.device)
    ptrs = (ctypes.c_void_p * out.world)(*out.ptrs)
    rc = lib.vxk_fp4_gemm_allgather(
        ptrs, ctypes.c_int(out.world), ctypes.c_int(out.rank), i64(col_offset), i64(out.buf.stride(0)),
        ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(wt_local.data_ptr()),
        ctypes.c_void_p(a_scale.data_ptr()), ctypes.c_void_p(w_scale_local.data_ptr()),
        ctypes.c_void_p(bias_local.data_ptr() if bias_local is not None else 0),
        i64(M), i64(n_local), i64(K), i64(a.stride(0)), i64(wt_local.stride(1)),
        i64(a_scale.stride(0)), i64(a_scale.stride(1)), i64(w_scale_local.stride(0)), i64(w_scale_local.stride(1)),
        ctypes.c_int(dt), ctypes.c_int(bias_dt), ctypes.c_void_p(ws.data_ptr()), i64(ws_bytes),
        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        raise RuntimeError("vxk_fp4_gemm_allgather: " + lib.vxk_last_error().decode())
    if sync:
        out.barrier()
    return out.buf
