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
