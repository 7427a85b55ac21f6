"""Parity of the indexer-K cache ops and convert_fp8 (torch.ops._C_cache_ops.* -> C ABI) against the CPU
oracle and the committed outputs of the reference's own torch refs.  Shapes follow
tests/test_indexer_k_quant_and_cache.py and tests/test_cp_gather_indexer_k_quant_cache.py of the reference.
Bar: every byte of the cache (codes AND fp32 scales) and of the gathered tensors identical.
"""
import os

import pytest
import torch

from oracle import cache as ocache

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
DEV = "cuda"
E4 = torch.float8_e4m3fn


@pytest.fixture(scope="module")
def ops():
    from vllm_xpu_kernels_b200 import ops as _ops
    return _ops


def build_paging(seqs, bs):
    per = [(n + bs - 1) // bs for n in seqs]
    maxb, nblk = max(per), sum(per)
    table, slots, nxt = [], [], 0
    for n, nb in zip(seqs, per):
        blocks = list(range(nxt, nxt + nb))
        nxt += nb
        table.append(blocks + [0] * (maxb - nb))
        slots += [blocks[p // bs] * bs + p % bs for p in range(n)]
    cu = [0]
    for n in seqs:
        cu.append(cu[-1] + n)
    return (torch.tensor(table, dtype=torch.int32), torch.tensor(slots, dtype=torch.int64),
            torch.tensor(cu, dtype=torch.int32), nblk)


@pytest.mark.parametrize("num_tokens", [1, 8, 17, 64, 1000])
@pytest.mark.parametrize("head_dim", [128, 256, 512])
@pytest.mark.parametrize("scale_fmt", ["ue8m0", "fp8e4m3"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_indexer_k_quant_and_cache(ops, num_tokens, head_dim, scale_fmt, dtype):
    """tests/test_indexer_k_quant_and_cache.py:134-190 (+ shuffled slots, skipped tokens, f16)."""
    G, bs = 128, 16
    groups = head_dim // G
    nblk = (num_tokens + bs - 1) // bs + 1
    cs = head_dim + groups * 4
    torch.manual_seed(0)
    k = torch.randn(num_tokens, head_dim, dtype=dtype)
    slots = torch.randperm(nblk * bs)[:num_tokens].long()
    if num_tokens > 4:
        slots[3] = -1                                    # padding token: must not touch the cache
    ref = torch.zeros(nblk, bs, cs, dtype=torch.uint8)
    ocache.indexer_k_quant_and_cache(k, ref, slots, G, scale_fmt)
    got = torch.zeros(nblk, bs, cs, dtype=torch.uint8, device=DEV)
    ops.indexer_k_quant_and_cache(k.to(DEV), got, slots.to(DEV), G, scale_fmt)
    assert torch.equal(got.cpu(), ref)


@pytest.mark.parametrize("seqs", [[1], [3], [6], [3, 4], [5, 4], [4, 3], [4, 4], [7, 8, 5], [300, 1, 77]])
@pytest.mark.parametrize("head_dim,G", [(128, 128), (256, 128), (512, 128), (24, 24)])
@pytest.mark.parametrize("bs", [4, 16])
@pytest.mark.parametrize("scale_fmt", ["ue8m0", "fp8e4m3"])
def test_cp_gather_indexer_k_quant_cache(ops, seqs, head_dim, G, bs, scale_fmt):
    """tests/test_cp_gather_indexer_k_quant_cache.py:89-230: quantise into the paged cache, gather back."""
    torch.manual_seed(0)
    T = sum(seqs)
    groups = head_dim // G
    cs = head_dim + groups * 4
    table, slots, cu, nblk = build_paging(seqs, bs)
    k = torch.randn(T, head_dim, dtype=torch.bfloat16)
    ref_cache = torch.zeros(nblk, bs, cs, dtype=torch.uint8)
    ocache.indexer_k_quant_and_cache(k, ref_cache, slots, G, scale_fmt)
    cache = torch.zeros(nblk, bs, cs, dtype=torch.uint8, device=DEV)
    ops.indexer_k_quant_and_cache(k.to(DEV), cache, slots.to(DEV), G, scale_fmt)
    assert torch.equal(cache.cpu(), ref_cache)
    rk = torch.zeros(T, head_dim, dtype=torch.uint8)
    rs = torch.zeros(T, groups * 4, dtype=torch.uint8)
    ocache.cp_gather_indexer_k_quant_cache(ref_cache, rk, rs, table, cu)
    dk = torch.zeros(T, head_dim, dtype=torch.uint8, device=DEV)
    ds = torch.zeros(T, groups * 4, dtype=torch.uint8, device=DEV)
    ops.cp_gather_indexer_k_quant_cache(cache, dk, ds, table.to(DEV), cu.to(DEV))
    assert torch.equal(dk.cpu(), rk) and torch.equal(ds.cpu(), rs)
    # round trip property: gathered codes * scales reproduce k to e4m3 precision
    deq = dk.cpu().view(E4).float().reshape(T, groups, G) * rs.view(torch.float32).reshape(T, groups, 1)
    torch.testing.assert_close(deq.reshape(T, head_dim), k.float(), atol=0.3, rtol=0.07)


def test_cache_ops_golden(ops):
    for c in torch.load(os.path.join(GOLD, "f3_indexer_k_cache.pt"), weights_only=False):
        cache = torch.zeros_like(c["cache"]).to(DEV)
        ops.indexer_k_quant_and_cache(c["k"].to(DEV), cache, c["slot_mapping"].to(DEV), c["G"], c["fmt"])
        assert torch.equal(cache.cpu(), c["cache"]), (c["fmt"], c["k"].dtype, c["G"])
        dk = torch.zeros_like(c["dst_k"]).to(DEV)
        ds = torch.zeros_like(c["dst_scale"]).to(DEV)
        ops.cp_gather_indexer_k_quant_cache(cache, dk, ds, c["block_table"].to(DEV), c["cu_seq_lens"].to(DEV))
        assert torch.equal(dk.cpu(), c["dst_k"]) and torch.equal(ds.cpu(), c["dst_scale"])


def test_indexer_full_size_round_trip(ops):
    """Size-independent property at serving scale (64k tokens, DeepSeek-V3.2 indexer head_dim 128):
    scatter then gather returns, for every token, exactly the standalone 1x128 quantisation of its row."""
    T, hd, G, bs = 65536, 128, 128, 64
    torch.manual_seed(1)
    k = torch.randn(T, hd, dtype=torch.bfloat16, device=DEV)
    seqs = [T // 4] * 4
    table, slots, cu, nblk = build_paging(seqs, bs)
    perm = torch.randperm(nblk)                     # shuffle the physical blocks
    table = perm[table.long()].to(torch.int32)
    slots = perm[(slots // bs)] * bs + slots % bs
    cache = torch.zeros(nblk, bs, hd + 4, dtype=torch.uint8, device=DEV)
    ops.indexer_k_quant_and_cache(k, cache, slots.to(DEV), G, "ue8m0")
    dk = torch.zeros(T, hd, dtype=torch.uint8, device=DEV)
    ds = torch.zeros(T, 4, dtype=torch.uint8, device=DEV)
    ops.cp_gather_indexer_k_quant_cache(cache, dk, ds, table.to(DEV), cu.to(DEV))
    idx = torch.tensor([0, 1, 63, 64, T // 2, T - 1])
    q, s = ocache.indexer_group_quant(k[idx.to(DEV)].cpu(), G, True)
    assert torch.equal(dk[idx.to(DEV)].cpu(), q.view(torch.uint8))
    assert torch.equal(ds[idx.to(DEV)].cpu().view(torch.float32), s)
    # and against our own a4 op on the whole tensor (same arithmetic up to the amax floor, none active here)
    q4, s4 = ops.per_token_group_quant_fp8(k, G, eps=1e-4 / 448.0, use_ue8m0=True)
    assert torch.equal(dk, q4.view(torch.uint8))


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("scale", [1.0, 0.5, 0.0123])
@pytest.mark.parametrize("kv", ["fp8", "fp8_e4m3", "auto"])
def test_convert_fp8(ops, dtype, scale, kv):
    torch.manual_seed(0)
    x = torch.randn(7, 16, 33, dtype=dtype) * 3
    if kv == "auto":
        xd = (x.abs() * 20)
        out = torch.empty(x.shape, dtype=torch.uint8, device=DEV)
        ops.convert_fp8(out, xd.to(DEV), scale, kv)
        assert torch.equal(out.cpu(), ocache.convert_fp8(xd, scale, kv, torch.uint8))
        back = torch.empty(x.shape, dtype=dtype, device=DEV)
        ops.convert_fp8(back, out, scale, kv)
        assert torch.equal(back.cpu(), ocache.convert_fp8(out.cpu(), scale, kv, dtype))
        return
    q = torch.empty(x.shape, dtype=torch.uint8, device=DEV)
    ops.convert_fp8(q, x.to(DEV), scale, kv)
    rq = ocache.convert_fp8(x, scale, kv, torch.uint8)
    assert torch.equal(q.cpu(), rq)
    deq = torch.empty(x.shape, dtype=dtype, device=DEV)
    ops.convert_fp8(deq, q, scale, kv)
    assert torch.equal(deq.cpu(), ocache.convert_fp8(rq, scale, kv, dtype))


def test_cache_op_errors(ops):
    k = torch.randn(4, 128, dtype=torch.bfloat16, device=DEV)
    cache = torch.zeros(1, 16, 132, dtype=torch.uint8, device=DEV)
    slots = torch.arange(4, device=DEV)
    with pytest.raises(RuntimeError, match="divisible by quant_block_size"):
        ops.indexer_k_quant_and_cache(k, cache, slots, 96, "ue8m0")
    with pytest.raises(RuntimeError, match="same device"):
        ops.indexer_k_quant_and_cache(k, cache.cpu(), slots, 128, "ue8m0")
    with pytest.raises(RuntimeError, match="Unsupported data type"):
        ops.convert_fp8(torch.empty(4, dtype=torch.uint8, device=DEV), torch.randn(4, device=DEV), 1.0, "fp8_e5m2")
