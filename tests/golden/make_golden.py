"""Generate tests/golden/*.pt from the REFERENCE's own Python reference functions.

Run in the build container only (needs /root/reference, which does not exist
on the GPU box):   python tests/golden/make_golden.py

The reference's torch reference implementations live inside its test modules,
which import the XPU extension modules at import time and hard-code
``device="xpu"``.  This script stubs those extension modules with empty ones,
redirects the "xpu" device to the CPU, imports the test modules from
/root/reference unmodified, and calls their reference functions on seeded
inputs.  Inputs and outputs are stored (fp8 as uint8 views) so that
tests/test_oracle_golden.py can pin oracle/ to them without the reference.
"""
import os
import sys
import types

import torch

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def _stub_reference_extensions():
    pkg = types.ModuleType("vllm_xpu_kernels")
    pkg.__path__ = []  # namespace-like: no real submodules are imported
    sys.modules["vllm_xpu_kernels"] = pkg
    for name in ("_C", "_moe_C", "_xpu_C", "_vllm_fa2_C"):
        m = types.ModuleType(f"vllm_xpu_kernels.{name}")
        sys.modules[f"vllm_xpu_kernels.{name}"] = m
        setattr(pkg, name, m)
    sys.path.insert(0, REF)


def u8(t):
    return t.contiguous().view(torch.uint8).clone()


def main():
    _stub_reference_extensions()
    import tests.test_fp8_quant as t_fp8
    import tests.test_fused_norm_quant as t_norm
    import tests.test_fused_quant_activation as t_act
    import tests.test_fused_silu_mul_block_quant as t_silu_blk
    import tests.test_fused_silu_mul_mxfp4_quant as t_silu_fp4
    from tests.ops import fp8_quant_op, mx_utils

    # device="xpu" -> cpu
    t_fp8.as_float32_tensor = lambda x: torch.as_tensor(x, dtype=torch.float32)

    E4, E5 = torch.float8_e4m3fn, torch.float8_e5m2
    g = {}

    # ---- exhaustive E2M1 table over every bf16 bit pattern (mx_utils :152-266)
    allbits = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16)
    allbf16 = allbits.view(torch.bfloat16)
    finite = torch.isfinite(allbf16.float())
    codes = mx_utils._f32_to_floatx_unpacked(
        torch.where(finite, allbf16.float(), torch.zeros(())), 2, 1)
    g["e2m1_all_bf16"] = {"codes": codes, "finite": finite}
    dec = mx_utils._floatx_unpacked_to_f32(torch.arange(16, dtype=torch.uint8), 2, 1)
    g["e2m1_decode"] = {"values": dec}

    # ---- a1 dynamic per-token (test_fp8_quant.py:270-304 inputs)
    cases = []
    for dtype in (torch.bfloat16, torch.float16, torch.float32):
        for qd in (E4, E5):
            for (T, H) in ((7, 67), (5, 1024), (3, 1031)):
                for use_ub in (False, True):
                    torch.manual_seed(0)
                    x = torch.rand(T, H, dtype=dtype) + 1e-6
                    ub = torch.mean(x).to(torch.float32) if use_ub else None
                    q, s = t_fp8.ref_dynamic_per_token_quant(x, qd, ub)
                    cases.append(dict(x=x, qdtype=str(qd), ub=ub, q=u8(q), s=s))
    g["a1_dynamic_per_token"] = cases

    # ---- a3 dynamic per-tensor / a2 static with that scale (:28-45, :401-425)
    cases = []
    for dtype in (torch.bfloat16, torch.float16, torch.float32):
        for qd in (E4, E5):
            torch.manual_seed(0)
            x = torch.rand(11, 520, dtype=dtype)
            q, s = t_fp8.ref_dynamic_per_tensor_fp8_quant(x, qd)
            cases.append(dict(x=x, qdtype=str(qd), q=u8(q), s=s))
    g["a3_dynamic_per_tensor"] = cases

    # ---- a2 static group shapes via scaled_quantize (fp8_quant_op.py:301-348)
    cases = []
    for gs in ((-1, -1), (-1, 1), (1, -1), (1, 128), (128, 128), (4, 256)):
        torch.manual_seed(0)
        x = torch.rand(128, 256, dtype=torch.bfloat16)
        q, s = fp8_quant_op.scaled_quantize(x, gs, E4, compute_dtype=torch.float32)
        cases.append(dict(x=x, group_shape=gs, q=u8(q), s=s))
    g["a2_static_scaled_quantize"] = cases

    # ---- a4 per-token-group (ref_per_block_quant :86-142; to_mxfp for ue8m0)
    cases = []
    for dtype in (torch.float32, torch.bfloat16, torch.float16):
        for G in (32, 64, 128):
            torch.manual_seed(0)
            x = torch.rand(9, 256, dtype=dtype) + 1e-6
            q, s = t_fp8.ref_per_block_quant(x, 1, G)
            cases.append(dict(x=x, G=G, q=u8(q), s=s))
    g["a4_per_block"] = cases
    cases = []
    for dtype in (torch.float32, torch.bfloat16):
        torch.manual_seed(0)
        x = torch.rand(9, 256, dtype=dtype) + 1e-6
        s, q = mx_utils.to_mxfp(x)
        cases.append(dict(x=x, q=u8(q), s_e8m0=u8(s)))
    g["a4_mxfp8"] = cases

    # ---- a5 mxfp4 (to_mxfp mxfp4; known values test_mxfp4_quant.py:113-142)
    cases = []
    for dtype in (torch.float32, torch.bfloat16):
        for (T, H) in ((4, 64), (8, 512)):
            torch.manual_seed(0)
            x = torch.rand(T, H, dtype=dtype) + 1e-6
            s, q = mx_utils.to_mxfp(x, block_size=32, format="mxfp4")
            cases.append(dict(x=x, q=u8(q), s_e8m0=u8(s)))
        torch.manual_seed(1)
        x = (torch.randn(16, 256, dtype=dtype) * 3.0)
        s, q = mx_utils.to_mxfp(x, block_size=32, format="mxfp4")
        cases.append(dict(x=x, q=u8(q), s_e8m0=u8(s)))
    for val in (0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0):
        x = torch.zeros(1, 32, dtype=torch.float32)
        x[0, 0] = val if val > 0 else 0.5
        s, q = mx_utils.to_mxfp(x, block_size=32, format="mxfp4")
        cases.append(dict(x=x, q=u8(q), s_e8m0=u8(s)))
    # mid-points and an all-zero group
    x = torch.zeros(2, 64, dtype=torch.float32)
    x[0, :9] = torch.tensor([6.0, 0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0, -5.0])
    x[1, 32:40] = torch.tensor([-6.0, -0.25, 0.1, 7e-3, 3.0, -3.5, 4.5, 5.5])
    s, q = mx_utils.to_mxfp(x, block_size=32, format="mxfp4")
    cases.append(dict(x=x, q=u8(q), s_e8m0=u8(s)))
    g["a5_mxfp4"] = cases

    # ---- a6..a9 fused rms-norm (test_fused_norm_quant.py:37-125,448-456)
    class Layer:
        pass

    cases = []
    for dtype in (torch.bfloat16, torch.float16):
        for (T, H) in ((7, 128), (5, 1024)):
            for add_res in (False, True):
                torch.manual_seed(0)
                layer = Layer()
                layer.variance_epsilon = 1e-6
                layer.weight = torch.empty(H, dtype=dtype).normal_(mean=1.0, std=0.1)
                x = torch.randn(T, H, dtype=dtype) * (1.0 / H)
                res = torch.randn_like(x) * (1.0 / H) if add_res else None
                normed, res_out = t_norm._ref_rms_norm(layer, x, res)
                c = dict(x=x, w=layer.weight, res=res, res_out=res_out, normed=normed)
                for qd in (E4, torch.int8):
                    q, s = t_norm._ref_per_token_quant(normed, qd)
                    c[f"tok_{qd}"] = (u8(q), s)
                    for G in (64, 128):
                        q, s = t_norm._ref_per_group_quant(normed, G, qd)
                        c[f"grp{G}_{qd}"] = (u8(q), s)
                q, s = t_norm._ref_per_group_quant(normed, 32, E4, scale_ue8m0=True)
                c["grp32_ue8m0"] = (u8(q), s)
                ub = torch.tensor(0.9 * float(normed.abs().max()))
                q, s = t_norm._ref_per_token_quant(normed, E4, ub)
                c["tok_ub"] = (ub, u8(q), s)
                c["static"] = u8(t_norm._ref_static_fp8_quant(normed, torch.tensor(1.0)))
                c["static_0p02"] = u8(t_norm._ref_static_fp8_quant(normed, torch.tensor(0.02)))
                s, q = mx_utils.to_mxfp(normed.float().contiguous(), block_size=32,
                                        format="mxfp4")
                c["mxfp4"] = (u8(q), u8(s))
                cases.append(c)
    g["a6_a9_rms_norm"] = cases

    # ---- a10 silu_and_mul_quant (test_fused_quant_activation.py:28-40)
    cases = []
    for dtype in (torch.bfloat16, torch.float16):
        for qd in (E4, E5):
            torch.manual_seed(0)
            x = torch.randn(7, 2 * 128, dtype=dtype)
            scale = torch.tensor([0.5], dtype=torch.float32)
            q = t_act.ref_silu_and_mul_quant(x, scale, qd)
            cases.append(dict(x=x, scale=scale, qdtype=str(qd), q=u8(q)))
    g["a10_silu_static"] = cases

    # ---- a11 silu_and_mul_per_block_quant (test_fused_silu_mul_block_quant.py:49-79)
    cases = []
    for dtype in (torch.bfloat16, torch.float16):
        for qd in (E4, torch.int8):
            for G in (64, 128):
                torch.manual_seed(0)
                H = 1024
                x = torch.randn(16, 2 * H, dtype=dtype) * (1.0 / H)
                q, s = t_silu_blk.ref_silu_and_mul_per_block_quant(x, qd, G)
                cases.append(dict(x=x, G=G, qdtype=str(qd), q=u8(q), s=s))
    g["a11_silu_block"] = cases

    # ---- a12 silu_and_mul_mxfp4_quant (test_fused_silu_mul_mxfp4_quant.py:22-41)
    cases = []
    for dtype in (torch.bfloat16, torch.float16):
        torch.manual_seed(0)
        H = 1024
        x = torch.randn(7, 2 * H, dtype=dtype) * (1.0 / H)
        q, s_f32, _ = t_silu_fp4._ref_silu_and_mul_mxfp4_quant(x)
        cases.append(dict(x=x, q=u8(q), s=s_f32))
    g["a12_silu_mxfp4"] = cases

    # ---- weight-side helpers used to build GEMM operands
    cases = []
    for ue in (False, True):
        torch.manual_seed(1234)
        w = torch.randn(256, 512, dtype=torch.bfloat16) / 10.0
        q, s = fp8_quant_op.fp8_block_quant_2d(w, 128, 128, use_ue8m0=ue)
        deq = fp8_quant_op.fp8_block_dequant_2d(q, s, 128, 128, torch.bfloat16)
        cases.append(dict(w=w, ue8m0=ue, q=u8(q), s=u8(s) if ue else s, deq=deq))
    g["fp8_block_quant_2d"] = cases

    # ---- a13 / a14 GEMM: the reference tests' expected outputs
    # (test_fp8_gemm_onednn.py:172-214 per-channel; test_fp4_gemm_onednn.py:49-78)
    torch.manual_seed(1234)
    m, n, k = 8, 64, 256
    a = torch.randn(m, k, dtype=torch.bfloat16) / 10.0
    w = torch.randn(n, k, dtype=torch.bfloat16) / 10.0
    aq, asc = t_fp8.ref_dynamic_per_token_quant(a, E4)
    wq, wsc = t_fp8.ref_dynamic_per_token_quant(w, E4)
    a_hp = aq.to(torch.bfloat16) * asc.to(torch.bfloat16)
    w_hp = wq.to(torch.bfloat16) * wsc.to(torch.bfloat16)
    g["a13_fp8_gemm_per_channel"] = dict(aq=u8(aq), asc=asc, wq=u8(wq), wsc=wsc,
                                         out=torch.matmul(a_hp, w_hp.t()))
    torch.manual_seed(1234)
    a = torch.randn(m, k, dtype=torch.bfloat16) * 0.01
    w = torch.randn(n, k, dtype=torch.bfloat16) * 0.01
    sa, qa = mx_utils.to_mxfp(a, format="mxfp4")
    sw, qw = mx_utils.to_mxfp(w, format="mxfp4")
    a_hp = mx_utils.from_blocked_format(
        mx_utils._floatx_unpacked_to_f32(mx_utils.unpack_uint4(qa), 2, 1), sa, 32)
    w_hp = mx_utils.from_blocked_format(
        mx_utils._floatx_unpacked_to_f32(mx_utils.unpack_uint4(qw), 2, 1), sw, 32)
    g["a14_fp4_gemm"] = dict(qa=u8(qa), sa=u8(sa), qw=u8(qw), sw=u8(sw),
                             a_hp=a_hp, w_hp=w_hp,
                             out=torch.matmul(a_hp, w_hp.t()))

    # ---- (f)#1 fp8_gemm_w8a16: the three expected outputs of the reference's tests
    # (test_fp8_gemm_onednn.py:88-126 per-tensor, :224-267 per-channel, :367-410 128x128 block),
    # built from the reference's own helpers and the literal formulas of those tests
    cases = []
    for dtype in (torch.bfloat16, torch.float16):
        m, n, k = 8, 256, 512
        torch.manual_seed(1234)
        a = torch.randn(m, k, dtype=dtype) / 10.0
        w = torch.randn(n, k, dtype=dtype) / 10.0
        # per tensor: weight * (1/scale) -> fp8 (static quant, :100-104), dequantised in `dtype`
        s_t = torch.tensor(4.0)
        wq_t = (w.float() * (1.0 / s_t)).clamp(-448, 448).to(E4)
        out_t = torch.matmul(a, (wq_t.to(dtype) * s_t.to(dtype)).t())
        cases.append(dict(kind="tensor", a=a, wq=u8(wq_t), scale=s_t.reshape(1), out=out_t))
        # per channel: dynamic per-token quant of W[N,K] (:237-239), scale [N,1]
        wq_c, s_c = t_fp8.ref_dynamic_per_token_quant(w, E4)
        out_c = torch.matmul(a, (wq_c.to(dtype) * s_c.to(dtype)).t())
        cases.append(dict(kind="channel", a=a, wq=u8(wq_c), scale=s_c, out=out_c))
        # block: fp8_block_quant_2d / fp8_block_dequant_2d of tests/ops/fp8_quant_op.py (:384-393)
        wq_b, s_b = fp8_quant_op.fp8_block_quant_2d(w, 128, 128, fp8_dtype=E4)
        w_deq = fp8_quant_op.fp8_block_dequant_2d(wq_b, s_b, 128, 128, dtype)
        out_b = torch.matmul(a, w_deq.t())
        cases.append(dict(kind="block", a=a, wq=u8(wq_b), scale=s_b, out=out_b))
    g["f1_fp8_gemm_w8a16"] = cases

    # ---- (f)#3 indexer-K cache ops: the reference tests' own torch refs
    # (tests/test_indexer_k_quant_and_cache.py:95-131, tests/test_cp_gather_indexer_k_quant_cache.py:34-86)
    import tests.test_indexer_k_quant_and_cache as t_idx
    import tests.test_cp_gather_indexer_k_quant_cache as t_gat
    cases = []
    for dtype in (torch.float32, torch.bfloat16):
        for fmt in ("ue8m0", "fp8e4m3"):
            for (seqs, hd, G, bs) in (([17], 128, 128, 16), ([3, 4], 256, 128, 4), ([7, 8, 5], 512, 128, 16),
                                      ([5, 4], 24, 24, 4)):
                torch.manual_seed(0)
                T = sum(seqs)
                k = torch.randn(T, hd, dtype=dtype)
                groups = hd // G
                cs = hd + groups * 4
                # block table / slots as the reference test builds them (:117-150)
                per = [(n + bs - 1) // bs for n in seqs]
                maxb, nblk = max(per), sum(per)
                table, slots, nxt = [], [], 0
                for n, nb in zip(seqs, per):
                    blocks = list(range(nxt, nxt + nb))
                    nxt += nb
                    table.append(blocks + [0] * (maxb - nb))
                    slots += [blocks[p // bs] * bs + p % bs for p in range(n)]
                block_table = torch.tensor(table, dtype=torch.int32)
                slot_mapping = torch.tensor(slots, dtype=torch.int64)
                cu = torch.tensor([0] + list(torch.tensor(seqs).cumsum(0)), dtype=torch.int32)
                cache = torch.zeros(nblk, bs, cs, dtype=torch.uint8)
                t_idx.ref_indexer_k_quant_and_cache(k, cache, slot_mapping, G, fmt)
                dst_k = torch.zeros(T, hd, dtype=torch.uint8)
                dst_s = torch.zeros(T, groups * 4, dtype=torch.uint8)
                t_gat.ref_cp_gather_indexer_k_quant_cache(cache, dst_k, dst_s, block_table, cu)
                cases.append(dict(k=k, G=G, block_size=bs, fmt=fmt, slot_mapping=slot_mapping, block_table=block_table,
                                  cu_seq_lens=cu, cache=cache.clone(), dst_k=dst_k, dst_scale=dst_s))
    g["f3_indexer_k_cache"] = cases

    # ---- (f)#4 MoE grouped GEMM: the reference test's own dequantize_mxfp4 (tests/fused_moe/test_grouped_gemm.py:341-377)
    # called unmodified (its device="xpu" constants redirected to the CPU), and the literal expected-output loops of
    # test_xe_grouped_gemm / _fp8 / _mxfp4 (:131-146, :175-218, :424-442)
    fmi = types.ModuleType("vllm_xpu_kernels.fused_moe_interface")
    fmi.cutlass_grouped_gemm = fmi.cutlass_grouped_gemm_xe2 = None
    sys.modules["vllm_xpu_kernels.fused_moe_interface"] = fmi
    import tests.fused_moe.test_grouped_gemm as t_gg

    class _CpuTorch:
        def __getattr__(self, name):
            return getattr(torch, name)

        @staticmethod
        def tensor(*a, **kw):
            kw.pop("device", None)
            return torch.tensor(*a, **kw)
    t_gg.torch = _CpuTorch()

    def expected(a, w_kn, bias, rows):
        ref, pre = [], 0
        for i in range(len(rows)):
            r = int(rows[i])
            if r == 0:
                continue
            acc = a[pre:pre + r].to(torch.float32) @ w_kn[i].to(torch.float32)
            if bias is not None:
                acc += bias[i]
            ref.append(acc.to(a.dtype))
            pre += r
        return torch.cat(ref, dim=0)

    cases = []
    for dtype in (torch.bfloat16, torch.float16):
        for has_bias in (False, True):
            torch.manual_seed(7)
            E, m, n, k = 4, 24, 64, 128
            rows = torch.tensor([5, 0, 12, 7], dtype=torch.int32)
            a = torch.randn(m, k, dtype=dtype)
            bias = torch.randn(E, n, dtype=dtype) * 100 if has_bias else None
            # bf16 recipe
            b16 = torch.randn(E, k, n, dtype=dtype)
            cases.append(dict(kind="bf16", a=a, b=b16, scales=None, bias=bias, rows=rows,
                              out=expected(a, b16, bias, rows)))
            # fp8 recipe: static quant x * (1/s) (scaled_fp8_quant with a given scale), power-of-two scales (:176-190)
            for qd in (E4, E5):
                sc = torch.pow(2.0, torch.randint(-3, 4, (E,)).float())
                bq = torch.stack([(b16[i].float() * (1.0 / sc[i])).clamp(-torch.finfo(qd).max, torch.finfo(qd).max).to(qd)
                                  for i in range(E)])
                deq = torch.stack([(bq[i].to(torch.float32) * sc[i]).to(dtype) for i in range(E)])
                cases.append(dict(kind="fp8", qdtype=str(qd), a=a, b=u8(bq), scales=sc, bias=bias, rows=rows,
                                  out=expected(a, deq, bias, rows)))
            # mxfp4 recipe (:392-442)
            bq4 = torch.randint(0, 0xff, (E, n, k // 2)).to(torch.uint8)
            sc4 = torch.randint(0, 0x7f, (E, n, k // 32), dtype=torch.uint8)
            w16 = torch.stack([t_gg.dequantize_mxfp4(bq4[i], sc4[i], 32, dtype) for i in range(E)])
            cases.append(dict(kind="mxfp4", a=a, b=bq4, scales=sc4, bias=bias, rows=rows, w16=w16,
                              out=expected(a, w16.transpose(1, 2), bias, rows)))
    g["f4_grouped_gemm"] = cases

    # ---- silu on EVERY 16-bit input, through the reference's own ref function
    # (test_fused_quant_activation.py:28-40: ref_silu_and_mul_quant = F.silu(gate) * up in x.dtype;
    # with up = 1 and scale = 1 the e4m3 cast is the only other step, so the table is taken from
    # F.silu directly -- the same call the ref makes).  The CUDA kernels' silu -> T rounding is held
    # to this table input by input (tests/test_silu_exhaustive_gpu.py).
    import torch.nn.functional as F
    tab = {}
    for dtype in (torch.bfloat16, torch.float16):
        allv = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(dtype)
        tab[str(dtype)] = F.silu(allv).view(torch.int16).clone()
    g["silu_all_16bit"] = tab

    for name, obj in g.items():
        torch.save(obj, os.path.join(OUT, name + ".pt"))
        print(f"{name}: {os.path.getsize(os.path.join(OUT, name + '.pt'))/1024:.1f} KiB")


if __name__ == "__main__":
    main()
