"""CPU-only checks (no GPU): the C-ABI library loads and exports every symbol include/vxk_b200.h
declares, the product never routes through oracle/, the operator schemas are the reference's, and
the multi-GPU partitioning (world_size 2, gloo) reassembles the unsharded result."""
import ctypes
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vllm_xpu_kernels_b200")
LIB = os.path.join(PKG, "_lib", "libvxk_b200.so")

needs_lib = pytest.mark.skipif(not os.path.exists(LIB), reason="extension not built (python -m vllm_xpu_kernels_b200.build)")


@needs_lib
def test_c_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "vxk_b200.h")).read()
    names = sorted(set(re.findall(r"\b(vxk_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 20
    lib = ctypes.CDLL(LIB)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/vxk_b200.h but not exported"
    lib.vxk_abi_version.restype = ctypes.c_int
    assert lib.vxk_abi_version() == 8
    lib.vxk_last_error.restype = ctypes.c_char_p
    assert isinstance(lib.vxk_last_error(), bytes)


@needs_lib
def test_c_abi_argument_errors_without_a_gpu():
    """Argument validation happens before any CUDA call, so it is testable on the CPU."""
    lib = ctypes.CDLL(LIB)
    lib.vxk_last_error.restype = ctypes.c_char_p
    i64 = ctypes.c_int64
    rc = lib.vxk_per_token_group_quant_mxfp4(None, None, None, i64(4), i64(64), i64(64), ctypes.c_double(0), i64(2),
                                             i64(1), 2, None)
    assert rc == -1 and b"group_size == 32" in lib.vxk_last_error()
    rc = lib.vxk_silu_and_mul_per_block_quant(None, None, None, i64(4), i64(512), i64(256), None, i64(2), i64(1), 0, 2,
                                              0, None)
    assert rc == -1 and b"group_size must be 32, 64 or 128" in lib.vxk_last_error()
    rc = lib.vxk_dynamic_per_token_scaled_fp8_quant(None, None, None, None, i64(4), i64(64), 2, 2, None)
    assert rc == -1 and b"fp8" in lib.vxk_last_error()


@needs_lib
def test_ops_registered_with_reference_schemas_and_no_cpu_kernel():
    import vllm_xpu_kernels_b200._C  # noqa: F401
    import vllm_xpu_kernels_b200._xpu_C  # noqa: F401
    expect = {
        "_C::dynamic_per_token_scaled_fp8_quant":
            "_C::dynamic_per_token_scaled_fp8_quant(Tensor($0! -> ) result, Tensor input, Tensor($1! -> ) scale, Tensor? scale_ub) -> ()",
        "_C::silu_and_mul_mxfp4_quant":
            "_C::silu_and_mul_mxfp4_quant(Tensor($0! -> ) out, Tensor input, Tensor($1! -> ) scales, int group_size, float eps=1e-10) -> ()",
        "_xpu_C::fp8_gemm":
            "_xpu_C::fp8_gemm(Tensor A, Tensor B, ScalarType? out_dtype, Tensor? A_scale_, Tensor? B_scale_, Tensor? bias_) -> Tensor",
        "_xpu_C::fp4_gemm":
            "_xpu_C::fp4_gemm(Tensor A, Tensor B, Tensor A_scale, Tensor B_scale, ScalarType? out_dtype, Tensor? bias_) -> Tensor",
    }
    for name, schema in expect.items():
        ns, op = name.split("::")
        got = str(getattr(getattr(torch.ops, ns), op).default._schema)
        assert got.replace(" ", "") == schema.replace(" ", ""), got
    for op in ("rms_norm_dynamic_per_token_quant", "rms_norm_per_block_quant", "rms_norm_mxfp4_quant",
               "rms_norm_static_fp8_quant", "fused_add_rms_norm_static_fp8_quant", "silu_and_mul_quant",
               "silu_and_mul_per_block_quant", "static_scaled_fp8_quant", "dynamic_scaled_fp8_quant",
               "per_token_group_fp8_quant", "per_token_group_quant_mxfp4"):
        assert hasattr(torch.ops._C, op)
    # no CPU fallback: a CPU tensor must be refused by the dispatcher
    with pytest.raises(NotImplementedError):
        torch.ops._C.silu_and_mul_quant(torch.empty(2, 8, dtype=torch.float8_e4m3fn),
                                        torch.zeros(2, 16, dtype=torch.bfloat16), torch.ones(1))


def test_product_never_imports_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may touch oracle/."""
    bad = []
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M) or "oracle/" in txt and f.endswith(".py") and "import" in txt and "from oracle" in txt:
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad


def test_missing_extension_fails_loudly(tmp_path, monkeypatch):
    from vllm_xpu_kernels_b200 import _loader
    monkeypatch.setattr(_loader, "LIB_PATH", str(tmp_path / "nope.so"))
    monkeypatch.setattr(_loader, "_cdll", None)
    with pytest.raises(ImportError, match="no fallback"):
        _loader.cdll()


def test_shard_index_arithmetic():
    from vllm_xpu_kernels_b200 import shard
    for n, world in ((57344, 8), (10240, 8), (8192, 4), (1536, 8), (36864, 8), (300, 2), (129, 4)):
        cover = []
        for r in range(world):
            n0, n1 = shard.column_shard(n, r, world)
            assert n0 == n1 or (n0 % 128 == 0 and (n1 % 128 == 0 or n1 == n) and n0 < n1)
            cover += list(range(n0, n1))
        assert cover == list(range(n))
    for rows, world in ((8192, 8), (7, 2), (1, 4), (83, 8)):
        cover = []
        for r in range(world):
            r0, r1 = shard.row_shard(rows, r, world)
            cover += list(range(r0, r1))
        assert cover == list(range(rows))


WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from oracle import gemm as og, quant as oq, formats
from vllm_xpu_kernels_b200 import shard
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
torch.manual_seed(0)
E4 = torch.float8_e4m3fn
M, N, K = 24, 384, 256
x = torch.randn(M, K, dtype=torch.bfloat16) / 10
w = torch.randn(N, K, dtype=torch.bfloat16) / 10
# (1) token-row sharded activation quantisation: no collective, concatenation == unsharded
r0, r1 = shard.row_shard(M, rank, world)
q_loc, s_loc = oq.dynamic_per_token_scaled_fp8_quant(x[r0:r1], E4)
q_full, s_full = oq.dynamic_per_token_scaled_fp8_quant(x, E4)
assert torch.equal(q_loc.view(torch.uint8), q_full[r0:r1].view(torch.uint8)) and torch.equal(s_loc, s_full[r0:r1])
# (2) N-sharded per-channel FP8 GEMM + all-gather == unsharded GEMM
wq, ws = oq.dynamic_per_token_scaled_fp8_quant(w, E4)
w_loc, ws_loc, (n0, n1) = shard.shard_weight(wq, ws, rank, world, "per_channel")
out_loc = og.fp8_gemm_ref(q_full, w_loc.t(), torch.bfloat16, s_full, ws_loc)
full = shard.all_gather_columns(out_loc.float(), N).to(torch.bfloat16)
ref = og.fp8_gemm_ref(q_full, wq.t(), torch.bfloat16, s_full, ws)
assert torch.equal(full, ref), (full - ref).abs().max()
# (3) N-sharded MXFP4 GEMM ([N, K/32] scales shard with the rows of W)
sa, qa = formats.mx_quantize(x * 0.1, "mxfp4"); sw, qw = formats.mx_quantize(w * 0.1, "mxfp4")
qw_loc, sw_loc, _ = shard.shard_weight(qw.view(torch.uint8), sw.view(torch.uint8), rank, world, "mx")
out4 = og.fp4_gemm_ref(qa, qw_loc.t(), sa, sw_loc.view(torch.float8_e8m0fnu), torch.bfloat16)
full4 = shard.all_gather_columns(out4.float(), N).to(torch.bfloat16)
assert torch.equal(full4, og.fp4_gemm_ref(qa, qw.t(), sa, sw, torch.bfloat16))
# (4) block-scaled weights [K/128, N/128] split on whole blocks
wq2, ws2 = og.fp8_block_quant_2d(w, 128, 128)
_, ws2_loc, (n0, n1) = shard.shard_weight(wq2, ws2.t().contiguous(), rank, world, "block")
assert ws2_loc.shape == (K // 128, (n1 - n0) // 128)
dist.destroy_process_group()
print("ok", rank)
'''


def test_world_size_2_gloo_sharding(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29653", str(script), ROOT],
                       capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert r.stdout.count("ok") == 2
