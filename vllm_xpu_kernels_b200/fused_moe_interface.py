"""Grouped-GEMM entry points of the reference's ``vllm_xpu_kernels/fused_moe_interface.py`` (:63-94) on B200.

Only the grouped GEMM is mirrored here: same function names and argument order, the op
behind them is ``torch.ops._xpu_C.cutlass_grouped_gemm_interface`` (reference schema,
csrc/xpu/torch_bindings.cpp:41-53) -> C ABI ``vxk_moe_grouped_gemm``.  Recipes served: "bf16", "fp8", "mxfp4"
(``_get_recipe``, :32-45); "int4" is refused by the op (INT4 GEMMs are out of scope), "mxfp8" /
"fp8block" never reach the grouped GEMM in the reference either (``_should_use_ref_fused_moe``, :26-29).
"""
import torch

from . import _xpu_C  # noqa: F401  (registers torch.ops._xpu_C.*)


def cutlass_grouped_gemm_xe2(input_A, input_B, scales, bias, output, num_rows_per_expert, n, k, num_experts):
    """fused_moe_interface.py:82-94.  num_rows_per_expert: int32 device tensor [num_experts]."""
    torch.ops._xpu_C.cutlass_grouped_gemm_interface(
        ptr_A=input_A, ptr_B=input_B, ptr_scales=scales, ptr_bias=bias, ptr_D=output,
        rows_per_expert=num_rows_per_expert, N=n, K=k, num_experts=num_experts)


def cutlass_grouped_gemm(input_A, input_B, bias, output, expert_token_count, n, k, num_experts):
    """fused_moe_interface.py:63-79: host list of rows per expert, no scales."""
    rows = torch.tensor(expert_token_count, dtype=torch.int32, device=input_A.device)
    cutlass_grouped_gemm_xe2(input_A, input_B, None, bias, output, rows, n, k, num_experts)
