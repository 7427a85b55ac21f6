"""``import vllm_xpu_kernels_b200._xpu_C`` registers ``torch.ops._xpu_C.fp8_gemm`` /
``fp4_gemm`` for CUDA tensors (reference: ``import vllm_xpu_kernels._xpu_C``,
csrc/xpu/torch_bindings.cpp)."""
from ._loader import load_ops

load_ops()
