"""``import vllm_xpu_kernels_b200._C`` registers ``torch.ops._C.*`` for CUDA tensors
(reference: ``import vllm_xpu_kernels._C``, csrc/torch_bindings.cpp)."""
from ._loader import load_ops

load_ops()
