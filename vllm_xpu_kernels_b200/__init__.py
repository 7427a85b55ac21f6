"""vllm_xpu_kernels_b200 -- B200 (sm_100a) implementation of the quantized-linear
hot path of vllm-xpu-kernels behind the reference's own operator surface.

    import vllm_xpu_kernels_b200._C       # registers torch.ops._C.* (CUDA key)
    import vllm_xpu_kernels_b200._xpu_C   # registers torch.ops._xpu_C.*

mirrors ``import vllm_xpu_kernels._C`` / ``._xpu_C`` of the reference
(/root/reference/tests/register_ops.py:6-8).  ``ops`` holds the thin Python
wrappers the reference keeps in tests/register_ops.py and tests/ops/*.py.

There is no CPU or library fallback: importing ``_C`` / ``_xpu_C`` raises if the
compiled extension is missing (build it with ``python -m vllm_xpu_kernels_b200.build``).
"""
__version__ = "0.1.0"
