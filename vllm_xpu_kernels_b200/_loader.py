"""Loads libvxk_b200.so (C ABI, ctypes handle) and the torch operator shim."""
from __future__ import annotations

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libvxk_b200.so")
SHIM_PATH = os.path.join(_HERE, "_lib", "_vxk_torch_ops.so")

ABI_VERSION = 8   # == VXK_ABI_VERSION of include/vxk_b200.h

_lock = threading.Lock()
_cdll = None
_ops_loaded = False


class ExtensionMissing(ImportError):
    pass


def cdll() -> ctypes.CDLL:
    """ctypes handle on the C ABI (include/vxk_b200.h)."""
    global _cdll
    with _lock:
        if _cdll is None:
            if not os.path.exists(LIB_PATH):
                raise ExtensionMissing(
                    f"{LIB_PATH} not found: the CUDA extension is not built "
                    "(run `python -m vllm_xpu_kernels_b200.build`). There is no fallback path.")
            lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
            lib.vxk_last_error.restype = ctypes.c_char_p
            lib.vxk_launch_count.restype = ctypes.c_uint64
            lib.vxk_abi_version.restype = ctypes.c_int
            if lib.vxk_abi_version() != ABI_VERSION:
                raise ExtensionMissing(
                    f"{LIB_PATH} exports C ABI version {lib.vxk_abi_version()}, this package needs {ABI_VERSION}: "
                    "stale build (run `python -m vllm_xpu_kernels_b200.build --force`).")
            _cdll = lib
    return _cdll


def load_ops() -> None:
    """Registers torch.ops._C.* and torch.ops._xpu_C.* (CUDA dispatch key)."""
    global _ops_loaded
    cdll()
    with _lock:
        if _ops_loaded:
            return
        if not os.path.exists(SHIM_PATH):
            raise ExtensionMissing(
                f"{SHIM_PATH} not found: the torch operator shim is not built "
                "(run `python -m vllm_xpu_kernels_b200.build`). There is no fallback path.")
        import torch
        torch.ops.load_library(SHIM_PATH)
        _ops_loaded = True


def launch_count() -> int:
    return int(cdll().vxk_launch_count())
