"""In-tree build of the two shared objects of the package (sm_100a only).

  _lib/libvxk_b200.so     nvcc only, extern "C" ABI of include/vxk_b200.h, no torch
  _lib/_vxk_torch_ops.so  g++ + torch headers: TORCH_LIBRARY(_C / _xpu_C) shim -> C ABI

nvcc cross-compiles without a GPU, so this runs in the build container; the
.so files travel to the GPU box with the repository snapshot.
Usage:  python -m vllm_xpu_kernels_b200.build [--force]
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "_lib")
OBJDIR = os.path.join(ROOT, "build", "obj")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_COMMON = ARCH + ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
                      "--expt-relaxed-constexpr", "-Xptxas", "-v"]
# per-file extra flags.  Quant kernels: no implicit FMA contraction (bit-exactness).
CU_SOURCES = {
    "vxk_api.cu": [],
    "quant_row.cu": ["-fmad=false"],
    "quant_row_f32.cu": ["-fmad=false"],
    "quant_row_f16.cu": ["-fmad=false"],
    "quant_row_bf16.cu": ["-fmad=false"],
    "quant_flat.cu": ["-fmad=false"],
    "quant_cache.cu": ["-fmad=false"],
    "gemm_fallback.cu": [],
    "gemm_tc.cu": [],
    "gemm_w8a16.cu": [],
    "gemm_grouped.cu": [],
}
# every header under csrc/ feeds every object's digest (a header missing here means stale objects)
HEADERS = sorted(f for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))) + [os.path.join(ROOT, "include", "vxk_b200.h")]


def _digest(paths, extra=""):
    h = hashlib.sha256(extra.encode())
    for p in paths:
        p = p if os.path.isabs(p) else os.path.join(CSRC, p)
        if os.path.exists(p):
            with open(p, "rb") as f:
                h.update(f.read())
    return h.hexdigest()[:16]


def _run(cmd, log):
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError(f"build step failed: {' '.join(cmd[:3])} ... (log: {log})")


def _compile_cu(name, flags, force):
    src = os.path.join(CSRC, name)
    tag = _digest([name] + HEADERS, " ".join(NVCC_COMMON + flags))
    obj = os.path.join(OBJDIR, f"{name}.{tag}.o")
    if force or not os.path.exists(obj):
        for old in os.listdir(OBJDIR):          # drop stale objects of this source
            if old.startswith(name + ".") and not old.startswith(f"{name}.{tag}."):
                os.remove(os.path.join(OBJDIR, old))
        _run([NVCC] + NVCC_COMMON + flags + ["-c", src, "-o", obj], obj + ".log")
    return obj


def build(force: bool = False, verbose: bool = False) -> dict:
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    with ThreadPoolExecutor(max_workers=len(CU_SOURCES)) as ex:
        objs = list(ex.map(lambda kv: _compile_cu(kv[0], kv[1], force), CU_SOURCES.items()))
    lib = os.path.join(LIBDIR, "libvxk_b200.so")
    stamp = os.path.join(LIBDIR, ".libvxk.stamp")
    want = _digest(objs)
    if force or not os.path.exists(lib) or not os.path.exists(stamp) or open(stamp).read() != want:
        _run([NVCC] + ARCH + ["-shared", "-o", lib] + objs + ["-Xlinker", "-soname,libvxk_b200.so"],
             os.path.join(OBJDIR, "link_libvxk.log"))
        open(stamp, "w").write(want)

    # torch shim
    import torch
    from torch.utils import cpp_extension as ce
    shim_src = os.path.join(CSRC, "torch_shim.cpp")
    shim = os.path.join(LIBDIR, "_vxk_torch_ops.so")
    sstamp = os.path.join(LIBDIR, ".shim.stamp")
    swant = _digest([shim_src, os.path.join(ROOT, "include", "vxk_b200.h")], torch.__version__)
    if force or not os.path.exists(shim) or not os.path.exists(sstamp) or open(sstamp).read() != swant:
        inc = []
        for p in ce.include_paths("cuda"):
            inc += ["-isystem", p]
        tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
        abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
        cmd = (["g++", "-O2", "-std=c++17", "-fPIC", "-shared", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
                shim_src, "-o", shim] + inc +
               ["-L" + tlib, "-L" + LIBDIR, "-lvxk_b200", "-lc10", "-lc10_cuda", "-ltorch_cpu",
                "-ltorch_cuda", "-ltorch", "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + tlib,
                "-Wl,--no-as-needed"])
        _run(cmd, os.path.join(OBJDIR, "shim.log"))
        open(sstamp, "w").write(swant)
    if verbose:
        print("built", lib, shim)
    return {"lib": lib, "shim": shim}


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
