#!/usr/bin/env python
"""Opcode histogram of the shipped kernels (cuobjdump -sass on libvxk_b200.so), grouped by kernel family:
the evidence that the GEMMs run on tcgen05 / TMA (UTCHMMA / UTCQMMA / UTCOMMA, UTCCP, UTMALDG, LDTM, UTCBAR)
and that the quantisation kernels use the packed fp32x2 / conversion instructions DESIGN.md describes.
    python tools/sass_histogram.py > profiles/sass_r02.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "vllm_xpu_kernels_b200", "_lib", "libvxk_b200.so")
FAMILIES = [
    ("gemm_tc2_kernel (CTA-pair FP8 / MXFP8 / MXFP4)", r"gemm_tc2_kernel"),
    ("gemm_tc2_blk_kernel (fp32-block FP8)", r"gemm_tc2_blk_kernel"),
    ("gemm_tc_kernel (single CTA)", r"gemm_tc_kernel"),
    ("gemm_grouped_kernel (MoE grouped GEMM)", r"gemm_grouped"),
    ("gemm_w8a16", r"w8a16|w16"),
    ("gemm_skinny (small-M FP8 / MXFP4)", r"skinny"),
    ("row_quant_kernel (a1, a6-a9)", r"row_quant_kernel"),
    ("flat_quant_kernel (a4, a5, a10-a12)", r"flat_quant_kernel"),
    ("cache ops (f3)", r"indexer|cp_gather|convert_fp8"),
]
KEY = ["UTCHMMA", "UTCQMMA", "UTCOMMA", "UTCIMMA", "UTCCP", "UTCBAR", "UTMALDG", "UBLKCP", "LDTM", "STTM", "SYNCS", "HMMA", "IMMA",
       "FFMA2", "FMUL2", "FADD2", "F2FP", "F2F", "DFMA", "MUFU", "LDG", "STG", "LDS", "STS", "SHFL", "BAR", "ACQBULK",
       "ELECT", "UCGABAR_ARV", "UCGABAR_WAIT", "USETMAXREG"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    fam_hist = {name: collections.Counter() for name, _ in FAMILIES}
    fam_kernels = {name: 0 for name, _ in FAMILIES}
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = None
            for name, rx in FAMILIES:
                if re.search(rx, m.group(1)):
                    cur = name
                    fam_kernels[name] += 1
                    break
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            fam_hist[cur][m.group(1)] += 1
    print("# SASS opcode histogram of vllm_xpu_kernels_b200/_lib/libvxk_b200.so (sm_100a), static instruction counts")
    print("# summed over every instantiation of a kernel family; produced by tools/sass_histogram.py")
    for name, _ in FAMILIES:
        h = fam_hist[name]
        total = sum(h.values())
        print(f"\n## {name}: {fam_kernels[name]} kernels, {total} instructions")
        print("   key opcodes: " + ", ".join(f"{k} {h[k]}" for k in KEY if h[k]))
        print("   top 12:      " + ", ".join(f"{k} {v}" for k, v in h.most_common(12)))


if __name__ == "__main__":
    sys.exit(main())
