#!/usr/bin/env python
"""SASS evidence for profiles/: per-kernel mnemonic histograms of libb200ddl.so plus the full
listing of the flagship kernels (runs on the CPU box: `cuobjdump -sass`).

What to look for (B200_PROFILING.md "What proves a Blackwell-native kernel"):
  tcgen05.mma -> UTCHMMA, tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG, commit -> UTCBAR,
  mbarrier -> SYNCS, multimem.* -> MULTIMEM (shown as LDGMC / STG...MC / REDG variants), red.global.add -> RED.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "databricks_distributed_deep_learning_b200", "ops", "libb200ddl.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEY = re.compile(r"^(UTC\w+|UTMA\w+|LDTM|STTM|SYNCS|UBLKCP|MULTIMEM|LDGMC|RED|ATOM|HMMA|LDGSTS|LDG|STG|LDS|STS|BAR|MEMBAR|ERRBAR|UTCBAR|FFMA|MUFU|SHFL|CCTL)")
FULL = ("gemm_bf16_kernelILi256ELb0ELb0ELi4ELi2", "attn_fwd_kernelILi1", "attn_bwd_kernel", "allreduce_nvls_kernelILb1",
        "allreduce_twoshot_kernelI13__nv_bfloat16S1_Li8", "gemm2sm_bf16_kernelILb0ELb0", "stem_fwd_kernel", "stem_wgrad_kernel", "conv1x1_fwd_kernel", "attn_fwd_stream_kernel",
        "conv_halo_kernelILb0ELi2ELb1", "conv_halo_kernelILb1ELi2ELb0", "conv_data_kernelILb0", "conv_data_kernelILb1", "conv_wgrad_kernel")


def main():
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = re.split(r"\n\s*Function : ", txt)[1:]
    lines = ["# SASS mnemonic histogram per kernel of libb200ddl.so (sm_100a)", f"# kernels: {len(kernels)}", ""]
    for k in kernels:
        name = k.split("\n", 1)[0].strip()
        ops = collections.Counter()
        total = 0
        for m in re.finditer(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", k):
            total += 1
            op = m.group(1)
            if KEY.match(op):
                ops[op.split(".")[0] if op.startswith(("FFMA", "LDS", "STS", "SHFL", "MUFU")) else op] += 1
        short = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()[:160]
        flag = " ".join(f"{o}:{c}" for o, c in sorted(ops.items(), key=lambda t: -t[1]) if re.match(r"UTC|UTMA|LDTM|STTM|SYNCS|MULTIMEM|LDGMC|RED|UBLKCP|HMMA", o))
        lines.append(f"{short}\n    instructions={total}  blackwell/async ops: {flag or '-'}")
        for tag in FULL:
            if tag in name:
                with open(os.path.join(OUT, f"{tag}.sass"), "w") as f:
                    # instruction text only (drop the encoding words) to keep the listing small
                    body = [re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", ln).rstrip() for ln in k.split("\n")]
                    body = [ln for ln in body if ln.strip() and not re.fullmatch(r"\s*/\* 0x[0-9a-f]{16} \*/", ln)]
                    f.write("Function : " + "\n".join(body) + "\n")
    with open(os.path.join(OUT, "histogram.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines[:12]))


if __name__ == "__main__":
    sys.exit(main())
