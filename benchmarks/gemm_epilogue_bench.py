#!/usr/bin/env python
"""Cost of the fused GEMM epilogues at the BERT-base FFN shape (device-timed, L2 flushed)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from databricks_distributed_deep_learning_b200.ops import gemm as G  # noqa: E402
from databricks_distributed_deep_learning_b200.utils.timing import time_kernel  # noqa: E402


def main():
    bf = torch.bfloat16
    rows = []
    for (M, N, K) in ((16384, 3072, 768), (16384, 768, 3072), (8192, 4096, 1024)):
        a = torch.randn(M, K, device="cuda").to(bf)
        b = torch.randn(N, K, device="cuda").to(bf)
        bias = torch.randn(N, device="cuda").to(bf)
        res = torch.randn(M, N, device="cuda").to(bf)
        pre = torch.randn(M, N, device="cuda").to(bf)
        out = torch.empty(M, N, device="cuda", dtype=bf)
        flops = 2.0 * M * N * K
        cases = {
            "none": dict(),
            "bias": dict(bias=bias),
            "bias+residual": dict(bias=bias, residual=res),
            "bias+gelu+aux": dict(bias=bias, aux=pre, epilogue=G.EPI_GELU),
            "bias+gelu": dict(bias=bias, epilogue=G.EPI_GELU),
            "dgelu": dict(aux=pre, epilogue=G.EPI_DGELU),
        }
        r = {"M": M, "N": N, "K": K}
        for name, kw in cases.items():
            t = time_kernel(lambda: G.gemm_nt(a, b, M, N, K, out=out, **kw), warmup=3, iters=15)
            r[name] = round(flops / t["median_ms"] / 1e9, 1)
        rows.append(r)
        print(json.dumps(r), flush=True)
    with open(os.path.join(ROOT, "gpurun_out", "gemm_epilogue_bench.json"), "w") as f:
        json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
