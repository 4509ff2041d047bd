#!/usr/bin/env python
"""tcgen05 GEMM throughput at the BERT shapes vs torch.matmul (cuBLAS), device-timed.

Roofline denominator: MEASURED_PEAKS.json `bf16_tflops` (cuBLAS 8192^3 burst)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from databricks_distributed_deep_learning_b200.ops import gemm as G  # noqa: E402
from databricks_distributed_deep_learning_b200.utils.timing import load_measured_peaks, time_kernel  # noqa: E402


def main():
    peaks = load_measured_peaks(ROOT)
    tok = int(os.environ.get("TOKENS", 16384))
    shapes = [("qkv", tok, 2304, 768), ("attn_out", tok, 768, 768), ("ffn1", tok, 3072, 768), ("ffn2", tok, 768, 3072),
              ("large_qkv", 8192, 3072, 1024), ("large_ffn1", 8192, 4096, 1024), ("large_ffn2", 8192, 1024, 4096),
              ("square8k", 8192, 8192, 8192)]
    rows = []
    for name, M, N, K in shapes:
        a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        b = torch.randn(N, K, device="cuda").to(torch.bfloat16)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        flops = 2.0 * M * N * K
        res = {}
        for bn in (128, 256):
            t = time_kernel(lambda: G.gemm_nt(a, b, M, N, K, out=out, block_n=bn), warmup=3, iters=20)
            res[f"ours_bn{bn}_tflops"] = flops / t["median_ms"] / 1e9
        res["ours_2sm_tflops"] = res["ours_bn256_tflops"]  # mode 0 = CTA pairs (cta_group::2)
        for mode, key in ((1, "ours_1cta_tflops"), (2, "ours_mcast_tflops")):
            G.set_cluster_mode(mode)
            t = time_kernel(lambda: G.gemm_nt(a, b, M, N, K, out=out, block_n=256), warmup=3, iters=20)
            res[key] = flops / t["median_ms"] / 1e9
        G.set_cluster_mode(0)
        t = time_kernel(lambda: torch.matmul(a, b.t(), out=out), warmup=3, iters=20)
        res["cublas_tflops"] = flops / t["median_ms"] / 1e9
        # backward-shaped calls
        dy = torch.randn(M, N, device="cuda").to(torch.bfloat16)
        t = time_kernel(lambda: G.gemm_nt(dy, b, M, K, N, b_mn=True), warmup=3, iters=10)
        res["ours_dgrad_tflops"] = flops / t["median_ms"] / 1e9
        sk = G.pick_split_k(N, K, M)
        t = time_kernel(lambda: G.gemm_nt(dy, a, N, K, M, a_mn=True, b_mn=True, split_k=sk), warmup=3, iters=10)
        res["ours_wgrad_tflops"] = flops / t["median_ms"] / 1e9
        res["wgrad_split_k"] = sk
        best = max(res["ours_bn128_tflops"], res["ours_bn256_tflops"], res.get("ours_2sm_tflops", 0.0))
        res.update(name=name, M=M, N=N, K=K, frac_of_measured_peak=best / peaks["bf16_tflops"], vs_cublas=best / res["cublas_tflops"])
        rows.append(res)
        print(json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in res.items()}), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "gemm_bench.json"), "w") as f:
        json.dump({"peaks": peaks, "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
