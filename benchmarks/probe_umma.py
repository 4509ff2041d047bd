#!/usr/bin/env python
"""Runs csrc/probe_umma.cu: which row shifts of a swizzled shared-memory operand does tcgen05.mma read correctly?"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from databricks_distributed_deep_learning_b200.ops import _lib as L  # noqa: E402


def main():
    torch.manual_seed(0)
    X = torch.randn(256, 64, device="cuda").to(torch.bfloat16)
    B = torch.eye(64, device="cuda").to(torch.bfloat16)
    D = torch.empty(128, 64, dtype=torch.float32, device="cuda")
    for mn in (0, 1):
        for variant in (0, 1):
            ok = []
            for shift in list(range(0, 17)) + [23, 57, 58, 59, 64]:
                D.fill_(float("nan"))
                L.call("ddl_probe_umma_shift", X.data_ptr(), B.data_ptr(), D.data_ptr(), shift, variant, mn, L.stream_ptr())
                torch.cuda.synchronize()
                if mn == 0:
                    want = X[shift:shift + 128].float()
                    good = torch.equal(D, want)
                else:
                    want = X[shift:shift + 64].float().t()  # D[m][n] = X[shift + n][m]
                    good = torch.equal(D[:64], want)
                ok.append((shift, bool(good)))
            print(f"mn={mn} variant={variant}: " + " ".join(f"{s}:{'ok' if g else 'BAD'}" for s, g in ok), flush=True)


if __name__ == "__main__":
    main()
