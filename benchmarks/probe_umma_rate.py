#!/usr/bin/env python
"""csrc/probe_umma.cu rate probe: SM cycles per tcgen05.mma (M = 128, K = 16, operands in shared memory) versus N and
versus what surrounds the instruction in a k-block loop (commit, fence, barrier waits)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from databricks_distributed_deep_learning_b200.ops import _lib as L  # noqa: E402

NAMES = {0: "one thread, divergent region", 1: "uniform warp + elect", 3: "+ commit per 4 MMAs", 4: "+ fence::after_thread_sync per group",
         5: "+ wait for the commit of 6 groups ago", 6: "+ wait for the commit of the previous group"}
out = torch.zeros(148, 2, dtype=torch.int64, device="cuda")
for mode in (0, 1, 3, 4, 5, 6):
    for n in (32, 64, 128, 256):
        for it in range(2):
            L.call("ddl_probe_umma_rate", out.data_ptr(), n, 2048, 148, mode, 2, L.stream_ptr())
            torch.cuda.synchronize()
        o = out.float().median(0).values
        print(f"mode {mode} ({NAMES[mode]:45s}) N {n:3d}: {o[1].item()/512:7.1f} cycles per group of 4 MMAs (nominal {n * 2})", flush=True)
