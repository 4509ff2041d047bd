#!/usr/bin/env python
"""csrc/probe_umma.cu drain probe: bytes per SM cycle that `nwarps` epilogue warps get out of TMEM with tcgen05.ld
(32 lanes x 32 columns = 4 KB per instruction), alone and while another warp keeps the tensor pipe busy with 128 x 256 x 16 MMAs."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from databricks_distributed_deep_learning_b200.ops import _lib as L  # noqa: E402

out = torch.zeros(148, 2, dtype=torch.int64, device="cuda")
iters = 512
for nw in (4, 8, 16):
    for mma in (0, 1):
        for _ in range(2):
            out.zero_()
            L.call("ddl_probe_tmem_drain", out.data_ptr(), nw, iters, mma, 148, L.stream_ptr())
            torch.cuda.synchronize()
        o = out.float().median(0).values
        cyc, mmas = o[0].item(), o[1].item()
        print(f"{nw:2d} reader warps, MMAs {'on ' if mma else 'off'}: {cyc / iters:7.1f} cycles per 4 KB load per warp = "
              f"{nw * 4096 * iters / cyc:6.1f} B/cycle per SM" + (f"; {mmas:.0f} MMAs issued meanwhile = {cyc / max(mmas, 1):.0f} cycles per MMA (nominal 128)" if mma else ""),
              flush=True)
