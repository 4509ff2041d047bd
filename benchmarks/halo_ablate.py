#!/usr/bin/env python
"""Timing experiments on the halo-tile kernel of the 64 -> 64 channel 3x3 layer (weights resident, so every ablation of
ddl_conv_set_debug terminates): python benchmarks/halo_ablate.py MT ABLATE [ABLATE ...]   (results of ablated runs are garbage)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from databricks_distributed_deep_learning_b200.ops import _lib as L  # noqa: E402


def main():
    lib = L.lib()
    bf, cl = torch.bfloat16, torch.channels_last
    mt = int(sys.argv[1])
    L.call("ddl_conv_set_halo", 1, mt, 14)
    N, cin, hw, cout = 256, 64, 56, 64
    x = torch.randn(N, cin, hw, hw, device="cuda").to(bf).contiguous(memory_format=cl)
    w = torch.randn(cout, cin, 3, 3, device="cuda").to(bf).contiguous(memory_format=cl)
    y = torch.empty(N, cout, hw, hw, device="cuda", dtype=bf).contiguous(memory_format=cl)
    geom = (N, hw, hw, cin, cout, 3, 3, 1, 1)
    part = torch.empty(lib.ddl_conv_fwd_groups(*geom) * 2 * cout, dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for ab in [int(v) for v in sys.argv[2:]]:
        L.call("ddl_conv_set_debug", None, ab)
        ts = []
        for it in range(8):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            L.call("ddl_conv_fwd", x.data_ptr(), w.data_ptr(), y.data_ptr(), part.data_ptr(), *geom, L.stream_ptr())
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
        ts = sorted(ts[3:])
        print(f"mt {mt} ablate {ab:3d}: {ts[len(ts) // 2]:7.1f} us", flush=True)
    L.call("ddl_conv_set_debug", None, 0)


if __name__ == "__main__":
    main()
