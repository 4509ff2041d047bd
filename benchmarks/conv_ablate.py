#!/usr/bin/env python
"""Timing experiments on one convolution layer's forward kernel with the ablation switches of ddl_conv_set_debug
(1 = epilogue reads the accumulator but neither stages nor stores, 2 = no activation loads; results are garbage):

    python benchmarks/conv_ablate.py CIN HW COUT K STRIDE ABLATE [ABLATE ...]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from databricks_distributed_deep_learning_b200.ops import _lib as L  # noqa: E402


def main():
    lib = L.lib()
    bf, cl = torch.bfloat16, torch.channels_last
    cin, hw, cout, k, s = (int(v) for v in sys.argv[1:6])
    N, pad = 256, k // 2
    oh = (hw + 2 * pad - k) // s + 1
    x = torch.randn(N, cin, hw, hw, device="cuda").to(bf).contiguous(memory_format=cl)
    w = torch.randn(cout, cin, k, k, device="cuda").to(bf).contiguous(memory_format=cl)
    y = torch.empty(N, cout, oh, oh, device="cuda", dtype=bf).contiguous(memory_format=cl)
    geom = (N, hw, hw, cin, cout, k, k, s, pad)
    part = torch.empty(lib.ddl_conv_fwd_groups(*geom) * 2 * cout, dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for ab in [int(v) for v in sys.argv[6:]]:
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
        # one more run with the cycle counters on (im2col / tiled data kernel only)
        dbg = torch.zeros(296, 8, dtype=torch.int64, device="cuda")
        L.call("ddl_conv_set_debug", dbg.data_ptr(), ab)
        L.call("ddl_conv_fwd", x.data_ptr(), w.data_ptr(), y.data_ptr(), part.data_ptr(), *geom, L.stream_ptr())
        torch.cuda.synchronize()
        L.call("ddl_conv_set_debug", None, ab)
        d = dbg.cpu()
        d = d[d[:, 0] > 0].float()
        extra = ""
        if len(d):
            m = [int(v) for v in d.median(0).values.tolist()]
            extra = (f" | cycles (median over {len(d)} CTAs): MMA warp total {m[0]}, waiting for a free accumulator {m[1]}, for operands {m[2]}; "
                     f"epilogue warp total {m[3]}, waiting for a finished accumulator {m[4]}")
        print(f"{cin}x{hw}^2 -> {cout} k{k}/s{s} ablate {ab}: {ts[len(ts) // 2]:7.1f} us{extra}", flush=True)
    L.call("ddl_conv_set_debug", None, 0)


if __name__ == "__main__":
    main()
