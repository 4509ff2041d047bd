#!/usr/bin/env python
"""Where does the CTA-pair convolution kernel wait?  Runs the forward kernel of a few layers with the wait-cycle counters
(ddl_conv_set_debug) on and prints, per layer, the median over CTAs of: producer total / wait-A-slot / wait-B-slot and
MMA-issuer total / wait-A-data / wait-B-data / wait-accumulator (SM cycles)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from databricks_distributed_deep_learning_b200.ops import _lib as L  # noqa: E402


def main():
    lib = L.lib()
    bf, cl = torch.bfloat16, torch.channels_last
    mode = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    ablate = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    print(f"pair mode {mode}, ablate {ablate}")
    L.call("ddl_conv_set_tuning", 0, mode)
    dbg = torch.zeros(296, 8, dtype=torch.int64, device="cuda")
    for cin, hw, cout, k, s in [(128, 28, 128, 3, 1), (256, 14, 256, 3, 1), (1024, 14, 512, 1, 1), (256, 14, 1024, 1, 1), (512, 7, 512, 3, 1)]:
        N, pad = 256, k // 2
        x = torch.randn(N, cin, hw, hw, device="cuda").to(bf).contiguous(memory_format=cl)
        w = torch.randn(cout, cin, k, k, device="cuda").to(bf).contiguous(memory_format=cl)
        oh = (hw + 2 * pad - k) // s + 1
        y = torch.empty(N, cout, oh, oh, device="cuda", dtype=bf).contiguous(memory_format=cl)
        geom = (N, hw, hw, cin, cout, k, k, s, pad)
        part = torch.empty(lib.ddl_conv_fwd_groups(*geom) * 2 * cout, dtype=torch.float32, device="cuda")
        for it in range(3):
            dbg.zero_()
            L.call("ddl_conv_set_debug", dbg.data_ptr() if it == 2 else None, ablate)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            L.call("ddl_conv_fwd", x.data_ptr(), w.data_ptr(), y.data_ptr(), part.data_ptr(), *geom, L.stream_ptr())
            b.record()
            torch.cuda.synchronize()
        L.call("ddl_conv_set_debug", None, 0)
        d = dbg.cpu()
        prod = d[d[:, 0] > 0][:, 0:3].float()
        mma = d[d[:, 3] > 0][:, [3, 4, 5, 6, 7, 0]].float()
        med = lambda t: [int(v) for v in t.median(0).values.tolist()] if len(t) else []
        print(f"{cin}x{hw}^2->{cout} k{k}: {a.elapsed_time(b)*1e3:.1f} us | producer total/waitA/waitB {med(prod)} | "
              f"mma total/waitA/waitB/waitAcc/issue/commit {med(mma)}", flush=True)


if __name__ == "__main__":
    main()
