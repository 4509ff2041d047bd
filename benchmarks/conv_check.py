#!/usr/bin/env python
"""Development check of the implicit-GEMM convolutions: every ResNet-50 shape, all three kernels, errors against fp32
autograd, no stop at the first failure (the pytest version is tests/test_gpu_conv.py).  --time also times them at the
training batch against the library convolution.

    python benchmarks/conv_check.py [--batch 3] [--time 256]
"""
import argparse
import os
import sys
import traceback

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=3)
    ap.add_argument("--only", type=int, default=-1)
    ap.add_argument("--im2col", type=int, default=1)
    ap.add_argument("--halo", default="1,0,28", help="halo-tile kernel for 3x3 / stride 1: on,mt,min_w")
    ap.add_argument("--stay", type=int, default=1, help="resident weight panels in the data kernel")
    ap.add_argument("--only3x3", type=int, default=0, help="only the 3x3 / stride-1 layers")
    ap.add_argument("--pair", type=int, default=0, help="0 = one-CTA data kernel, 1 = CTA pairs + halo tiles, 2 = CTA pairs")
    args = ap.parse_args()
    from test_gpu_conv import RESNET50_CONVS, conv_case

    from databricks_distributed_deep_learning_b200.ops import _lib as L

    L.call("ddl_conv_set_tuning", 0, args.pair)
    L.call("ddl_conv_set_im2col", args.im2col)
    L.call("ddl_conv_set_stay", args.stay)
    L.call("ddl_conv_set_halo", *[int(v) for v in args.halo.split(",")])
    print(f"pair mode {args.pair}", flush=True)
    bad = 0
    for i, (cin, hw, cout, k, s) in enumerate(RESNET50_CONVS):
        if args.only >= 0 and i != args.only:
            continue
        if args.only3x3 and not (k == 3 and s == 1):
            continue
        name = f"{cin}x{hw}^2 -> {cout} k{k}/s{s}"
        try:
            e = conv_case(args.batch, cin, hw, cout, k, s)
            ok = all(e[q] < 8e-3 for q in ("fwd", "dgrad", "wgrad")) and e["sum"] < 2e-3 and e["sumsq"] < 2e-3 and e.get("ws_dgrad", 0) < 8e-3
            bad += 0 if ok else 1
            print(f"{i:2d} {name:30s} " + " ".join(f"{k2}={v:.2e}" for k2, v in e.items()) + ("" if ok else "   <-- FAIL"), flush=True)
        except Exception:
            bad += 1
            print(f"{i:2d} {name:30s} EXCEPTION\n{traceback.format_exc()}", flush=True)
            try:
                torch.cuda.synchronize()
            except Exception:
                print("CUDA context is broken; stopping", flush=True)
                break
    print(f"conv_check: {bad} failing shapes")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
