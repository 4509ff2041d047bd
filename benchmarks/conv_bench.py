#!/usr/bin/env python
"""Per-layer timing of the ResNet-50 convolutions at per-GPU batch 256:
library convolution (cuDNN through ATen, bf16 NHWC, benchmark mode) forward / dgrad / wgrad, against the
memory and tensor-core floors and the implicit-GEMM tcgen05 kernels of this repo (csrc/conv_igemm.cu: forward with
the BatchNorm-statistics epilogue, dgrad, wgrad; csrc/conv1x1.cu: weight-stationary forward of the skinny 1x1 layers).

    python benchmarks/conv_bench.py [--batch 256] [--json gpurun_out/conv_bench.json]

CUDA events, 3 warm-up + 10 timed runs per point, L2 flushed between runs by a 256 MB memset.
"""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# (count, Cin, HW_in, Cout, k, stride)
LAYERS = [
    (1, 3, 224, 64, 7, 2), (1, 64, 56, 64, 1, 1), (3, 64, 56, 64, 3, 1), (4, 64, 56, 256, 1, 1), (2, 256, 56, 64, 1, 1),
    (1, 256, 56, 128, 1, 1), (1, 128, 56, 128, 3, 2), (4, 128, 28, 512, 1, 1), (1, 256, 56, 512, 1, 2), (3, 512, 28, 128, 1, 1),
    (3, 128, 28, 128, 3, 1), (1, 512, 28, 256, 1, 1), (1, 256, 28, 256, 3, 2), (6, 256, 14, 1024, 1, 1), (1, 512, 28, 1024, 1, 2),
    (5, 1024, 14, 256, 1, 1), (5, 256, 14, 256, 3, 1), (1, 1024, 14, 512, 1, 1), (1, 512, 14, 512, 3, 2), (3, 512, 7, 2048, 1, 1),
    (1, 1024, 14, 2048, 1, 2), (2, 2048, 7, 512, 1, 1), (2, 512, 7, 512, 3, 1),
]


def timeit(fn, flush, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--json", default=os.path.join(ROOT, "gpurun_out", "conv_bench.json"))
    ap.add_argument("--im2col", type=int, default=1)
    ap.add_argument("--halo", default="1,0,28", help="halo-tile kernel for 3x3 / stride 1: on,mt,min_w")
    ap.add_argument("--ablate", type=int, default=0, help="timing experiments (results are garbage): see ddl_conv_set_debug")
    ap.add_argument("--no-lib", type=int, default=0, help="skip the library convolutions")
    ap.add_argument("--stay", type=int, default=1, help="resident weight panels in the data kernel")
    ap.add_argument("--only3x3", type=int, default=0, help="only the 3x3 / stride-1 layers")
    ap.add_argument("--pair", type=int, default=0, help="forward / data-gradient kernel: 0 one CTA, 1 CTA pairs + halo tiles, 2 CTA pairs")
    ap.add_argument("--wgrad-rows", type=int, default=0)
    args = ap.parse_args()
    from databricks_distributed_deep_learning_b200.ops import _lib as L0

    L0.call("ddl_conv_set_tuning", args.wgrad_rows, args.pair)
    L0.call("ddl_conv_set_im2col", args.im2col)
    L0.call("ddl_conv_set_stay", args.stay)
    L0.call("ddl_conv_set_halo", *[int(v) for v in args.halo.split(",")])
    L0.call("ddl_conv_set_debug", None, args.ablate)
    torch.backends.cudnn.benchmark = True
    dev = torch.device("cuda")
    bf = torch.bfloat16
    peaks = {"hbm_gbs": 6551.4, "bf16_tflops": 1397.3}
    try:
        peaks.update({k: v for k, v in json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).items() if k in ("hbm_gbs",)})
        peaks["bf16_tflops"] = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]
    except Exception:
        pass
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    N = args.batch
    rows = []
    tot = {"cudnn_fwd": 0.0, "cudnn_dgrad": 0.0, "cudnn_wgrad": 0.0, "floor": 0.0}
    print(f"{'layer':32s} {'x':>2s} {'fwd':>8s} {'dgrad':>8s} {'wgrad':>8s} {'floor':>7s} | {'n.fwd':>8s} {'n.dgrad':>8s} {'n.wgrad':>8s} {'ws.fwd /dgrad':>16s}   (us per call; n.* = csrc/conv_igemm.cu, n.fwd / ws.fwd include the BN statistics)")
    for cnt, cin, hw, cout, k, s in LAYERS:
        if args.only3x3 and not (k == 3 and s == 1):
            continue
        pad = k // 2
        x = torch.randn(N, cin, hw, hw, device=dev).to(bf).contiguous(memory_format=torch.channels_last)
        w = torch.randn(cout, cin, k, k, device=dev).to(bf).contiguous(memory_format=torch.channels_last)
        y = F.conv2d(x, w, None, s, pad)
        dy = torch.randn_like(y).contiguous(memory_format=torch.channels_last)
        ho = y.shape[2]
        if args.no_lib:
            t_f = t_d = t_w = 0.0
        else:
            t_f = timeit(lambda: F.conv2d(x, w, None, s, pad), flush)
        t_d = 0.0 if args.no_lib else timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, (s, s), (pad, pad), (1, 1), False, (0, 0), 1, (True, False, False)), flush) if cin > 3 else 0.0
        t_w = 0.0 if args.no_lib else timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, (s, s), (pad, pad), (1, 1), False, (0, 0), 1, (False, True, False)), flush)
        flops = 2.0 * N * ho * ho * cout * cin * k * k
        bytes_io = 2.0 * (x.numel() + y.numel())
        floor = max(flops / (peaks["bf16_tflops"] * 1e12), bytes_io / (peaks["hbm_gbs"] * 1e9)) * 1e6
        row = {"cin": cin, "hw": hw, "cout": cout, "k": k, "stride": s, "count": cnt, "cudnn_fwd_us": t_f, "cudnn_dgrad_us": t_d,
               "cudnn_wgrad_us": t_w, "floor_us": floor, "gflop": flops / 1e9, "mbytes": bytes_io / 1e6}
        g = ["", "", "", ""]
        from databricks_distributed_deep_learning_b200.ops import _lib as L

        lib = L.lib()
        geom = (N, hw, hw, cin, cout, k, k, s, pad)
        if lib.ddl_conv_supported(*geom):
            st = L.stream_ptr(dev)
            yy = torch.empty_like(y)
            part = torch.empty(lib.ddl_conv_fwd_groups(*geom) * 2 * cout, dtype=torch.float32, device=dev)
            dx = torch.empty_like(x)
            dw = torch.empty_like(w)
            scr = torch.empty(max(lib.ddl_conv_wgrad_scratch_bytes(*geom), 16), dtype=torch.uint8, device=dev)
            row["native_fwd_us"] = timeit(lambda: L.call("ddl_conv_fwd", x.data_ptr(), w.data_ptr(), yy.data_ptr(), part.data_ptr(), *geom, st), flush)
            row["native_dgrad_us"] = timeit(lambda: L.call("ddl_conv_dgrad", dy.data_ptr(), w.data_ptr(), dx.data_ptr(), *geom, st), flush)
            row["native_wgrad_us"] = timeit(lambda: L.call("ddl_conv_wgrad", x.data_ptr(), dy.data_ptr(), dw.data_ptr(), *dw.stride(), scr.data_ptr(),
                                                            scr.numel(), *geom, st), flush)
            g = [f"{row['native_fwd_us']:8.1f}", f"{row['native_dgrad_us']:8.1f}", f"{row['native_wgrad_us']:8.1f}", ""]
            for key in ("fwd", "dgrad", "wgrad"):
                tot["native_" + key] = tot.get("native_" + key, 0.0) + cnt * row[f"native_{key}_us"]
            M = N * hw * hw
            if k == 1 and s == 1 and lib.ddl_conv1x1_supported(M, cout, cin):
                p1 = torch.empty(lib.ddl_conv1x1_grid(M, cout) * 2 * cout, dtype=torch.float32, device=dev)
                row["ws_fwd_stats_us"] = timeit(lambda: L.call("ddl_conv1x1_fwd", x.data_ptr(), w.data_ptr(), yy.data_ptr(), p1.data_ptr(),
                                                               M, cout, cin, st), flush)
                g[3] = f"{row['ws_fwd_stats_us']:8.1f}"
            if k == 1 and s == 1 and lib.ddl_conv1x1_supported(M, cin, cout):
                row["ws_dgrad_us"] = timeit(lambda: L.call("ddl_conv1x1_dgrad", dy.data_ptr(), w.data_ptr(), dx.data_ptr(), M, cin, cout, st), flush)
                g[3] += f" /{row['ws_dgrad_us']:6.1f}"
            del yy, dx, dw, scr
        rows.append(row)
        tot["cudnn_fwd"] += cnt * t_f
        tot["cudnn_dgrad"] += cnt * t_d
        tot["cudnn_wgrad"] += cnt * t_w
        tot["floor"] += cnt * floor
        name = f"{cin}x{hw}^2 -> {cout} k{k}/s{s}"
        print(f"{name:32s} {cnt:2d} {t_f:8.1f} {t_d:8.1f} {t_w:8.1f} {floor:7.1f} | {g[0]:>8s} {g[1]:>8s} {g[2]:>8s} {(g[3] if len(g) > 3 else ''):>16s}")
        del x, w, y, dy
    print("totals per step (ms):", {k: round(v / 1e3, 3) for k, v in tot.items()})
    os.makedirs(os.path.dirname(args.json), exist_ok=True)
    json.dump({"batch": N, "peaks": peaks, "rows": rows, "totals_ms": {k: v / 1e3 for k, v in tot.items()}}, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
