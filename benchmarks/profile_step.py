#!/usr/bin/env python
"""Per-kernel device-time breakdown of one training step (torch.profiler / CUPTI).

Shares only - numbers taken under a profiler are never reported as benchmark values.
    python benchmarks/profile_step.py --workload bert-base [--impl ours|stock] [--batch N]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="bert-base")
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--dropout", type=float, default=None)
    ap.add_argument("--top", type=int, default=45)
    ap.add_argument("--trace", default="", help="kernel-name substring: list the ATen ops (with shapes and callers) that launch it")
    args = ap.parse_args()
    import bench

    from databricks_distributed_deep_learning_b200.config import get_workload
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29544")
    dev = pg.init_process_group(use_gpu=True)
    torch.backends.cudnn.benchmark = True
    wl = get_workload(args.workload)
    batch = args.batch or wl.per_gpu_batch
    ns = argparse.Namespace(workload=args.workload, dropout=args.dropout, no_graph=True, comm="auto", bucket_mb=25.0, reshard=False)
    runner = bench.OursRunner(ns, wl, args.workload, dev) if args.impl == "ours" else bench.StockRunner(ns, wl, args.workload, dev)
    loader = bench.build_batches(args.workload, batch, 0, 1, pool=1)
    b = bench.to_device(next(iter(loader)), dev)
    for _ in range(4):
        runner.step_device(b)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile

    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(2):
            runner.step_device(b)
        torch.cuda.synchronize()
    ev = prof.key_averages()
    rows = sorted(((e.key, e.device_time_total / 2.0, e.count // 2) for e in ev if e.device_time_total > 0), key=lambda r: -r[1])
    total = sum(r[1] for r in rows)
    print(f"# {args.workload} impl={args.impl} batch={batch}: sum of kernel time per step = {total/1e3:.3f} ms over {sum(r[2] for r in rows)} launches")
    for name, us, cnt in rows[: args.top]:
        print(f"{us/1e3:9.3f} ms {100*us/total:5.1f}%  x{cnt:<5d} {name[:150]}")
    out = os.path.join(ROOT, "gpurun_out", f"profile_{args.workload}_{args.impl}.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    with open(out, "w") as f:
        json.dump({"workload": args.workload, "impl": args.impl, "batch": batch, "total_ms": total / 1e3,
                   "kernels": [{"name": n, "ms": u / 1e3, "count": c} for n, u, c in rows]}, f, indent=1)
    if args.trace:
        trace_kernel(runner, b, args.trace)
    pg.destroy_process_group()


def trace_kernel(runner, b, needle: str) -> None:
    """Which ATen op (shapes, parent ops) launches the kernels whose name contains ``needle``."""
    from torch.profiler import ProfilerActivity, profile

    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
        runner.step_device(b)
        torch.cuda.synchronize()
    print(f"# ops launching kernels matching {needle!r}")
    for e in prof.events():
        ks = [k for k in (e.kernels or []) if needle in k.name]
        if not ks:
            continue
        chain, par = [], e.cpu_parent
        while par is not None and len(chain) < 6:
            chain.append(par.name)
            par = par.cpu_parent
        us = sum(k.duration for k in ks)
        print(f"{us:9.1f} us  {e.name}  shapes={e.input_shapes}  <- {' <- '.join(chain)}")


if __name__ == "__main__":
    main()
