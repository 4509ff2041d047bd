#!/usr/bin/env python
"""Gradient all-reduce bandwidth sweep, 1 KB - 1 GB (BASELINE.json:11).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/allreduce_sweep.py \
        [--max-mb 1024] [--out gpurun_out/allreduce_wN.json] [--write-table]

For every message size: our one-shot / two-shot / NVLS kernels at several CTA counts and
NCCL (``dist.all_reduce``) as the baseline.  Device-timed with CUDA events around a batch
of back-to-back launches, max over ranks.  algbw = bytes / time; busbw = algbw * 2(N-1)/N;
the roofline fractions use 900 GB/s (nominal, BASELINE.json:5) and the 770 GB/s measured
peer-copy figure of the profiling recipe.  ``--write-table`` stores the per-size winner in
``profiles/allreduce_table_wN.json``, which :class:`VariantSelector` loads (K4).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--min-kb", type=int, default=1)
    ap.add_argument("--out", default=None)
    ap.add_argument("--write-table", action="store_true")
    ap.add_argument("--provider", default=None, help="ipc | torch (torch also enables NVLS where exposed)")
    ap.add_argument("--dtype", default="bf16")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    from databricks_distributed_deep_learning_b200.parallel import process_group as pg
    from databricks_distributed_deep_learning_b200.parallel.collectives import PeerMemoryComm
    from databricks_distributed_deep_learning_b200.parallel.symm import SymmetricHeap

    dev = pg.init_process_group(use_gpu=True)
    rank, world = pg.rank(), pg.world_size()
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    esz = 2 if dtype == torch.bfloat16 else 4
    heap = SymmetricHeap(dev, provider=args.provider) if args.provider else None
    nvls_note = ""
    try:
        comm = PeerMemoryComm(dev, heap=heap)
    except Exception as e:  # provider not available: fall back to IPC and say so
        nvls_note = f"provider {args.provider} failed: {type(e).__name__}: {e}"
        comm = PeerMemoryComm(dev, heap=SymmetricHeap(dev, provider="ipc"))
    nvls = comm.heap.multicast_supported
    max_bytes = args.max_mb << 20
    buf_all = comm.alloc_flat(max_bytes // esz, dtype, dev)
    nccl_buf = torch.zeros(max_bytes // esz, dtype=dtype, device=dev)
    buf_all.normal_()

    def timeit(fn, nbytes):
        iters = 200 if nbytes <= (1 << 20) else (40 if nbytes <= (64 << 20) else 10)
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        pg.barrier()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
            fn()
        e.record()
        torch.cuda.synchronize()
        return pg.max_over_ranks(s.elapsed_time(e) / iters)  # ms

    rows = []
    size = args.min_kb << 10
    while size <= max_bytes:
        n = size // esz
        n = max(n // (8 * world) * (8 * world), 8 * world)
        nbytes = n * esz
        view = buf_all[:n]
        nview = nccl_buf[:n]
        cands = []
        one_max = 2 << 20
        for blocks in (4, 8, 16, 32, 64, 128):
            if nbytes <= one_max and blocks <= 64:
                cands.append(("oneshot", blocks))
            cands.append(("twoshot", blocks))
            if nvls:
                cands.append(("nvls", blocks))
        res = {}
        if nvls and dtype == torch.bfloat16 and nbytes >= (2 << 20) and world in (2, 4, 8):
            # hybrid: part of the message through the switch, the rest over peer loads / stores, one kernel
            for permille in (500, 600, 700, 800):
                for nvb, p2b in ((8, 32), (8, 64), (16, 48)):
                    key = f"hybrid{permille}/{nvb}+{p2b}"
                    try:
                        res[key] = {"ms": timeit(lambda: comm.all_reduce_hybrid_(view, permille, nvb, p2b), nbytes)}
                    except Exception as ex:
                        res[key] = {"error": str(ex)[:200]}
        for algo, blocks in cands:
            if algo == "oneshot" and nbytes > comm_oneshot_max(dtype, blocks):
                continue
            if nbytes < (64 << 10) and blocks > 16:
                continue
            try:
                ms = timeit(lambda: comm.all_reduce_mean_(view, algo=algo, num_blocks=blocks), nbytes)
            except Exception as ex:
                res[f"{algo}/{blocks}"] = {"error": str(ex)[:200]}
                continue
            res[f"{algo}/{blocks}"] = {"ms": ms}
        ms_nccl = timeit(lambda: dist.all_reduce(nview), nbytes)
        best_key = min((k for k in res if "ms" in res[k]), key=lambda k: res[k]["ms"])
        f = 2 * (world - 1) / world
        row = {
            "bytes": nbytes, "best": best_key, "best_ms": res[best_key]["ms"], "nccl_ms": ms_nccl,
            "best_busbw_gbs": nbytes / res[best_key]["ms"] / 1e6 * f, "nccl_busbw_gbs": nbytes / ms_nccl / 1e6 * f,
            "all": res,
        }
        row["frac_of_900"] = row["best_busbw_gbs"] / 900.0
        row["frac_of_770_measured_peer_copy"] = row["best_busbw_gbs"] / 770.0
        row["speedup_vs_nccl"] = ms_nccl / res[best_key]["ms"]
        rows.append(row)
        if rank == 0:
            print(f"{nbytes:>12d} B  best {best_key:<18s} {row['best_ms']*1e3:9.1f} us  busbw {row['best_busbw_gbs']:7.1f} GB/s "
                  f"({row['frac_of_900']*100:5.1f}% of 900) | nccl {ms_nccl*1e3:9.1f} us busbw {row['nccl_busbw_gbs']:7.1f} GB/s "
                  f"| x{row['speedup_vs_nccl']:.2f}", flush=True)
        size *= 2
    # correctness spot check at the end (sum of identical buffers == world * x for scale 1)
    chk = comm.alloc_flat(8 * world * 1024, torch.float32, dev)
    chk.fill_(float(rank + 1))
    pg.barrier()
    comm.all_reduce_mean_(chk, algo="twoshot", scale=1.0)
    torch.cuda.synchronize()
    ok = bool((chk == world * (world + 1) / 2).all())
    if rank == 0:
        out = {"world": world, "dtype": args.dtype, "provider": comm.heap.provider, "nvls": nvls, "note": nvls_note,
               "check_ok": ok, "rows": rows}
        path = args.out or os.path.join(ROOT, "gpurun_out", f"allreduce_w{world}.json")
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "w") as fh:
            json.dump(out, fh, indent=1)
        if args.write_table:
            table = []
            for r in rows:
                algo, blocks = r["best"].split("/")
                if algo.startswith("hybrid"):
                    nvb, p2b = blocks.split("+")
                    table.append({"max_bytes": r["bytes"], "algo": "hybrid", "permille": int(algo[6:]), "blocks": int(nvb), "p2p_blocks": int(p2b)})
                else:
                    table.append({"max_bytes": r["bytes"], "algo": algo, "blocks": int(blocks)})
            table[-1]["max_bytes"] = 1 << 62
            tpath = os.path.join(ROOT, "gpurun_out", f"allreduce_table_w{world}.json")
            with open(tpath, "w") as fh:
                json.dump({"world": world, "source": os.path.basename(path), "table": table}, fh, indent=1)
        print("check_ok", ok, "->", path)
    pg.barrier()
    pg.destroy_process_group()


def comm_oneshot_max(dtype, blocks):
    from databricks_distributed_deep_learning_b200.ops import _lib as L

    return L.lib().ddl_allreduce_oneshot_max_bytes(L.dtype_code(dtype), blocks)


if __name__ == "__main__":
    main()
