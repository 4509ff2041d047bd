#!/usr/bin/env python
"""Batch-1 ResNet-50 inference latency, the reference notebook's own experiment (notebooks/cv/onnx_experiments.py:90-104,
:173-184): PyTorch eager fp32 on the CPU (the reference's PyTorch path) against the B200 runner (bf16, BN folded, native
kernels, CUDA graph), with the reference's one-sample protocol and a warm-up + median protocol.

    python benchmarks/infer_latency.py [--json gpurun_out/infer_latency.json]
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
    ap.add_argument("--json", default=os.path.join(ROOT, "gpurun_out", "infer_latency.json"))
    args = ap.parse_args()
    from databricks_distributed_deep_learning_b200.utils import export as E

    torch.manual_seed(0)
    model = E.load_resnet50()
    model.eval()
    cats = E.load_categories()
    x = E.preprocess(E.sample_image(None, seed=0))
    out = {}
    for r in (E.EagerRunner(model, "cpu"), E.B200Runner(model)):
        res = E.measure(r, x, cats, warmup=5, samples=30)
        out[r.name] = {"reference_protocol_ms": res.reference_protocol_ms, "median_ms": res.median_ms, "min_ms": res.min_ms, "top1": res.top5[0]}
        print(r.name, out[r.name], flush=True)
    # device-only time of the graph replay (no H2D / D2H)
    b = E.B200Runner(model)
    b.prepare()
    for _ in range(5):
        b.graph.replay()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(50):
        b.graph.replay()
    e.record()
    torch.cuda.synchronize()
    out["b200_graph_replay_ms"] = s.elapsed_time(e) / 50
    print("graph replay only:", out["b200_graph_replay_ms"], "ms")
    os.makedirs(os.path.dirname(args.json), exist_ok=True)
    json.dump(out, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
