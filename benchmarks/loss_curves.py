#!/usr/bin/env python
"""Prints the loss curves of the Trainer and of stock DDP + autocast on the same batches (tests/_gpu_workers.py::w_trainer_vs_stock):
    python benchmarks/loss_curves.py resnet50|bert-base"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

if __name__ == "__main__":
    import _gpu_workers as GW
    from databricks_distributed_deep_learning_b200 import Distributor

    wl = sys.argv[1] if len(sys.argv) > 1 else "bert-base"
    out = Distributor(num_processes=1, use_gpu=True, timeout_s=600).run(GW.w_trainer_vs_stock, wl, 12)
    print("ours ", [round(v, 4) for v in out["ours"]])
    print("stock", [round(v, 4) for v in out["stock"]])
    print(out["mean_rel"], out["max_rel"])
