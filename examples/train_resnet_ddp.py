#!/usr/bin/env python
"""ResNet image classification, data-parallel (BASELINE.json:7-8) - the CV notebook's training counterpart.

CPU / gloo plumbing run (2 ranks, ResNet-18 on 32x32):   python examples/train_resnet_ddp.py --cpu --num-processes 2
B200 run (ResNet-50 on 224x224, one process per GPU):     python examples/train_resnet_ddp.py --num-processes 8
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def train_fn(model_name, image_size, num_classes, batch, steps, ckpt):
    from databricks_distributed_deep_learning_b200.engine import Trainer
    from databricks_distributed_deep_learning_b200.models import build_model
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg
    from databricks_distributed_deep_learning_b200.utils.checkpoint import save_checkpoint
    from databricks_distributed_deep_learning_b200.utils.data import SyntheticImageLoader
    from databricks_distributed_deep_learning_b200.utils.timing import JsonlLogger

    model = build_model(model_name, num_classes=num_classes, small_input=image_size <= 64)
    trainer = Trainer(model, lr=1e-3)  # bf16 + fused kernels + CUDA graph on GPU, fp32 eager on CPU
    loader = SyntheticImageLoader(batch, image_size=image_size, num_classes=num_classes, steps=steps,
                                  rank=pg.rank(), world_size=pg.world_size())
    # step records (loss, samples/s) plus one record with the per-bucket all-reduce time / bytes / bus bandwidth of this model
    out = trainer.fit(loader, log_every=max(1, steps // 5), logger=JsonlLogger(also_print=True), log_comm=True)
    if ckpt:
        save_checkpoint(ckpt, trainer.ddp, trainer.opt, step=out["steps"])  # rank 0 writes, everyone syncs
    return {"final_loss": out["losses"][-1] if out["losses"] else None, "world": pg.world_size()}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--num-processes", type=int, default=1)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--checkpoint", default="")
    a = ap.parse_args()
    from databricks_distributed_deep_learning_b200 import Distributor

    cfg = ("resnet18", 32, 10, 32) if a.cpu else ("resnet50", 224, 1000, 256)
    print(Distributor(num_processes=a.num_processes, use_gpu=not a.cpu).run(train_fn, *cfg, a.steps, a.checkpoint))
