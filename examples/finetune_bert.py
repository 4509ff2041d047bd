#!/usr/bin/env python
"""BERT sequence-classification fine-tune (BASELINE.json:9-10): DDP for BERT-base / seq 128, FSDP (ZeRO-3) for
BERT-large / seq 512; Hugging Face safetensors in and out.

    python examples/finetune_bert.py --model bert-base --num-processes 8
    python examples/finetune_bert.py --model bert-large --parallelism fsdp --num-processes 8 --save out_dir
    python examples/finetune_bert.py --model bert-tiny --cpu --num-processes 2        # gloo plumbing
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def train_fn(model_name, parallelism, seq_len, batch, steps, init, save):
    from databricks_distributed_deep_learning_b200.engine import Trainer
    from databricks_distributed_deep_learning_b200.models import build_model
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg
    from databricks_distributed_deep_learning_b200.utils.checkpoint import load_safetensors, save_pretrained
    from databricks_distributed_deep_learning_b200.utils.data import SyntheticTokenLoader

    model = build_model(model_name, num_labels=2)
    if init:
        load_safetensors(init, model, strict=False)  # e.g. a Hugging Face model.safetensors
    trainer = Trainer(model, lr=2e-5, weight_decay=0.01, parallelism=parallelism)
    loader = SyntheticTokenLoader(batch, seq_len=seq_len, vocab_size=model.config.vocab_size, min_len=seq_len // 2, steps=steps,
                                  rank=pg.rank(), world_size=pg.world_size())
    out = trainer.fit(loader)
    if save:
        save_pretrained(save, trainer.ddp)  # gathers the shards under FSDP; rank 0 writes model.safetensors
    return out["losses"][-3:]


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="bert-base", choices=["bert-tiny", "bert-base", "bert-large"])
    ap.add_argument("--parallelism", default="ddp", choices=["ddp", "fsdp"])
    ap.add_argument("--num-processes", type=int, default=1)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--init", default="")
    ap.add_argument("--save", default="")
    a = ap.parse_args()
    from databricks_distributed_deep_learning_b200 import Distributor

    seq, batch = {"bert-tiny": (128, 8), "bert-base": (128, 128), "bert-large": (512, 16)}[a.model]
    print(Distributor(num_processes=a.num_processes, use_gpu=not a.cpu).run(train_fn, a.model, a.parallelism, seq, batch, a.steps, a.init, a.save))
