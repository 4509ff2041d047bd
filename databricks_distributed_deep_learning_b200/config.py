"""Small dataclass configs, one per workload (SURVEY.md §5 "Config / flag system").

The reference hard-codes every literal (image size at notebooks/cv/onnx_experiments.py:29-30,
opset :38, paths :36/:48/:198/:215, tolerances :144).  The four workload configs below are
the ones BASELINE.json:7-10 names.
"""
from __future__ import annotations

from dataclasses import asdict, dataclass, field
from typing import Any, Dict, Optional


@dataclass
class OptimConfig:
    lr: float = 1e-3
    betas: tuple = (0.9, 0.999)
    eps: float = 1e-8
    weight_decay: float = 0.01


@dataclass
class WorkloadConfig:
    name: str
    model: str
    per_gpu_batch: int
    parallelism: str = "ddp"          # "ddp" | "fsdp"
    image_size: Optional[int] = None
    num_classes: Optional[int] = None
    seq_len: Optional[int] = None
    dtype: str = "bf16"
    bucket_cap_mb: float = 25.0
    optim: OptimConfig = field(default_factory=OptimConfig)
    unit: str = "img/s"

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)


WORKLOADS: Dict[str, WorkloadConfig] = {
    # BASELINE.json:7 - plumbing, CPU/gloo world_size=2
    "resnet18-cifar": WorkloadConfig("resnet18-cifar", "resnet18", 32, image_size=32, num_classes=10, dtype="fp32"),
    # BASELINE.json:8 - the headline CV config
    "resnet50": WorkloadConfig("resnet50", "resnet50", 256, image_size=224, num_classes=1000),
    # BASELINE.json:9
    "bert-base": WorkloadConfig("bert-base", "bert-base", 128, seq_len=128, num_classes=2, unit="seq/s",
                                optim=OptimConfig(lr=2e-5)),
    # BASELINE.json:10
    "bert-large": WorkloadConfig("bert-large", "bert-large", 16, parallelism="fsdp", seq_len=512, num_classes=2,
                                 unit="seq/s", optim=OptimConfig(lr=2e-5)),
}


def get_workload(name: str) -> WorkloadConfig:
    if name not in WORKLOADS:
        raise KeyError(f"unknown workload {name!r}; choose from {sorted(WORKLOADS)}")
    return WORKLOADS[name]
