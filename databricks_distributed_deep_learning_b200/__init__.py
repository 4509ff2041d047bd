"""b200-ddl: a Blackwell-native data-parallel training framework.

Capabilities follow BASELINE.json (launcher, DDP-compatible wrapper, FSDP/ZeRO-3 mode,
rank-0 state_dict / safetensors checkpoints, ResNet + BERT workloads on synthetic data)
plus what the mounted reference notebook really does (export + single-image inference
comparison).  See DESIGN.md for the architecture.
"""
from .launcher import DistributedError, Distributor, HorovodRunner, TorchDistributor  # noqa: F401
from .parallel import process_group  # noqa: F401
from .parallel.ddp import DDP, DistributedDataParallel  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):  # lazy: these pull in heavier modules
    if name in ("FullyShardedDataParallel", "FSDP"):
        from .parallel.fsdp import FullyShardedDataParallel

        return FullyShardedDataParallel
    if name == "Trainer":
        from .engine import Trainer

        return Trainer
    if name == "FusedAdamW":
        from .optim import FusedAdamW

        return FusedAdamW
    raise AttributeError(name)
