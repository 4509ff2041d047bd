"""tcgen05/TMEM GEMM front end (K8) - filled in by csrc/gemm_tcgen05.cu."""
from __future__ import annotations

import torch


def usable(x: torch.Tensor, weight: torch.Tensor) -> bool:
    return False


def linear(x, weight, bias=None, activation=None):
    raise NotImplementedError
