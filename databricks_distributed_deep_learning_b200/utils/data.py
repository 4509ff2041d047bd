"""Synthetic in-memory loaders (target layer T7).

BASELINE.json:5 replaces the notebooks' Spark/Petastorm pipeline by "synthetic
in-memory loaders of the same shape"; the mounted reference only ever loads one JPEG
(notebooks/cv/onnx_experiments.py:47-66).  Shapes provided:

* CIFAR shape   [B, 3, 32, 32]  fp32, labels in [0, 10)        (BASELINE.json:7)
* ImageNet shape[B, 3, 224, 224] fp32, labels in [0, 1000)     (BASELINE.json:8; the
  reference's input shape at onnx_experiments.py:29-31)
* token batches [B, S] int64 ids + attention mask + labels, S = 128 / 512
  (BASELINE.json:9-10)

B200-first details: batches are materialised once in *pinned* host memory (a ring of
``pool`` distinct batches, so consecutive steps do not re-read the same lines) and every
rank draws a disjoint, seed-determined shard, which is what a DistributedSampler gives.
:class:`DevicePrefetcher` overlaps the host->device copy of batch k+1 with step k on a
copy stream.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Iterator, Optional, Tuple

import torch


def _gen(seed: int, rank: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(seed * 1_000_003 + rank * 7919 + 17)
    return g


def _pin(t: torch.Tensor, pin: bool) -> torch.Tensor:
    if pin and torch.cuda.is_available():
        return t.pin_memory()
    return t


@dataclass
class SyntheticImageLoader:
    """Endless (or ``steps``-long) stream of (images, labels) for one rank."""

    batch_size: int
    image_size: int = 224
    num_classes: int = 1000
    channels: int = 3
    steps: Optional[int] = None
    rank: int = 0
    world_size: int = 1
    seed: int = 0
    pool: int = 4
    pin_memory: bool = True
    dtype: torch.dtype = torch.float32

    def __post_init__(self):
        g = _gen(self.seed, self.rank)
        self._x, self._y = [], []
        for _ in range(self.pool):
            x = torch.randn(self.batch_size, self.channels, self.image_size, self.image_size, generator=g)
            y = torch.randint(0, self.num_classes, (self.batch_size,), generator=g)
            self._x.append(_pin(x.to(self.dtype), self.pin_memory))
            self._y.append(_pin(y, self.pin_memory))

    def __len__(self) -> int:
        return self.steps if self.steps is not None else 1 << 30

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        i = 0
        while self.steps is None or i < self.steps:
            yield self._x[i % self.pool], self._y[i % self.pool]
            i += 1

    @property
    def bytes_per_batch(self) -> int:
        return self._x[0].numel() * self._x[0].element_size() + self._y[0].numel() * self._y[0].element_size()


def cifar_shape_loader(batch_size: int, **kw) -> SyntheticImageLoader:
    kw.setdefault("image_size", 32)
    kw.setdefault("num_classes", 10)
    return SyntheticImageLoader(batch_size=batch_size, **kw)


def imagenet_shape_loader(batch_size: int, **kw) -> SyntheticImageLoader:
    kw.setdefault("image_size", 224)
    kw.setdefault("num_classes", 1000)
    return SyntheticImageLoader(batch_size=batch_size, **kw)


@dataclass
class SyntheticTokenLoader:
    """Endless stream of dicts {input_ids, attention_mask, token_type_ids, labels}.

    ``min_len`` < ``seq_len`` produces right-padded sequences (HF tokenizer layout) so
    the padding-mask path of the attention kernel is exercised.
    """

    batch_size: int
    seq_len: int = 128
    vocab_size: int = 30522
    num_labels: int = 2
    min_len: Optional[int] = None
    steps: Optional[int] = None
    rank: int = 0
    world_size: int = 1
    seed: int = 0
    pool: int = 4
    pin_memory: bool = True

    def __post_init__(self):
        g = _gen(self.seed, self.rank)
        self._b = []
        lo = self.seq_len if self.min_len is None else self.min_len
        for _ in range(self.pool):
            ids = torch.randint(min(1000, self.vocab_size // 2), self.vocab_size, (self.batch_size, self.seq_len), generator=g)
            lens = torch.randint(lo, self.seq_len + 1, (self.batch_size,), generator=g)
            mask = (torch.arange(self.seq_len)[None, :] < lens[:, None]).long()
            ids = ids * mask  # [PAD] = 0
            ids[:, 0] = min(101, self.vocab_size - 1)  # [CLS]
            batch = {
                "input_ids": _pin(ids, self.pin_memory),
                "attention_mask": _pin(mask, self.pin_memory),
                "token_type_ids": _pin(torch.zeros_like(ids), self.pin_memory),
                "labels": _pin(torch.randint(0, self.num_labels, (self.batch_size,), generator=g), self.pin_memory),
            }
            self._b.append(batch)

    def __len__(self) -> int:
        return self.steps if self.steps is not None else 1 << 30

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        i = 0
        while self.steps is None or i < self.steps:
            yield self._b[i % self.pool]
            i += 1

    @property
    def bytes_per_batch(self) -> int:
        return sum(t.numel() * t.element_size() for t in self._b[0].values())


class DevicePrefetcher:
    """Copies batch k+1 host->device on a side stream while step k computes.

    Yields device tensors whose storage is reused every other step (double buffer),
    which also makes them valid static inputs for a captured CUDA graph.
    """

    def __init__(self, loader, device: torch.device, depth: int = 2):
        self.loader = loader
        self.device = device
        self.depth = depth
        self.stream = torch.cuda.Stream(device=device) if device.type == "cuda" else None
        self._slots = None

    def _to_device(self, batch, slot):
        def cp(dst, src):
            dst.copy_(src, non_blocking=True)
            return dst

        if isinstance(batch, dict):
            if slot is None:
                return {k: torch.empty_like(v, device=self.device) for k, v in batch.items()}
            return {k: cp(slot[k], v) for k, v in batch.items()}
        if slot is None:
            return tuple(torch.empty_like(v, device=self.device) for v in batch)
        return tuple(cp(d, v) for d, v in zip(slot, batch))

    def __iter__(self):
        it = iter(self.loader)
        if self.stream is None:
            for b in it:
                yield b
            return
        first = next(it, None)
        if first is None:
            return
        if self._slots is None:
            self._slots = [self._to_device(first, None) for _ in range(self.depth)]
        events = [torch.cuda.Event() for _ in range(self.depth)]
        cur = torch.cuda.current_stream(self.device)
        i = 0
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            nxt = self._to_device(first, self._slots[0])
            events[0].record(self.stream)
        while nxt is not None:
            cur.wait_event(events[i % self.depth])
            out = nxt
            b = next(it, None)
            i += 1
            if b is not None:
                # slot i%depth was consumed `depth` steps ago; order the overwrite after it
                self.stream.wait_stream(cur)
                with torch.cuda.stream(self.stream):
                    nxt = self._to_device(b, self._slots[i % self.depth])
                    events[i % self.depth].record(self.stream)
            else:
                nxt = None
            yield out
