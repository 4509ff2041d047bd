"""ResNet-18 / ResNet-50 (v1.5) for the CV workload.

Capability source: the reference builds ``torchvision.models.resnet50`` at
notebooks/cv/onnx_experiments.py:19 (25,557,032 parameters in 161 tensors, SURVEY.md
§2.A A3) and BASELINE.json:7-8 add ResNet-18 on 32x32 inputs and ResNet-50 training.
Parameter and buffer names are identical to torchvision's so that checkpoints move in
both directions (``conv1.weight``, ``layer1.0.bn1.running_mean``, ``fc.bias`` ...).

Two execution paths share the same parameters:

* ``impl="torch"``  - plain ATen ops (CPU plumbing config, export, oracle in tests);
* ``impl="fused"``  - NHWC bf16 activations; every conv -> BN -> (+residual) -> ReLU group
  goes through :func:`..ops.functional.conv_bn_act`, whose batch-norm statistics,
  normalise, residual add and ReLU are hand-written sm_100a kernels
  (``csrc/bn_act.cu``) and whose 1x1 convolutions run on the tcgen05 GEMM
  (``csrc/gemm_tcgen05.cu``) (K9 / K14 in SURVEY.md §2.E.3).
"""
from __future__ import annotations

import os
from typing import List, Optional, Type

import torch
import torch.nn as nn
import torch.nn.functional as F


def _conv_bn_act(x, conv: nn.Conv2d, bn: nn.BatchNorm2d, relu: bool, residual, impl: str):
    if impl == "fused" and x.is_cuda:
        from ..ops import functional as OF

        return OF.conv_bn_act(x, conv, bn, relu=relu, residual=residual)
    y = bn(conv(x))
    if residual is not None:
        y = y + residual
    return F.relu(y) if relu else y


def _fork(x, impl: str):
    """(main, skip) aliases of a block input.  On the fused path the two gradients are summed inside the
    backward kernel of the op that produced ``x`` (ops.functional.residual_fork) instead of by an ATen add."""
    if impl == "fused" and x.is_cuda:
        from ..ops import functional as OF

        return OF.residual_fork(x)
    return x, x


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.downsample = downsample
        self.impl = "torch"

    def forward(self, x):
        x, identity = _fork(x, self.impl)
        if self.downsample is not None:
            identity = _conv_bn_act(identity, self.downsample[0], self.downsample[1], False, None, self.impl)
        out = _conv_bn_act(x, self.conv1, self.bn1, True, None, self.impl)
        return _conv_bn_act(out, self.conv2, self.bn2, True, identity, self.impl)


class Bottleneck(nn.Module):
    """1x1 -> 3x3 (stride here: the v1.5 variant torchvision ships) -> 1x1."""

    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * 4)
        self.downsample = downsample
        self.impl = "torch"

    def forward(self, x):
        x, identity = _fork(x, self.impl)
        if self.downsample is not None:
            identity = _conv_bn_act(identity, self.downsample[0], self.downsample[1], False, None, self.impl)
        out = _conv_bn_act(x, self.conv1, self.bn1, True, None, self.impl)
        out = _conv_bn_act(out, self.conv2, self.bn2, True, None, self.impl)
        return _conv_bn_act(out, self.conv3, self.bn3, True, identity, self.impl)


class ResNet(nn.Module):
    def __init__(
        self,
        block: Type[nn.Module],
        layers: List[int],
        num_classes: int = 1000,
        small_input: bool = False,
        impl: str = "torch",
    ):
        """``small_input`` swaps the 7x7/2 stem + max-pool for a 3x3/1 stem, the usual
        CIFAR-shape (32x32) adaptation used for BASELINE.json:7."""
        super().__init__()
        self.inplanes = 64
        self.small_input = small_input
        # fused path, opt-in: zero-pad the 3-channel input (and stem weight) to 8 channels (ops.functional.to_nhwc_bf16).
        # Measured on B200 (profiles/resnet50_step_r1.md): cuDNN then picks a 256x64 tcgen05 implicit GEMM that takes
        # 2.8 ms against 1.47 ms for the 3-channel legacy kernel, so the default stays 0.
        self.stem_pad = int(os.environ.get("DDL_STEM_PAD", "0"))
        if small_input:
            self.conv1 = nn.Conv2d(3, 64, 3, 1, 1, bias=False)
        else:
            self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.layer1 = self._make_layer(block, 64, layers[0], 1)
        self.layer2 = self._make_layer(block, 128, layers[1], 2)
        self.layer3 = self._make_layer(block, 256, layers[2], 2)
        self.layer4 = self._make_layer(block, 512, layers[3], 2)
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
        self.set_impl(impl)

    def _make_layer(self, block, planes, blocks, stride):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(
                nn.Conv2d(self.inplanes, planes * block.expansion, 1, stride, bias=False),
                nn.BatchNorm2d(planes * block.expansion),
            )
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes))
        return nn.Sequential(*layers)

    def set_impl(self, impl: str) -> "ResNet":
        if impl not in ("torch", "fused"):
            raise ValueError(impl)
        self.impl = impl
        for m in self.modules():
            if isinstance(m, (BasicBlock, Bottleneck)):
                m.impl = impl
        return self

    def forward(self, x):
        if self.impl == "fused" and x.is_cuda:
            from ..ops import functional as OF

            with OF.batches_counted_once(self):  # one kernel for all num_batches_tracked counters
                return self._forward(x)
        return self._forward(x)

    def _forward(self, x):
        fused = self.impl == "fused" and x.is_cuda
        if fused:
            from ..ops import functional as OF

            if OF.stem_usable(x, self.conv1):
                # native stem: space-to-depth + tcgen05 implicit GEMM, BatchNorm statistics from its epilogue
                y, partials = OF.stem_conv(x, self.conv1.weight)
                if not self.small_input and OF.bn_relu_maxpool_usable(y, self.bn1):
                    # BatchNorm + ReLU + max pool in one pass: the normalised 112 x 112 activation is never written
                    return self._trunk(OF.bn_relu_maxpool(y, self.bn1, partials), fused, pooled=True)
                x = OF.bn_act(y, self.bn1, relu=True, partials=partials)
                return self._trunk(x, fused)
            x = OF.to_nhwc_bf16(x, pad_channels=self.stem_pad)
            if x.shape[1] != self.conv1.in_channels:
                # input zero-padded to 8 channels: pad the 3-channel stem weight the same way
                c = self.conv1
                y = F.conv2d(x, OF.pad_in_channels(c.weight, x.shape[1]), None, c.stride, c.padding)
                x = OF.bn_act(y, self.bn1, relu=True)
            else:
                x = _conv_bn_act(x, self.conv1, self.bn1, True, None, self.impl)
        else:
            x = _conv_bn_act(x, self.conv1, self.bn1, True, None, self.impl)
        return self._trunk(x, fused)

    def _trunk(self, x, fused: bool, pooled: bool = False):
        if fused:
            from ..ops import functional as OF
        if not self.small_input and not pooled:
            x = OF.maxpool3x3s2(x) if fused else F.max_pool2d(x, 3, 2, 1)
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        if fused:
            return OF.linear(OF.global_avgpool(x), self.fc.weight, self.fc.bias)
        return self.fc(torch.flatten(F.adaptive_avg_pool2d(x, 1), 1))


def resnet18(num_classes: int = 1000, small_input: bool = False, impl: str = "torch") -> ResNet:
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes, small_input, impl)


def resnet50(num_classes: int = 1000, small_input: bool = False, impl: str = "torch") -> ResNet:
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes, small_input, impl)
