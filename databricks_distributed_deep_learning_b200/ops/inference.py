"""Inference-only forward of a BatchNorm-folded ResNet on the native kernels ("inference-side native
equivalent"; what the reference does at notebooks/cv/onnx_experiments.py:173-174 with the BN constant folding of :39).

Every layer is: implicit-GEMM convolution on tcgen05 (``csrc/conv_igemm.cu``; the 7x7 stem through ``csrc/stem_conv.cu``)
followed by one pass ``y = relu(conv + bias [+ residual])`` (``ddl_bn_apply`` with scale 1 and the folded bias as shift).
No autograd, no library convolution; NHWC bf16 activations throughout.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import _lib as L
from . import functional as OF


class FoldedConv:
    """Weights of one folded convolution in the layout the kernels take."""

    def __init__(self, conv: nn.Conv2d, device, dtype=torch.bfloat16):
        self.stride, self.pad = conv.stride[0], conv.padding[0]
        self.cout, self.cin, self.r, self.s = conv.weight.shape
        w = conv.weight.detach().to(device, dtype)
        self.weight = w.contiguous(memory_format=torch.channels_last)  # [Cout][R][S][C] in memory
        self.shift = conv.bias.detach().to(device, torch.float32).contiguous()
        self.scale = torch.ones(self.cout, dtype=torch.float32, device=device)

    def __call__(self, x: torch.Tensor, relu: bool, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        n, c, h, w = x.shape
        oh = (h + 2 * self.pad - self.r) // self.stride + 1
        ow = (w + 2 * self.pad - self.s) // self.stride + 1
        dev = x.device
        st = L.stream_ptr(dev)
        y = torch.empty((n, self.cout, oh, ow), dtype=x.dtype, device=dev, memory_format=torch.channels_last)
        geom = (n, h, w, c, self.cout, self.r, self.s, self.stride, self.pad)
        if not L.lib().ddl_conv_supported(*geom):
            raise L.DdlError(f"native inference: convolution shape {geom} is outside ddl_conv_supported")
        L.call("ddl_conv_fwd", x.data_ptr(), self.weight.data_ptr(), y.data_ptr(), None, *geom, st)
        out = torch.empty_like(y, memory_format=torch.channels_last)
        L.call("ddl_bn_apply", y.data_ptr(), L.ptr(residual), self.scale.data_ptr(), self.shift.data_ptr(), out.data_ptr(),
               n * oh * ow, self.cout, int(relu), st)
        return out


class FoldedResNet:
    """Forward of a ``utils.export.fold_batchnorm`` model (Bottleneck or BasicBlock ResNet) on the native kernels."""

    def __init__(self, folded: nn.Module, device="cuda", dtype=torch.bfloat16):
        from ..models.resnet import BasicBlock, Bottleneck

        self.device, self.dtype = torch.device(device), dtype
        stem = folded.conv1
        if tuple(stem.weight.shape) != (64, 3, 7, 7) or stem.stride != (2, 2):
            raise L.DdlError("native inference expects the 7x7 / stride-2 ImageNet stem")
        self.stem_w = stem.weight.detach().to(self.device, dtype).contiguous(memory_format=torch.channels_last)
        self.stem_shift = stem.bias.detach().to(self.device, torch.float32).contiguous()
        self.stem_scale = torch.ones(64, dtype=torch.float32, device=self.device)
        self.blocks = []
        for layer in (folded.layer1, folded.layer2, folded.layer3, folded.layer4):
            for blk in layer:
                convs = [FoldedConv(blk.conv1, self.device, dtype), FoldedConv(blk.conv2, self.device, dtype)]
                if isinstance(blk, Bottleneck):
                    convs.append(FoldedConv(blk.conv3, self.device, dtype))
                elif not isinstance(blk, BasicBlock):
                    raise L.DdlError(f"unknown block type {type(blk)}")
                down = FoldedConv(blk.downsample[0], self.device, dtype) if blk.downsample is not None else None
                self.blocks.append((convs, down))
        self.fc_w = folded.fc.weight.detach().to(self.device, dtype).contiguous()
        self.fc_b = folded.fc.bias.detach().to(self.device, dtype).contiguous()

    @torch.no_grad()
    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        """x: [N, 3, H, W] fp32 or bf16, NCHW contiguous on the device.  Returns bf16 logits [N, classes]."""
        lib = L.lib()
        n, _, h, w = x.shape
        dev = x.device
        st = L.stream_ptr(dev)
        # stem: space-to-depth + tcgen05 implicit GEMM, then bias + ReLU
        z = torch.empty(lib.ddl_stem_z_bytes(n, h, w) // 2, dtype=torch.bfloat16, device=dev)
        L.call("ddl_stem_s2d", x.data_ptr(), L.dtype_code(x.dtype), z.data_ptr(), n, h, w, st)
        y = torch.empty((n, 64, h // 2, w // 2), dtype=torch.bfloat16, device=dev, memory_format=torch.channels_last)
        part = torch.empty(lib.ddl_stem_fwd_grid(n, h, w) * 2 * 64, dtype=torch.float32, device=dev)
        w2 = torch.empty(4 * 64 * 64, dtype=torch.bfloat16, device=dev)
        L.call("ddl_stem_conv_fwd", z.data_ptr(), self.stem_w.data_ptr(), *self.stem_w.stride(), w2.data_ptr(), y.data_ptr(), part.data_ptr(),
               n, h, w, st)
        a = torch.empty_like(y, memory_format=torch.channels_last)
        L.call("ddl_bn_apply", y.data_ptr(), None, self.stem_scale.data_ptr(), self.stem_shift.data_ptr(), a.data_ptr(),
               n * (h // 2) * (w // 2), 64, 1, st)
        a = OF.maxpool3x3s2(a)
        for convs, down in self.blocks:
            identity = down(a, relu=False) if down is not None else a
            out = a
            for c in convs[:-1]:
                out = c(out, relu=True)
            a = convs[-1](out, relu=True, residual=identity)
        pooled = OF.global_avgpool(a)
        return OF.linear(pooled, self.fc_w, self.fc_b)
