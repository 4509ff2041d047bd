"""Autograd-aware wrappers around the native sm_100a kernels (target layer T5).

Every function here launches hand-written kernels from ``libb200ddl.so`` through the C ABI
(``include/b200ddl.h``).  There is no eager fallback on CUDA tensors: if the library is
missing the call raises.  The plain-PyTorch expression of each op (used as the fp32 oracle
in ``tests/``) is given in each docstring.
"""
from __future__ import annotations

import itertools
import os
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.autograd.function import once_differentiable

from . import _lib as L

# ----------------------------------------------------------------------------- RNG state
_salt_counter = itertools.count(1)
_seed_words = {}


def seed_word(device: torch.device) -> torch.Tensor:
    """Device-resident uint64 mixed into every dropout mask; bump it once per step (works
    under CUDA-graph replay, where host-side seeds would be frozen)."""
    key = (device.type, device.index)
    if key not in _seed_words:
        _seed_words[key] = torch.zeros(1, dtype=torch.int64, device=device)
    return _seed_words[key]


def advance_seed(device: torch.device) -> None:
    seed_word(device).add_(1)


def manual_seed(seed: int, device: torch.device) -> None:
    seed_word(device).fill_(int(seed))


def _next_salt() -> int:
    return (next(_salt_counter) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF


def _scratch(nbytes: int, device) -> torch.Tensor:
    return torch.empty(max(nbytes, 16), dtype=torch.uint8, device=device)


def _req_bf16(*ts):
    for t in ts:
        if t is not None and t.dtype != torch.bfloat16:
            raise L.DdlError(f"native op needs bfloat16 tensors, got {t.dtype}")


# ----------------------------------------------------------------------------- cross entropy
class _CrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        logits = logits.contiguous()
        B, Cn = logits.shape
        row = torch.empty(B, dtype=torch.float32, device=logits.device)
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        dlogits = torch.empty_like(logits)
        L.call("ddl_cross_entropy", logits.data_ptr(), L.dtype_code(logits.dtype), labels.data_ptr(), row.data_ptr(),
               loss.data_ptr(), dlogits.data_ptr(), B, Cn, 1.0 / B, int(ignore_index), L.stream_ptr(logits.device))
        ctx.save_for_backward(dlogits)
        return loss

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_out):
        (dlogits,) = ctx.saved_tensors
        return dlogits * grad_out.to(dlogits.dtype), None, None


def cross_entropy(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """``F.cross_entropy(logits.float(), labels)`` (mean).  Loss and d(logits) come out of one
    pass over the logits (K12).  Rows equal to ``ignore_index`` contribute 0 and the mean is
    over all rows."""
    return _CrossEntropy.apply(logits, labels.contiguous(), ignore_index)


# ----------------------------------------------------------------------------- layer norm
class _AddLayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, gamma, beta, eps, p, res_slot=None):
        ctx.res_slot = res_slot
        _req_bf16(x, residual, gamma, beta)
        x = x.contiguous()
        rows, H = x.shape
        dev = x.device
        need_z = residual is not None or p > 0
        z = torch.empty_like(x) if need_z else None
        y = torch.empty_like(x)
        mean = torch.empty(rows, dtype=torch.float32, device=dev)
        rstd = torch.empty(rows, dtype=torch.float32, device=dev)
        salt = _next_salt() if p > 0 else 0
        sw = seed_word(dev) if p > 0 else None
        L.call("ddl_add_layernorm_fwd", x.data_ptr(), L.ptr(None if residual is None else residual.contiguous()), gamma.data_ptr(),
               beta.data_ptr(), L.ptr(z), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), rows, H, float(eps), float(p), salt,
               L.ptr(sw), L.stream_ptr(dev))
        ctx.save_for_backward(z if need_z else x, mean, rstd, gamma)
        ctx.p, ctx.salt, ctx.has_res = p, salt, residual is not None
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        z, mean, rstd, gamma = ctx.saved_tensors
        rows, H = z.shape
        dev = z.device
        dy = dy.contiguous()
        dz = torch.empty_like(z)
        dx = torch.empty_like(z) if ctx.p > 0 else None
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
        sw = seed_word(dev) if ctx.p > 0 else None
        if ctx.needs_input_grad[0]:
            # the column sums of the first input's gradient come out of the same pass: if that input was produced by a
            # linear layer, this IS its bias gradient (ops/gemm.py picks it up instead of running a column-sum pass)
            from . import gemm as G

            dxsum = torch.empty_like(gamma)
            scratch = _scratch(L.lib().ddl_layernorm_bwd_colsum_scratch_bytes(rows, H), dev)
            L.call("ddl_layernorm_bwd_colsum", dy.data_ptr(), z.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(),
                   dz.data_ptr(), L.ptr(dx), dgamma.data_ptr(), dbeta.data_ptr(), dxsum.data_ptr(), L.DDL_BF16, scratch.data_ptr(),
                   scratch.numel(), rows, H, float(ctx.p), ctx.salt, L.ptr(sw), L.stream_ptr(dev))
            G.offer_colsum(dx if dx is not None else dz, dxsum)
        else:
            scratch = _scratch(L.lib().ddl_layernorm_bwd_scratch_bytes(rows, H), dev)
            L.call("ddl_layernorm_bwd", dy.data_ptr(), z.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(), dz.data_ptr(),
                   L.ptr(dx), dgamma.data_ptr(), dbeta.data_ptr(), L.DDL_BF16, scratch.data_ptr(), scratch.numel(), rows, H,
                   float(ctx.p), ctx.salt, L.ptr(sw), L.stream_ptr(dev))
        dres = dz if (ctx.has_res and ctx.needs_input_grad[1]) else None
        if dres is not None and ctx.res_slot is not None:
            ctx.res_slot.put(dres)  # added in the epilogue of the data-gradient GEMM that owns the slot (ops/gemm.py::GradSlot)
            dres = None
        return (dx if dx is not None else dz), dres, dgamma, dbeta, None, None, None


def dropout_add_layernorm(x, residual, gamma, beta, eps: float = 1e-12, p: float = 0.0, residual_grad_slot=None) -> torch.Tensor:
    """``F.layer_norm(F.dropout(x, p) + residual, (H,), gamma, beta, eps)`` on [rows, H] bf16 (K10).
    ``residual_grad_slot``: the gradient of ``residual`` is handed to that :class:`ops.gemm.GradSlot` instead of autograd."""
    return _AddLayerNorm.apply(x, residual, gamma, beta, eps, p, residual_grad_slot)


def layernorm(x, gamma, beta, eps: float = 1e-12) -> torch.Tensor:
    return _AddLayerNorm.apply(x, None, gamma, beta, eps, 0.0)


class _EmbedLayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, tts, word, pos, typ, gamma, beta, eps, p):
        _req_bf16(word, pos, typ, gamma, beta)
        B, S = ids.shape
        H = word.shape[1]
        dev = word.device
        rows = B * S
        z = torch.empty(rows, H, dtype=torch.bfloat16, device=dev)
        y = torch.empty_like(z)
        mean = torch.empty(rows, dtype=torch.float32, device=dev)
        rstd = torch.empty(rows, dtype=torch.float32, device=dev)
        salt = _next_salt() if p > 0 else 0
        sw = seed_word(dev) if p > 0 else None
        ids, tts = ids.contiguous(), tts.contiguous()
        L.call("ddl_embed_layernorm_fwd", ids.data_ptr(), tts.data_ptr(), word.data_ptr(), pos.data_ptr(), typ.data_ptr(),
               gamma.data_ptr(), beta.data_ptr(), z.data_ptr(), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), rows, S, H, float(eps),
               float(p), salt, L.ptr(sw), L.stream_ptr(dev))
        ctx.save_for_backward(ids, tts, z, mean, rstd, gamma)
        ctx.p, ctx.salt, ctx.sizes = p, salt, (word.shape[0], pos.shape[0], typ.shape[0], B, S)
        return y.view(B, S, H)

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        ids, tts, z, mean, rstd, gamma = ctx.saved_tensors
        V, P, T, B, S = ctx.sizes
        rows, H = z.shape
        dev = z.device
        dy = dy.contiguous().view(rows, H)
        if ctx.p > 0:
            masked = torch.empty_like(dy)
            L.call("ddl_dropout", dy.data_ptr(), masked.data_ptr(), dy.numel(), float(ctx.p), ctx.salt, seed_word(dev).data_ptr(),
                   L.stream_ptr(dev))
            dy = masked
        dz = torch.empty_like(z)
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
        scratch = _scratch(L.lib().ddl_layernorm_bwd_scratch_bytes(rows, H), dev)
        L.call("ddl_layernorm_bwd", dy.data_ptr(), z.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(), dz.data_ptr(),
               None, dgamma.data_ptr(), dbeta.data_ptr(), L.DDL_BF16, scratch.data_ptr(), scratch.numel(), rows, H, 0.0, 0, None,
               L.stream_ptr(dev))
        # table gradients (csrc/embed_bwd.cu): word = fp32 scatter-add + one cast, position / type = deterministic sums
        bf = torch.bfloat16
        dword_f = torch.zeros(V, H, dtype=torch.float32, device=dev)
        dword = torch.empty(V, H, dtype=bf, device=dev)
        dpos_full = torch.empty(P, H, dtype=bf, device=dev)
        dtype_ = torch.empty(T, H, dtype=bf, device=dev)
        scr = _scratch(L.lib().ddl_embed_bwd_scratch_bytes(H, T), dev)
        st = L.stream_ptr(dev)
        L.call("ddl_embed_bwd", dz.data_ptr(), ids.data_ptr(), tts.data_ptr(), dword_f.data_ptr(), dpos_full.data_ptr(), dtype_.data_ptr(),
               scr.data_ptr(), scr.numel(), B, S, H, V, P, T, st)
        L.call("ddl_f32_to_bf16", dword_f.data_ptr(), dword.data_ptr(), V * H, 0, st)
        return None, None, dword, dpos_full, dtype_, dgamma, dbeta, None, None


def embed_layernorm(ids, token_type_ids, word, pos, typ, gamma, beta, eps: float, p: float) -> torch.Tensor:
    """``dropout(LN(word[ids] + pos[arange(S)] + typ[tt]))`` -> [B, S, H] bf16."""
    return _EmbedLayerNorm.apply(ids, token_type_ids, word, pos, typ, gamma, beta, eps, p)


# ----------------------------------------------------------------------------- residual fork
# A residual block reads its input twice (main branch + skip), so autograd sums two gradients of that tensor
# with an elementwise kernel (3 passes over the activation) before handing the sum to the op that produced it.
# When the producer is one of our kernels that can take a second gradient (BN+residual+ReLU backward, max-pool
# backward), `residual_fork` hands both gradients over un-summed and the producer's backward adds them while it
# reads them anyway.  The second gradient travels in `_pending_addend`, keyed by the first one's storage; the
# table must be empty again when the backward pass ends, otherwise a gradient would have been dropped and the
# pass raises.
_FUSED_RESIDUAL_GRAD = os.environ.get("DDL_FUSED_RESIDUAL_GRAD", "1") != "0"
_pending_addend = {}


def _take_addend(g: torch.Tensor) -> Optional[torch.Tensor]:
    return _pending_addend.pop((g.data_ptr(), tuple(g.shape)), None) if _pending_addend else None


def _check_addends_consumed() -> None:
    if _pending_addend:
        n = len(_pending_addend)
        _pending_addend.clear()
        raise L.DdlError(f"residual_fork: {n} skip-branch gradient(s) were not consumed by the producer's backward "
                         "(the forked tensor has other consumers?); set DDL_FUSED_RESIDUAL_GRAD=0")


def _accepts_second_grad(t: torch.Tensor) -> torch.Tensor:
    """Tag the output of an op whose backward looks for a pending second gradient."""
    t._ddl_second_grad_ok = True
    return t


class _ResidualFork(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.view_as(x), x.view_as(x)

    @staticmethod
    @once_differentiable
    def backward(ctx, g_main, g_skip):
        if g_main is None or g_skip is None:
            return g_main if g_skip is None else g_skip
        cl = torch.channels_last
        if not (g_main.is_contiguous(memory_format=cl) and g_skip.is_contiguous(memory_format=cl)) or g_main.dtype != g_skip.dtype:
            return g_main + g_skip
        if not _pending_addend:
            torch.autograd.Variable._execution_engine.queue_callback(_check_addends_consumed)
        _pending_addend[(g_main.data_ptr(), tuple(g_main.shape))] = g_skip
        return g_main


def residual_fork(x: torch.Tensor):
    """(x_main, x_skip): two aliases of ``x`` whose gradients are summed inside the backward kernel of the op
    that produced ``x`` (when it can; otherwise this is the identity and autograd sums as usual)."""
    if (_FUSED_RESIDUAL_GRAD and getattr(x, "_ddl_second_grad_ok", False) and x.requires_grad and torch.is_grad_enabled()
            and x.dim() == 4):
        return _ResidualFork.apply(x)
    return x, x


# ----------------------------------------------------------------------------- batch norm
_batches_counted_by_model = False


class batches_counted_once:
    """``with batches_counted_once(model):`` - the ``num_batches_tracked`` counters of every training-mode BatchNorm of
    ``model`` are advanced by ONE multi-tensor kernel here instead of one tiny kernel per layer inside :func:`bn_act`
    (53 launches per ResNet-50 step)."""

    def __init__(self, model: nn.Module):
        # (collected on every call: buffers are replaced by model.to(...) / load_state_dict(assign=True))
        self.counters = [m.num_batches_tracked for m in model.modules()
                         if isinstance(m, nn.modules.batchnorm._BatchNorm) and m.training and m.num_batches_tracked is not None]

    def __enter__(self):
        global _batches_counted_by_model
        self.prev = _batches_counted_by_model
        if self.counters and not self.prev:
            torch._foreach_add_(self.counters, 1)
        _batches_counted_by_model = True
        return self

    def __exit__(self, *exc):
        global _batches_counted_by_model
        _batches_counted_by_model = self.prev
        return False


class _BnAct(torch.autograd.Function):
    """x: NCHW-shaped tensor in channels_last memory (i.e. an NHWC matrix [M, C])."""

    @staticmethod
    def forward(ctx, x, residual, gamma, beta, running_mean, running_var, training, momentum, eps, relu, sums=None, partials=None):
        _req_bf16(x, residual, gamma, beta)
        N, Cc, H, W = x.shape
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        if residual is not None and not residual.is_contiguous(memory_format=torch.channels_last):
            residual = residual.contiguous(memory_format=torch.channels_last)
        M = N * H * W
        dev = x.device
        y = torch.empty_like(x, memory_format=torch.channels_last)
        scale = torch.empty(Cc, dtype=torch.float32, device=dev)
        shift = torch.empty(Cc, dtype=torch.float32, device=dev)
        st = L.stream_ptr(dev)
        mask = None
        if training:
            mean = torch.empty(Cc, dtype=torch.float32, device=dev)
            invstd = torch.empty(Cc, dtype=torch.float32, device=dev)
            if partials is not None:
                # per-CTA partial statistics written by the stem convolution's epilogue: no extra pass over x
                mask = torch.empty(M * Cc // 8, dtype=torch.uint8, device=dev) if relu else None
                L.call("ddl_bn_act_fwd_train_from_partials", x.data_ptr(), L.ptr(residual), gamma.data_ptr(), beta.data_ptr(),
                       L.ptr(running_mean), L.ptr(running_var), y.data_ptr(), L.ptr(mask), mean.data_ptr(), invstd.data_ptr(),
                       scale.data_ptr(), shift.data_ptr(), partials.data_ptr(), partials.numel() // (2 * Cc), M, Cc, float(eps),
                       float(momentum), int(relu), st)
            elif sums is not None:
                # statistics already produced by the convolution's GEMM epilogue: no extra pass over x
                L.call("ddl_bn_act_fwd_train_from_sums", x.data_ptr(), L.ptr(residual), gamma.data_ptr(), beta.data_ptr(),
                       L.ptr(running_mean), L.ptr(running_var), y.data_ptr(), mean.data_ptr(), invstd.data_ptr(), scale.data_ptr(),
                       shift.data_ptr(), sums.data_ptr(), M, Cc, float(eps), float(momentum), int(relu), st)
            else:
                scratch = _scratch(L.lib().ddl_bn_scratch_bytes(M, Cc), dev)
                # one bit per element records the ReLU decision: backward reads it instead of y
                mask = torch.empty(M * Cc // 8, dtype=torch.uint8, device=dev) if relu else None
                L.call("ddl_bn_act_fwd_train_mask", x.data_ptr(), L.ptr(residual), gamma.data_ptr(), beta.data_ptr(),
                       L.ptr(running_mean), L.ptr(running_var), y.data_ptr(), L.ptr(mask), mean.data_ptr(), invstd.data_ptr(),
                       scale.data_ptr(), shift.data_ptr(), scratch.data_ptr(), scratch.numel(), M, Cc, float(eps), float(momentum),
                       int(relu), st)
            ctx.save_for_backward(x, y if (relu and mask is None) else None, gamma, mean, invstd, mask)
        else:
            L.call("ddl_bn_act_fwd_eval", x.data_ptr(), L.ptr(residual), gamma.data_ptr(), beta.data_ptr(), running_mean.data_ptr(),
                   running_var.data_ptr(), y.data_ptr(), scale.data_ptr(), shift.data_ptr(), M, Cc, float(eps), int(relu), st)
            ctx.save_for_backward(x, y if relu else None, gamma, scale, None, None)
        ctx.training, ctx.relu, ctx.has_res, ctx.M = training, relu, residual is not None, M
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, y, gamma, mean, invstd, mask = ctx.saved_tensors
        if not ctx.training:
            raise L.DdlError("backward through eval-mode fused BatchNorm is not supported; use impl='torch' for frozen BN")
        Cc = x.shape[1]
        dev = x.device
        dy2 = _take_addend(dy)
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        if dy2 is not None and not (ctx.relu and ctx.has_res):
            dy, dy2 = dy + dy2, None  # only the relu + residual kernel takes two gradients
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dres = None
        if ctx.has_res:
            dres = torch.empty_like(x, memory_format=torch.channels_last) if ctx.relu else dy
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
        scratch = _scratch(L.lib().ddl_bn_scratch_bytes(ctx.M, Cc), dev)
        L.call("ddl_bn_act_bwd_ex", dy.data_ptr(), L.ptr(dy2), L.ptr(y), L.ptr(mask), x.data_ptr(), gamma.data_ptr(), mean.data_ptr(),
               invstd.data_ptr(), dx.data_ptr(), L.ptr(dres) if ctx.relu else None, dgamma.data_ptr(), dbeta.data_ptr(), L.DDL_BF16,
               scratch.data_ptr(), scratch.numel(), ctx.M, Cc, int(ctx.relu), L.stream_ptr(dev))
        return dx, dres, dgamma, dbeta, None, None, None, None, None, None, None, None


def bn_act(x, bn: nn.BatchNorm2d, relu: bool = True, residual: Optional[torch.Tensor] = None,
           sums: Optional[torch.Tensor] = None, partials: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``relu(batch_norm(x) + residual)`` with torch's training/eval semantics (running stats
    updated with the unbiased variance and ``bn.momentum``)."""
    training = bn.training or bn.running_mean is None
    if bn.running_mean is not None and (bn.running_mean.dtype != torch.float32 or bn.running_var.dtype != torch.float32):
        raise L.DdlError("fused BatchNorm keeps running_mean / running_var in float32 (got "
                         f"{bn.running_mean.dtype}); convert the model with models.prepare_for_fused() instead of .to(bfloat16)")
    if training and bn.num_batches_tracked is not None and not _batches_counted_by_model:
        bn.num_batches_tracked.add_(1)
    y = _BnAct.apply(x, residual, bn.weight, bn.bias, bn.running_mean, bn.running_var, training,
                     0.1 if bn.momentum is None else bn.momentum, bn.eps, relu, sums if training else None,
                     partials if training else None)
    if training and relu and residual is not None:
        _accepts_second_grad(y)  # a block output: the next block may fork it (see residual_fork)
    return y


class _MaxPool3x3s2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _req_bf16(x)
        N, Cc, H, W = x.shape
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
        y = torch.empty((N, Cc, OH, OW), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        idx = torch.empty(N * OH * OW * Cc, dtype=torch.uint8, device=x.device)
        L.call("ddl_maxpool3x3s2_fwd", x.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W, Cc, L.stream_ptr(x.device))
        ctx.save_for_backward(idx)
        ctx.shape = (N, Cc, H, W)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        N, Cc, H, W = ctx.shape
        dy2 = _take_addend(dy)
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((N, Cc, H, W), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        L.call("ddl_maxpool3x3s2_bwd_ex", dy.data_ptr(), L.ptr(dy2), idx.data_ptr(), dx.data_ptr(), N, H, W, Cc, L.stream_ptr(dy.device))
        return dx


class _BnReluMaxPool(torch.autograd.Function):
    """maxpool3x3s2(relu(batch_norm(x))) for the ResNet stem in one pass (csrc/bn_act.cu): the normalised full-resolution
    activation is never materialised; statistics come from the convolution epilogue's per-CTA partials."""

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, momentum, eps, partials):
        _req_bf16(x, gamma, beta)
        N, Cc, H, W = x.shape
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        dev = x.device
        y = torch.empty((N, Cc, H // 2, W // 2), dtype=x.dtype, device=dev, memory_format=torch.channels_last)
        codes = torch.empty(N * (H // 2) * (W // 2) * Cc // 2, dtype=torch.uint8, device=dev)
        mean, invstd, scale, shift = (torch.empty(Cc, dtype=torch.float32, device=dev) for _ in range(4))
        L.call("ddl_bn_relu_maxpool_fwd_train", x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), L.ptr(running_mean), L.ptr(running_var),
               y.data_ptr(), codes.data_ptr(), mean.data_ptr(), invstd.data_ptr(), scale.data_ptr(), shift.data_ptr(), partials.data_ptr(),
               partials.numel() // (2 * Cc), N, H, W, Cc, float(eps), float(momentum), L.stream_ptr(dev))
        ctx.save_for_backward(x, gamma, beta, mean, invstd, codes, y)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, gamma, beta, mean, invstd, codes, y = ctx.saved_tensors
        N, Cc, H, W = x.shape
        dev = x.device
        dy2 = _take_addend(dy)
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
        scratch = _scratch(L.lib().ddl_bn_relu_maxpool_bwd_scratch_bytes(N, H, W, Cc), dev)
        L.call("ddl_bn_relu_maxpool_bwd", dy.data_ptr(), L.ptr(dy2), y.data_ptr(), codes.data_ptr(), x.data_ptr(), gamma.data_ptr(),
               beta.data_ptr(), mean.data_ptr(), invstd.data_ptr(), dx.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), L.DDL_BF16,
               scratch.data_ptr(), scratch.numel(), N, H, W, Cc, L.stream_ptr(dev))
        return dx, dgamma, dbeta, None, None, None, None, None


_BN_POOL_FUSED = os.environ.get("DDL_BN_POOL", "fused") == "fused"  # DDL_BN_POOL=split: bn_act followed by maxpool3x3s2


def bn_relu_maxpool_usable(x: torch.Tensor, bn: nn.BatchNorm2d) -> bool:
    if not (_BN_POOL_FUSED and x.is_cuda and x.dim() == 4 and x.dtype == torch.bfloat16):
        return False
    if not (bn.training and bn.running_mean is not None and bn.weight is not None and bn.bias is not None):
        return False
    N, Cc, H, W = x.shape
    return bool(L.lib().ddl_bn_relu_maxpool_supported(N, H, W, Cc))


def bn_relu_maxpool(x: torch.Tensor, bn: nn.BatchNorm2d, partials: torch.Tensor) -> torch.Tensor:
    """``max_pool2d(relu(batch_norm(x)), 3, 2, 1)`` in training mode with the statistics taken from ``partials`` (the stem
    convolution's epilogue); torch's running-statistics semantics as in :func:`bn_act`."""
    if bn.running_mean.dtype != torch.float32 or bn.running_var.dtype != torch.float32:
        raise L.DdlError("fused BatchNorm keeps running_mean / running_var in float32; convert the model with models.prepare_for_fused()")
    if bn.num_batches_tracked is not None and not _batches_counted_by_model:
        bn.num_batches_tracked.add_(1)
    return _accepts_second_grad(_BnReluMaxPool.apply(x, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                                                     0.1 if bn.momentum is None else bn.momentum, bn.eps, partials))


def maxpool3x3s2(x: torch.Tensor) -> torch.Tensor:
    """``F.max_pool2d(x, 3, 2, 1)`` on channels-last bf16 (K14)."""
    return _accepts_second_grad(_MaxPool3x3s2.apply(x))


class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _req_bf16(x)
        N, Cc, H, W = x.shape
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        y = torch.empty(N, Cc, dtype=x.dtype, device=x.device)
        L.call("ddl_avgpool_fwd", x.data_ptr(), y.data_ptr(), N, H * W, Cc, L.stream_ptr(x.device))
        ctx.shape = (N, Cc, H, W)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        N, Cc, H, W = ctx.shape
        dy = dy.contiguous()
        dx = torch.empty((N, Cc, H, W), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        L.call("ddl_avgpool_bwd", dy.data_ptr(), dx.data_ptr(), N, H * W, Cc, L.stream_ptr(dy.device))
        return dx


def global_avgpool(x: torch.Tensor) -> torch.Tensor:
    """``torch.flatten(F.adaptive_avg_pool2d(x, 1), 1)`` on channels-last bf16 -> [N, C]."""
    return _GlobalAvgPool.apply(x)


# ----------------------------------------------------------------------------- stem convolution (tcgen05)
_STEM_NATIVE = os.environ.get("DDL_STEM", "tcgen05") == "tcgen05"


def stem_usable(x: torch.Tensor, conv: nn.Conv2d) -> bool:
    """The native stem kernel takes the ResNet stem exactly: 7x7 / stride 2 / pad 3, 3 -> 64 channels, no bias, a
    contiguous NCHW fp32/bf16 network input that needs no gradient, even H and W, W/2 a multiple of 16."""
    if not (_STEM_NATIVE and x.is_cuda and x.dim() == 4 and x.shape[1] == 3 and x.is_contiguous() and not x.requires_grad
            and x.dtype in (torch.float32, torch.bfloat16)):
        return False
    H, W = x.shape[2], x.shape[3]
    return (conv.in_channels == 3 and conv.out_channels == 64 and conv.kernel_size == (7, 7) and conv.stride == (2, 2)
            and conv.padding == (3, 3) and conv.dilation == (1, 1) and conv.groups == 1 and conv.bias is None
            and conv.weight.dtype == torch.bfloat16 and H % 2 == 0 and W % 2 == 0 and H >= 8 and (W // 2) % 16 == 0 and W // 2 <= 256)


class _StemConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight):
        N, _, H, W = x.shape
        dev = x.device
        lib = L.lib()
        st = L.stream_ptr(dev)
        z = torch.empty(lib.ddl_stem_z_bytes(N, H, W) // 2, dtype=torch.bfloat16, device=dev)
        L.call("ddl_stem_s2d", x.data_ptr(), L.dtype_code(x.dtype), z.data_ptr(), N, H, W, st)
        y = torch.empty((N, 64, H // 2, W // 2), dtype=torch.bfloat16, device=dev, memory_format=torch.channels_last)
        partials = torch.empty(lib.ddl_stem_fwd_grid(N, H, W) * 2 * 64, dtype=torch.float32, device=dev)
        w2 = torch.empty(4 * 64 * 64, dtype=torch.bfloat16, device=dev)
        L.call("ddl_stem_conv_fwd", z.data_ptr(), weight.data_ptr(), *weight.stride(), w2.data_ptr(), y.data_ptr(), partials.data_ptr(),
               N, H, W, st)
        ctx.save_for_backward(z)
        ctx.geom = (N, H, W)
        ctx.wlike = weight  # only its shape / strides are used
        ctx.mark_non_differentiable(partials)
        ctx.set_materialize_grads(False)  # no zero-filled gradient tensor for the statistics output on every backward
        return y, partials

    @staticmethod
    @once_differentiable
    def backward(ctx, dy, _unused):
        if dy is None:  # (grads are not materialised: the convolution's output did not reach the loss)
            return (None,) * 2
        (z,) = ctx.saved_tensors
        N, H, W = ctx.geom
        dev = z.device
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dw = torch.empty_like(ctx.wlike)
        scratch = _scratch(L.lib().ddl_stem_wgrad_scratch_bytes(N, H, W), dev)
        L.call("ddl_stem_conv_wgrad", z.data_ptr(), dy.data_ptr(), dw.data_ptr(), *dw.stride(), scratch.data_ptr(), scratch.numel(),
               N, H, W, L.stream_ptr(dev))
        return None, dw


def stem_conv(x: torch.Tensor, weight: torch.Tensor):
    """``F.conv2d(x, weight, stride=2, padding=3)`` for the 3 -> 64 channel 7x7 stem on the tcgen05 kernel
    (``csrc/stem_conv.cu``).  Returns ``(y, partials)``: y is NHWC bf16 (NCHW-shaped, channels-last), ``partials`` the
    per-CTA BatchNorm statistics of y for :func:`bn_act` (``partials=``)."""
    return _StemConv.apply(x, weight)


def to_nhwc_bf16(x: torch.Tensor, pad_channels: int = 0) -> torch.Tensor:
    """Network input -> bf16 channels-last.  ``pad_channels=8`` (native kernel, ``csrc/pool.cu``) also
    zero-pads the channel dimension to 8 so that the stem convolution is a tensor-core implicit GEMM;
    use it together with :func:`pad_in_channels` on the weight."""
    if (pad_channels == 8 and x.is_cuda and x.dim() == 4 and x.shape[1] < 8 and not x.requires_grad
            and x.dtype in (torch.float32, torch.bfloat16) and x.is_contiguous()):
        N, C, H, W = x.shape
        y = torch.empty((N, 8, H, W), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
        L.call("ddl_nchw_to_nhwc8_bf16", x.data_ptr(), L.dtype_code(x.dtype), y.data_ptr(), N, C, H * W, L.stream_ptr(x.device))
        return y
    return x.to(dtype=torch.bfloat16, memory_format=torch.channels_last)


def pad_in_channels(weight: torch.Tensor, cin: int) -> torch.Tensor:
    """Zero-pad a convolution weight [Cout, Cin, kh, kw] to ``cin`` input channels (differentiable; the
    gradient of the padding is dropped), keeping the channels-last layout."""
    if weight.shape[1] >= cin:
        return weight
    w = F.pad(weight, (0, 0, 0, 0, 0, cin - weight.shape[1]))
    return w.contiguous(memory_format=torch.channels_last)


class _Conv1x1(torch.autograd.Function):
    """1x1 / stride-1 convolution on NHWC == GEMM [N*H*W, Cin] x [Cout, Cin]^T on the tcgen05 kernel, with the
    BatchNorm statistics (per-channel sum, sum of squares) accumulated in the GEMM epilogue."""

    @staticmethod
    def forward(ctx, x, weight, want_stats):
        from . import gemm as G

        N, Cin, H, W = x.shape
        Cout = weight.shape[0]
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        M = N * H * W
        x2 = x.permute(0, 2, 3, 1).reshape(M, Cin)          # view: channels-last memory is [M, Cin]
        w2 = weight.reshape(Cout, Cin)
        sums = torch.zeros(2, Cout, dtype=torch.float32, device=x.device) if want_stats else None
        y2 = G.gemm_nt(x2, w2, M, Cout, Cin, col_stats=sums)
        ctx.save_for_backward(x2, w2)
        ctx.shape = (N, Cin, H, W, Cout)
        y = y2.view(N, H, W, Cout).permute(0, 3, 1, 2)       # NCHW-shaped, channels-last strides
        if want_stats:
            ctx.mark_non_differentiable(sums)
            ctx.set_materialize_grads(False)
            return y, sums
        return y, None

    @staticmethod
    @once_differentiable
    def backward(ctx, dy, _dsums):
        if dy is None:  # (grads are not materialised: the convolution's output did not reach the loss)
            return (None,) * 3
        from . import gemm as G

        x2, w2 = ctx.saved_tensors
        N, Cin, H, W, Cout = ctx.shape
        M = N * H * W
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dy2 = dy.permute(0, 2, 3, 1).reshape(M, Cout)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = G.gemm_nt(dy2, w2, M, Cin, Cout, b_mn=True).view(N, H, W, Cin).permute(0, 3, 1, 2)
        if ctx.needs_input_grad[1]:
            dw = G._wgrad(dy2, x2).view(Cout, Cin, 1, 1)
        return dx, dw, None


def _is_gemm_conv(x, conv: nn.Conv2d) -> bool:
    return (_CONV1X1_GEMM and conv.kernel_size == (1, 1) and conv.stride == (1, 1) and conv.padding == (0, 0) and conv.groups == 1
            and conv.bias is None and x.dtype == torch.bfloat16 and conv.in_channels % 8 == 0 and conv.out_channels % 8 == 0)


import os as _os

# Measured on B200 (ResNet-50, batch 256): with the 1x1 convolutions on the tcgen05 GEMM the step takes 31.6 ms against
# 25.0 ms with the library convolution - at K = 64..512 the GEMM is bound by its epilogue (one k-block per 128 x 256
# tile), not by the tensor cores.  The GEMM route (with the BN-statistics epilogue) therefore stays opt-in:
# DDL_CONV1X1=gemm.
_CONV1X1_GEMM = _os.environ.get("DDL_CONV1X1", "cudnn") == "gemm"


# Weight-stationary tcgen05 kernel for the skinny 1x1 convolutions (csrc/conv1x1.cu): forward + BatchNorm statistics in
# one pass; DDL_CONV1X1_WS=0 sends these layers back to the library.
_CONV1X1_WS = _os.environ.get("DDL_CONV1X1_WS", "1") != "0"


def conv1x1_ws_usable(x: torch.Tensor, conv: nn.Conv2d) -> bool:
    if not (_CONV1X1_WS and x.is_cuda and x.dim() == 4 and x.dtype == torch.bfloat16 and conv.kernel_size == (1, 1)
            and conv.stride == (1, 1) and conv.padding == (0, 0) and conv.groups == 1 and conv.bias is None
            and conv.weight.dtype == torch.bfloat16):
        return False
    N_, _, H, W = x.shape
    return bool(L.lib().ddl_conv1x1_supported(N_ * H * W, conv.out_channels, conv.in_channels))


class _Conv1x1WS(torch.autograd.Function):
    """Forward on the weight-stationary kernel (also yields the BN statistics); both gradients from the library,
    which is within 1.05 - 1.4x of the memory floor on these layers (benchmarks/conv_bench.py)."""

    @staticmethod
    def forward(ctx, x, weight):
        Nb, Cin, H, W = x.shape
        Cout = weight.shape[0]
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        w2 = weight.reshape(Cout, Cin)  # [Cout, Cin, 1, 1] is [Cout][Cin] in memory in either layout
        if not w2.is_contiguous():
            w2 = w2.contiguous()
        M = Nb * H * W
        dev = x.device
        y = torch.empty((Nb, Cout, H, W), dtype=torch.bfloat16, device=dev, memory_format=torch.channels_last)
        partials = torch.empty(L.lib().ddl_conv1x1_grid(M, Cout) * 2 * Cout, dtype=torch.float32, device=dev)
        L.call("ddl_conv1x1_fwd", x.data_ptr(), w2.data_ptr(), y.data_ptr(), partials.data_ptr(), M, Cout, Cin, L.stream_ptr(dev))
        ctx.save_for_backward(x, weight)
        ctx.mark_non_differentiable(partials)
        ctx.set_materialize_grads(False)  # no zero-filled gradient tensor for the statistics output on every backward
        return y, partials

    @staticmethod
    @once_differentiable
    def backward(ctx, dy, _unused):
        if dy is None:  # (grads are not materialised: the convolution's output did not reach the loss)
            return (None,) * 2
        x, weight = ctx.saved_tensors
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx, dw, _ = torch.ops.aten.convolution_backward(dy, x, weight, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1,
                                                        [ctx.needs_input_grad[0], ctx.needs_input_grad[1], False])
        return dx, dw


# ----------------------------------------------------------------------------- implicit-GEMM convolutions (tcgen05)
# csrc/conv_igemm.cu: forward (+ BatchNorm statistics), data gradient and weight gradient of the 1x1 / 3x3 layers.
# DDL_CONV=cudnn sends every convolution except the stem back to the library (comparison runs only).
_CONV_NATIVE = _os.environ.get("DDL_CONV", "native") == "native"


def _conv_geom(x: torch.Tensor, conv: nn.Conv2d):
    N_, C, H, W = x.shape
    return (N_, H, W, C, conv.out_channels, conv.kernel_size[0], conv.kernel_size[1], conv.stride[0], conv.padding[0])


def conv_native_usable(x: torch.Tensor, conv: nn.Conv2d) -> bool:
    if not (_CONV_NATIVE and x.is_cuda and x.dim() == 4 and x.dtype == torch.bfloat16 and conv.weight.dtype == torch.bfloat16
            and conv.groups == 1 and conv.bias is None and conv.dilation == (1, 1) and conv.stride[0] == conv.stride[1]
            and conv.padding[0] == conv.padding[1] and conv.padding_mode == "zeros"):
        return False
    return bool(L.lib().ddl_conv_supported(*_conv_geom(x, conv)))


def _weight_krsc(weight: torch.Tensor) -> torch.Tensor:
    """The weight as [Cout][R][S][C] memory (channels-last); 1x1 weights are that in either layout."""
    Cout, C, R, S = weight.shape
    if weight.stride() == (R * S * C, 1, S * C, C) or (R == 1 and S == 1 and weight.stride(0) == C and weight.stride(1) == 1):
        return weight
    return weight.contiguous(memory_format=torch.channels_last)


# ---- weight gradients on a side stream -------------------------------------------------------------------------
# A convolution's weight gradient depends only on (x, dy) and nothing in the backward chain depends on it until the
# gradient is collected.  Issued on a second stream it becomes - inside the whole-step CUDA graph - a parallel branch: its
# persistent, latency-bound CTAs share the SMs with the small memory-bound BatchNorm-backward CTAs of the main branch.
# Its inputs and scratch (x, dy, scratch) are kept alive - and dw lives on as the parameter's .grad - until `join_side_work()` has made the consuming stream wait
# for it; the DDP / FSDP wrappers join before they collect a bucket's gradients, and a callback joins at the end of every
# backward pass.  DDL_WGRAD_STREAM=0 keeps the weight gradients on the main stream.
_WGRAD_STREAM = _os.environ.get("DDL_WGRAD_STREAM", "1") != "0"
_side_streams = {}
_side_pending = []


def _side_stream(device: torch.device) -> torch.cuda.Stream:
    key = (device.type, device.index)
    if key not in _side_streams:
        _side_streams[key] = torch.cuda.Stream(device)
    return _side_streams[key]


def join_side_work() -> None:
    """Make the current stream wait for every weight gradient still running on the side stream."""
    if not _side_pending:
        return
    cur = torch.cuda.current_stream()
    for ev, _keep in _side_pending:
        cur.wait_event(ev)
    _side_pending.clear()


class _ConvNative(torch.autograd.Function):
    """conv2d on NHWC bf16 through the implicit-GEMM kernels; returns (y, partial BatchNorm statistics of y)."""

    @staticmethod
    def forward(ctx, x, weight, stride, pad, use_ws):
        Nb, C, H, W = x.shape
        Cout, _, R, S = weight.shape
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        w = _weight_krsc(weight)
        geom = (Nb, H, W, C, Cout, R, S, stride, pad)
        OH, OW = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
        dev = x.device
        lib = L.lib()
        y = torch.empty((Nb, Cout, OH, OW), dtype=torch.bfloat16, device=dev, memory_format=torch.channels_last)
        if use_ws:
            M = Nb * H * W
            partials = torch.empty(lib.ddl_conv1x1_grid(M, Cout) * 2 * Cout, dtype=torch.float32, device=dev)
            L.call("ddl_conv1x1_fwd", x.data_ptr(), w.data_ptr(), y.data_ptr(), partials.data_ptr(), M, Cout, C, L.stream_ptr(dev))
        else:
            partials = torch.empty(lib.ddl_conv_fwd_groups(*geom) * 2 * Cout, dtype=torch.float32, device=dev)
            L.call("ddl_conv_fwd", x.data_ptr(), w.data_ptr(), y.data_ptr(), partials.data_ptr(), *geom, L.stream_ptr(dev))
        ctx.save_for_backward(x, weight)
        ctx.geom = geom
        ctx.mark_non_differentiable(partials)
        ctx.set_materialize_grads(False)  # no zero-filled gradient tensor for the statistics output on every backward
        return y, partials

    @staticmethod
    @once_differentiable
    def backward(ctx, dy, _unused):
        if dy is None:  # (grads are not materialised: the convolution's output did not reach the loss)
            return (None,) * 5
        x, weight = ctx.saved_tensors
        geom = ctx.geom
        dev = x.device
        st = L.stream_ptr(dev)
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x, memory_format=torch.channels_last)
            Nb, H, W, C, Cout, R, S, stride, pad = geom
            if _CONV1X1_WS and R == 1 and stride == 1 and L.lib().ddl_conv1x1_supported(Nb * H * W, C, Cout):
                # skinny 1x1 layer: weight-stationary kernel with the weight read MN-major
                L.call("ddl_conv1x1_dgrad", dy.data_ptr(), _weight_krsc(weight).data_ptr(), dx.data_ptr(), Nb * H * W, C, Cout, st)
            else:
                L.call("ddl_conv_dgrad", dy.data_ptr(), _weight_krsc(weight).data_ptr(), dx.data_ptr(), *geom, st)
        if ctx.needs_input_grad[1]:
            dw = torch.empty_like(weight)  # same (dense) layout as the parameter: the copy into the gradient bucket is flat
            scratch = _scratch(L.lib().ddl_conv_wgrad_scratch_bytes(*geom), dev)
            sk, sc, sr, ss = dw.stride()
            # The side stream is only safe when nothing on the main stream reads dw before the join: a leaf parameter whose
            # .grad is still None (AccumulateGrad then adopts dw without touching it).  The pending list must not hold dw
            # itself - an extra reference makes AccumulateGrad clone it on the main stream, before the side stream wrote it.
            if _WGRAD_STREAM and weight.is_leaf and weight.grad is None and not weight._backward_hooks:
                main, side = torch.cuda.current_stream(dev), _side_stream(dev)
                ready = torch.cuda.Event()
                ready.record(main)       # dy (and x) are complete on the main stream here
                side.wait_event(ready)
                L.call("ddl_conv_wgrad", x.data_ptr(), dy.data_ptr(), dw.data_ptr(), sk, sc, sr, ss, scratch.data_ptr(), scratch.numel(), *geom,
                       side.cuda_stream)
                done = torch.cuda.Event()
                done.record(side)
                if not _side_pending:
                    torch.autograd.Variable._execution_engine.queue_callback(join_side_work)
                _side_pending.append((done, (x, dy, scratch)))
            else:
                L.call("ddl_conv_wgrad", x.data_ptr(), dy.data_ptr(), dw.data_ptr(), sk, sc, sr, ss, scratch.data_ptr(), scratch.numel(), *geom, st)
        return dx, dw, None, None, None


def conv2d_native(x: torch.Tensor, conv: nn.Conv2d):
    """``F.conv2d(x, conv.weight, stride, padding)`` on the tcgen05 implicit-GEMM kernels.  Returns ``(y, partials)``;
    ``partials`` are the per-CTA BatchNorm statistics of y for :func:`bn_act` (``partials=``)."""
    use_ws = conv1x1_ws_usable(x, conv)
    return _ConvNative.apply(x, conv.weight, conv.stride[0], conv.padding[0], use_ws)


def conv_bn_act(x, conv: nn.Conv2d, bn: nn.BatchNorm2d, relu: bool = True, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """conv -> BN -> (+residual) -> ReLU on NHWC bf16.  The convolution (forward, data gradient, weight gradient) runs on
    the implicit-GEMM tcgen05 kernels of ``csrc/conv_igemm.cu`` (skinny 1x1 forwards on the weight-stationary kernel of
    ``csrc/conv1x1.cu``); its epilogue leaves the BatchNorm statistics of the output, so normalise + residual + ReLU
    (``csrc/bn_act.cu``) is one more pass over y.  Shapes the kernels do not take (channel counts that are not multiples
    of 64, grouped / dilated / biased convolutions) fall back to the library call."""
    if conv_native_usable(x, conv):
        y, partials = conv2d_native(x, conv)
        training = bn.training or bn.running_mean is None
        return bn_act(y, bn, relu=relu, residual=residual, partials=partials if training else None)
    if _is_gemm_conv(x, conv):
        training = bn.training or bn.running_mean is None
        y, sums = _Conv1x1.apply(x, conv.weight, training)
        return bn_act(y, bn, relu=relu, residual=residual, sums=sums)
    if conv1x1_ws_usable(x, conv):
        y, partials = _Conv1x1WS.apply(x, conv.weight)
        return bn_act(y, bn, relu=relu, residual=residual, partials=partials)
    y = conv2d(x, conv)
    return bn_act(y, bn, relu=relu, residual=residual)


def conv2d(x, conv: nn.Conv2d) -> torch.Tensor:
    """Library convolution: only for shapes outside ``ddl_conv_supported`` (e.g. the 3-channel CIFAR stem)."""
    return F.conv2d(x, conv.weight, None, conv.stride, conv.padding, conv.dilation, conv.groups)


# ----------------------------------------------------------------------------- linear / attention
def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, activation: Optional[str] = None,
           residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``act(x @ weight.T + bias) + residual``; runs on the tcgen05 GEMM (``ops/gemm.py``) for bf16
    CUDA operands whose inner dimensions are multiples of 8, otherwise on the library GEMM."""
    from . import gemm

    if gemm.usable(x, weight):
        return gemm.linear(x, weight, bias, activation, residual)
    y = F.linear(x, weight, bias)
    if activation == "gelu":
        y = F.gelu(y)
    elif activation == "relu":
        y = F.relu(y)
    return y if residual is None else y + residual


def attention_qkvpacked(qkv: torch.Tensor, seqlens: Optional[torch.Tensor], p: float = 0.0) -> torch.Tensor:
    """Self attention on packed projections ``qkv`` [B, S, 3, h, d] with right-padding given by
    ``seqlens`` [B]; returns [B, S, h*d]."""
    from . import attention

    return attention.attention_qkvpacked(qkv, seqlens, p)
