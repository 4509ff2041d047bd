"""Self-attention front end (K11): tcgen05 kernels in ``csrc/attention.cu``.

``attention_qkvpacked(qkv [B, S, 3, h, 64], seqlens [B] | None, p)`` -> ``[B, S, h*64]``.
The native path takes head_dim 64 and S in {128, 256, 384, 512} (BERT at seq 128 / 512);
anything else goes to the library (SDPA), which is also the oracle in the tests:

    q, k, v = qkv.unbind(2); mask = arange(S) < seqlens[:, None]
    softmax(q k^T / 8 + mask_bias) (dropout) v
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F
from torch.autograd.function import once_differentiable

from . import _lib as L


def native_ok(qkv: torch.Tensor) -> bool:
    B, S, three, h, d = qkv.shape
    return qkv.is_cuda and qkv.dtype == torch.bfloat16 and d == 64 and S % 128 == 0 and 128 <= S <= 512 and qkv.is_contiguous()


class _Attention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, seqlens, p):
        from .functional import _next_salt, seed_word

        B, S, _, h, d = qkv.shape
        dev = qkv.device
        out = torch.empty(B, S, h * d, dtype=torch.bfloat16, device=dev)
        lse = torch.empty(B, h, S, dtype=torch.float32, device=dev)
        sl = None if seqlens is None else seqlens.to(torch.int32).contiguous()
        salt = _next_salt() if p > 0 else 0
        sw = seed_word(dev) if p > 0 else None
        L.call("ddl_attention_fwd", qkv.data_ptr(), L.ptr(sl), out.data_ptr(), lse.data_ptr(), B, S, h, float(p), salt, L.ptr(sw),
               L.stream_ptr(dev))
        ctx.save_for_backward(qkv, out, lse, sl)
        ctx.p, ctx.salt = p, salt
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        from .functional import seed_word

        qkv, out, lse, sl = ctx.saved_tensors
        B, S, _, h, d = qkv.shape
        dev = qkv.device
        dout = dout.contiguous()
        dqkv = torch.empty_like(qkv)
        nbytes = L.lib().ddl_attention_bwd_scratch_bytes(B, S, h)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        sw = seed_word(dev) if ctx.p > 0 else None
        L.call("ddl_attention_bwd", qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), L.ptr(sl), dqkv.data_ptr(),
               scratch.data_ptr(), nbytes, B, S, h, float(ctx.p), ctx.salt, L.ptr(sw), L.stream_ptr(dev))
        return dqkv, None, None


def attention_reference(qkv: torch.Tensor, seqlens: Optional[torch.Tensor], p: float = 0.0) -> torch.Tensor:
    """Library path (SDPA): used for shapes the native kernel does not take, and as test oracle."""
    B, S, _, h, d = qkv.shape
    q, k, v = qkv.permute(2, 0, 3, 1, 4)
    mask = None
    if seqlens is not None:
        keep = torch.arange(S, device=qkv.device)[None, :] < seqlens[:, None]
        mask = keep[:, None, None, :]
    out = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=p)
    return out.transpose(1, 2).reshape(B, S, h * d)


def attention_qkvpacked(qkv: torch.Tensor, seqlens: Optional[torch.Tensor], p: float = 0.0) -> torch.Tensor:
    if native_ok(qkv):
        return _Attention.apply(qkv, seqlens, p)
    return attention_reference(qkv, seqlens, p)
