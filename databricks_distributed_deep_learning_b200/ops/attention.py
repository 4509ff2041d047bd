"""Self-attention front end (K11)."""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F


def attention_qkvpacked(qkv: torch.Tensor, seqlens: Optional[torch.Tensor], p: float = 0.0) -> torch.Tensor:
    """qkv [B, S, 3, h, d] -> [B, S, h*d].  Library path (SDPA) until the native kernel lands."""
    B, S, _, h, d = qkv.shape
    q, k, v = qkv.permute(2, 0, 3, 1, 4)
    mask = None
    if seqlens is not None:
        keep = torch.arange(S, device=qkv.device)[None, :] < seqlens[:, None]
        mask = keep[:, None, None, :]
    out = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=p)
    return out.transpose(1, 2).reshape(B, S, h * d)
