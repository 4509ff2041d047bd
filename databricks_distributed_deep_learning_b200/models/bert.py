"""BERT-base / BERT-large sequence classification for the NLP workload.

Capability source: BASELINE.json:5,9,10 ("Hugging Face BERT-style sequence-
classification fine-tuning", seq 128 DDP and seq 512 FSDP).  The mounted reference's
NLP directory is an empty README, so the architecture follows the
public BERT definition and the *checkpoint key names* follow Hugging Face's
``BertForSequenceClassification`` so that safetensors files move in both directions.

Internally query/key/value are ONE fused parameter (``qkv_weight`` [3H, H]) so that the
projection is a single GEMM with N = 3H (2304 / 3072); state-dict
hooks split / concatenate it to HF's ``attention.self.{query,key,value}`` names.

Execution paths:

* ``impl="torch"`` - ATen ops, fp32 or autocast; CPU plumbing + test oracle;
* ``impl="fused"`` - bf16; linears on the tcgen05/TMEM GEMM with bias/GELU epilogues,
  flash-style attention with the padding mask, fused bias+dropout+residual+LayerNorm,
  fused cross-entropy (K8, K10, K11, K12 in SURVEY.md §2.E.3).
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


# fused path: sum the two gradients of each residual stream in GEMM epilogues (ops/gemm.py::GradSlot); 0 = autograd adds
_FUSE_RESIDUAL_GRAD = os.environ.get("DDL_BERT_FUSE_RESIDUAL_GRAD", "1") != "0"


@dataclass
class BertConfig:
    vocab_size: int = 30522
    hidden_size: int = 768
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    intermediate_size: int = 3072
    max_position_embeddings: int = 512
    type_vocab_size: int = 2
    layer_norm_eps: float = 1e-12
    hidden_dropout_prob: float = 0.1
    attention_probs_dropout_prob: float = 0.1
    num_labels: int = 2
    initializer_range: float = 0.02

    @staticmethod
    def base(**kw) -> "BertConfig":
        return BertConfig(**kw)

    @staticmethod
    def large(**kw) -> "BertConfig":
        d = dict(hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, intermediate_size=4096)
        d.update(kw)
        return BertConfig(**d)

    @staticmethod
    def tiny(**kw) -> "BertConfig":
        d = dict(vocab_size=512, hidden_size=128, num_hidden_layers=2, num_attention_heads=2,
                 intermediate_size=512, max_position_embeddings=128)
        d.update(kw)
        return BertConfig(**d)


class BertEmbeddings(nn.Module):
    def __init__(self, cfg: BertConfig):
        super().__init__()
        self.word_embeddings = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        self.position_embeddings = nn.Embedding(cfg.max_position_embeddings, cfg.hidden_size)
        self.token_type_embeddings = nn.Embedding(cfg.type_vocab_size, cfg.hidden_size)
        self.LayerNorm = nn.LayerNorm(cfg.hidden_size, eps=cfg.layer_norm_eps)
        self.p_drop = cfg.hidden_dropout_prob
        self.impl = "torch"

    def forward(self, input_ids, token_type_ids=None):
        B, S = input_ids.shape
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        if self.impl == "fused" and input_ids.is_cuda:
            from ..ops import functional as OF

            return OF.embed_layernorm(
                input_ids, token_type_ids, self.word_embeddings.weight, self.position_embeddings.weight,
                self.token_type_embeddings.weight, self.LayerNorm.weight, self.LayerNorm.bias,
                self.LayerNorm.eps, self.p_drop if self.training else 0.0,
            )
        pos = torch.arange(S, device=input_ids.device)
        x = self.word_embeddings(input_ids) + self.position_embeddings(pos)[None] + self.token_type_embeddings(token_type_ids)
        x = self.LayerNorm(x)
        return F.dropout(x, self.p_drop, self.training)


class BertSelfAttention(nn.Module):
    """Fused-QKV self attention.  HF key names are produced by the state-dict hooks."""

    def __init__(self, cfg: BertConfig):
        super().__init__()
        H = cfg.hidden_size
        self.num_heads = cfg.num_attention_heads
        self.head_dim = H // cfg.num_attention_heads
        self.qkv_weight = nn.Parameter(torch.empty(3 * H, H))
        self.qkv_bias = nn.Parameter(torch.zeros(3 * H))
        self.p_drop = cfg.attention_probs_dropout_prob
        self.impl = "torch"
        self._register_state_dict_hook(self._split_qkv_hook)
        self.register_load_state_dict_pre_hook(self._merge_qkv_hook)

    @staticmethod
    def _split_qkv_hook(module, state_dict, prefix, local_metadata):
        w = state_dict.pop(prefix + "qkv_weight")
        b = state_dict.pop(prefix + "qkv_bias")
        H = w.shape[1]
        for i, name in enumerate(("query", "key", "value")):
            state_dict[f"{prefix}{name}.weight"] = w[i * H : (i + 1) * H]
            state_dict[f"{prefix}{name}.bias"] = b[i * H : (i + 1) * H]
        return state_dict

    @staticmethod
    def _merge_qkv_hook(module, state_dict, prefix, *args):
        if prefix + "query.weight" in state_dict:
            ws = [state_dict.pop(f"{prefix}{n}.weight") for n in ("query", "key", "value")]
            bs = [state_dict.pop(f"{prefix}{n}.bias") for n in ("query", "key", "value")]
            state_dict[prefix + "qkv_weight"] = torch.cat(ws, 0)
            state_dict[prefix + "qkv_bias"] = torch.cat(bs, 0)

    def forward(self, x, attn_mask_bias, seqlens=None):
        """x: [B, S, H].  attn_mask_bias: [B, S] additive (0 keep / -inf pad) or None."""
        B, S, H = x.shape
        if self.impl == "fused" and x.is_cuda:
            from ..ops import functional as OF

            qkv = OF.linear(x.view(B * S, H), self.qkv_weight, self.qkv_bias)  # [B*S, 3H]
            ctx = OF.attention_qkvpacked(
                qkv.view(B, S, 3, self.num_heads, self.head_dim), seqlens,
                self.p_drop if self.training else 0.0,
            )
            return ctx.view(B, S, H)
        qkv = F.linear(x, self.qkv_weight, self.qkv_bias).view(B, S, 3, self.num_heads, self.head_dim)
        q, k, v = qkv.permute(2, 0, 3, 1, 4)  # [B, h, S, d]
        mask = None if attn_mask_bias is None else attn_mask_bias[:, None, None, :].to(q.dtype)
        ctx = F.scaled_dot_product_attention(q, k, v, attn_mask=mask,
                                             dropout_p=self.p_drop if self.training else 0.0)
        return ctx.transpose(1, 2).reshape(B, S, H)


class BertSelfOutput(nn.Module):
    def __init__(self, cfg: BertConfig, in_features: Optional[int] = None):
        super().__init__()
        self.dense = nn.Linear(in_features or cfg.hidden_size, cfg.hidden_size)
        self.LayerNorm = nn.LayerNorm(cfg.hidden_size, eps=cfg.layer_norm_eps)
        self.p_drop = cfg.hidden_dropout_prob
        self.impl = "torch"

    def forward(self, h, residual):
        if self.impl == "fused" and h.is_cuda:
            from ..ops import functional as OF

            shp = residual.shape
            y = OF.linear(h.reshape(-1, h.shape[-1]), self.dense.weight, self.dense.bias)
            y = OF.dropout_add_layernorm(
                y, residual.reshape(-1, shp[-1]), self.LayerNorm.weight, self.LayerNorm.bias,
                self.LayerNorm.eps, self.p_drop if self.training else 0.0,
            )
            return y.view(shp)
        y = F.dropout(self.dense(h), self.p_drop, self.training)
        return self.LayerNorm(y + residual)


class BertAttention(nn.Module):
    def __init__(self, cfg: BertConfig):
        super().__init__()
        self.self = BertSelfAttention(cfg)
        self.output = BertSelfOutput(cfg)

    def forward(self, x, attn_mask_bias, seqlens=None):
        return self.output(self.self(x, attn_mask_bias, seqlens), x)


class BertIntermediate(nn.Module):
    def __init__(self, cfg: BertConfig):
        super().__init__()
        self.dense = nn.Linear(cfg.hidden_size, cfg.intermediate_size)
        self.impl = "torch"

    def forward(self, x):
        if self.impl == "fused" and x.is_cuda:
            from ..ops import functional as OF

            shp = x.shape
            y = OF.linear(x.reshape(-1, shp[-1]), self.dense.weight, self.dense.bias, activation="gelu")
            return y.view(*shp[:-1], -1)
        return F.gelu(self.dense(x))


class BertLayer(nn.Module):
    def __init__(self, cfg: BertConfig):
        super().__init__()
        self.attention = BertAttention(cfg)
        self.intermediate = BertIntermediate(cfg)
        self.output = BertSelfOutput(cfg, in_features=cfg.intermediate_size)
        self.impl = "torch"

    def forward(self, x, attn_mask_bias, seqlens=None):
        if self.impl == "fused" and x.is_cuda:
            return self._fused_forward(x, seqlens)
        a = self.attention(x, attn_mask_bias, seqlens)
        return self.output(self.intermediate(a), a)

    def _fused_forward(self, x, seqlens):
        """The whole encoder layer on native kernels: 4 tcgen05 GEMM launches forward
        (QKV, attention output, FFN up + GELU, FFN down), attention, 2 LayerNorm launches.
        With dropout off the residual adds ride in the GEMM epilogues; with dropout on they
        move into the dropout+residual+LayerNorm kernel (the mask must precede the add)."""
        from ..ops import functional as OF
        from ..ops import gemm as G

        B, S, H = x.shape
        x2 = x.reshape(B * S, H)
        sa, so = self.attention.self, self.attention.output
        p_h = so.p_drop if self.training else 0.0
        p_a = sa.p_drop if self.training else 0.0
        # the two gradients of each residual stream are summed in GEMM epilogues, not by autograd (ops/gemm.py::GradSlot)
        fuse = _FUSE_RESIDUAL_GRAD and torch.is_grad_enabled() and x2.requires_grad
        slot1 = G.GradSlot() if fuse else None
        slot2 = G.GradSlot() if fuse else None
        qkv = G.linear(x2, sa.qkv_weight, sa.qkv_bias, input_grad_slot=slot1)
        ctx = OF.attention_qkvpacked(qkv.view(B, S, 3, sa.num_heads, sa.head_dim), seqlens, p_a).reshape(B * S, H)
        ln1, ln2 = so.LayerNorm, self.output.LayerNorm
        if p_h == 0.0:
            z = G.linear(ctx, so.dense.weight, so.dense.bias, residual=x2, residual_grad_slot=slot1)
            a = OF.layernorm(z, ln1.weight, ln1.bias, ln1.eps)
            z2 = G.ffn(a, self.intermediate.dense.weight, self.intermediate.dense.bias, self.output.dense.weight,
                       self.output.dense.bias, residual=a)
            out = OF.layernorm(z2, ln2.weight, ln2.bias, ln2.eps)
        else:
            y = G.linear(ctx, so.dense.weight, so.dense.bias)
            a = OF.dropout_add_layernorm(y, x2, ln1.weight, ln1.bias, ln1.eps, p_h, residual_grad_slot=slot1)
            y2 = G.ffn(a, self.intermediate.dense.weight, self.intermediate.dense.bias, self.output.dense.weight,
                       self.output.dense.bias, input_grad_slot=slot2)
            out = OF.dropout_add_layernorm(y2, a, ln2.weight, ln2.bias, ln2.eps, p_h, residual_grad_slot=slot2)
        return out.view(B, S, H)


class BertEncoder(nn.Module):
    def __init__(self, cfg: BertConfig):
        super().__init__()
        self.layer = nn.ModuleList([BertLayer(cfg) for _ in range(cfg.num_hidden_layers)])

    def forward(self, x, attn_mask_bias, seqlens=None):
        for l in self.layer:
            x = l(x, attn_mask_bias, seqlens)
        return x


class BertPooler(nn.Module):
    def __init__(self, cfg: BertConfig):
        super().__init__()
        self.dense = nn.Linear(cfg.hidden_size, cfg.hidden_size)

    def forward(self, x):
        return torch.tanh(self.dense(x[:, 0]))


class BertModel(nn.Module):
    def __init__(self, cfg: BertConfig):
        super().__init__()
        self.embeddings = BertEmbeddings(cfg)
        self.encoder = BertEncoder(cfg)
        self.pooler = BertPooler(cfg)

    def forward(self, input_ids, attention_mask=None, token_type_ids=None):
        x = self.embeddings(input_ids, token_type_ids)
        bias, seqlens = None, None
        if attention_mask is not None:
            # right-padded masks (HF tokenizers' default) are summarised by their lengths for
            # the fused kernel; the additive bias serves the ATen path
            seqlens = attention_mask.sum(-1, dtype=torch.int32)
            bias = torch.zeros(attention_mask.shape, dtype=x.dtype, device=x.device)
            bias = bias.masked_fill(attention_mask == 0, float("-inf"))
        x = self.encoder(x, bias, seqlens)
        return x, self.pooler(x)


class BertForSequenceClassification(nn.Module):
    """``forward`` returns ``(loss, logits)`` when labels are given, else ``logits``."""

    def __init__(self, cfg: BertConfig, impl: str = "torch"):
        super().__init__()
        self.config = cfg
        self.bert = BertModel(cfg)
        self.p_drop = cfg.hidden_dropout_prob
        self.classifier = nn.Linear(cfg.hidden_size, cfg.num_labels)
        self.apply(self._init)
        self.set_impl(impl)

    def _init(self, m):
        std = self.config.initializer_range
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, 0.0, std)
            nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, 0.0, std)
        elif isinstance(m, nn.LayerNorm):
            nn.init.ones_(m.weight)
            nn.init.zeros_(m.bias)
        elif isinstance(m, BertSelfAttention):
            nn.init.normal_(m.qkv_weight, 0.0, std)
            nn.init.zeros_(m.qkv_bias)

    def set_impl(self, impl: str) -> "BertForSequenceClassification":
        if impl not in ("torch", "fused"):
            raise ValueError(impl)
        self.impl = impl
        for m in self.modules():
            if hasattr(m, "impl") and m is not self:
                m.impl = impl
        return self

    def forward(self, input_ids, attention_mask=None, token_type_ids=None, labels=None):
        _, pooled = self.bert(input_ids, attention_mask, token_type_ids)
        pooled = F.dropout(pooled, self.p_drop, self.training)
        logits = self.classifier(pooled)
        if labels is None:
            return logits
        if self.impl == "fused" and logits.is_cuda:
            from ..ops import functional as OF

            loss = OF.cross_entropy(logits, labels)
        else:
            loss = F.cross_entropy(logits.float(), labels)
        return loss, logits


def bert_base(num_labels: int = 2, impl: str = "torch", **kw) -> BertForSequenceClassification:
    return BertForSequenceClassification(BertConfig.base(num_labels=num_labels, **kw), impl)


def bert_large(num_labels: int = 2, impl: str = "torch", **kw) -> BertForSequenceClassification:
    return BertForSequenceClassification(BertConfig.large(num_labels=num_labels, **kw), impl)


def bert_tiny(num_labels: int = 2, impl: str = "torch", **kw) -> BertForSequenceClassification:
    return BertForSequenceClassification(BertConfig.tiny(num_labels=num_labels, **kw), impl)
