#!/usr/bin/env python
"""One small call of every single-GPU kernel family, for `compute-sanitizer --tool memcheck`."""
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from databricks_distributed_deep_learning_b200.ops import attention as A  # noqa: E402
from databricks_distributed_deep_learning_b200.ops import functional as OF  # noqa: E402
from databricks_distributed_deep_learning_b200.ops import gemm as G  # noqa: E402

bf = torch.bfloat16
torch.manual_seed(0)
# GEMM: CTA pair, single CTA, MN-major, split-K, epilogues, ragged edges
for (M, N, K) in ((256, 256, 128), (136, 136, 72), (384, 512, 256)):
    a = torch.randn(M, K, device="cuda").to(bf)
    b = torch.randn(N, K, device="cuda").to(bf)
    bias = torch.randn(N, device="cuda").to(bf)
    res = torch.randn(M, N, device="cuda").to(bf)
    pre = torch.empty(M, N, device="cuda", dtype=bf)
    G.gemm_nt(a, b, M, N, K, bias=bias, residual=res)
    G.gemm_nt(a, b, M, N, K, bias=bias, aux=pre, epilogue=G.EPI_GELU)
    G.gemm_nt(a, b, M, N, K, aux=pre, epilogue=G.EPI_DGELU)
    G.gemm_nt(a.t().contiguous(), b.t().contiguous(), M, N, K, a_mn=True, b_mn=True, split_k=2)
    st = torch.zeros(2, N, device="cuda")
    G.gemm_nt(a, b, M, N, K, col_stats=st)
    G.colsum(res)
# attention fwd + bwd, masked, dropout, one and several key tiles
for (B, S, h) in ((2, 128, 2), (1, 384, 2)):
    qkv = torch.randn(B, S, 3, h, 64, device="cuda").to(bf).requires_grad_()
    sl = torch.tensor([S // 2 + 3] + [S] * (B - 1), device="cuda")
    o = A.attention_qkvpacked(qkv, sl, 0.1)
    o.backward(torch.randn_like(o))
# LayerNorm family, cross entropy
x = torch.randn(300, 768, device="cuda").to(bf).requires_grad_()
r = torch.randn(300, 768, device="cuda").to(bf).requires_grad_()
g = torch.ones(768, device="cuda", dtype=bf).requires_grad_()
b2 = torch.zeros(768, device="cuda", dtype=bf).requires_grad_()
OF.dropout_add_layernorm(x, r, g, b2, 1e-12, 0.1).sum().backward()
lg = torch.randn(33, 1000, device="cuda").to(bf).requires_grad_()
OF.cross_entropy(lg, torch.randint(0, 1000, (33,), device="cuda")).backward()
# BatchNorm + pooling
bn = torch.nn.BatchNorm2d(64).cuda()
bn.weight.data, bn.bias.data = bn.weight.data.to(bf), bn.bias.data.to(bf)
xi = torch.randn(4, 64, 28, 28, device="cuda").to(bf).contiguous(memory_format=torch.channels_last).requires_grad_()
y = OF.maxpool3x3s2(OF.bn_act(xi, bn, relu=True, residual=xi.detach()))
OF.global_avgpool(y).sum().backward()
torch.cuda.synchronize()
print("sanitize target ok")
