#!/usr/bin/env python
"""Short single-GPU programs for `ncu --set full` captures (one kernel family per mode).

    python benchmarks/ncu_target.py gemm|attention|bn|ln|adamw|stem|conv1x1
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "gemm"
    torch.manual_seed(0)
    bf = torch.bfloat16
    if mode == "gemm":
        from databricks_distributed_deep_learning_b200.ops import gemm as G

        M, N, K = 16384, 3072, 768  # BERT-base FFN up-projection
        a = torch.randn(M, K, device="cuda").to(bf)
        b = torch.randn(N, K, device="cuda").to(bf)
        bias = torch.randn(N, device="cuda").to(bf)
        pre = torch.empty(M, N, device="cuda", dtype=bf)
        for _ in range(4):
            G.gemm_nt(a, b, M, N, K, bias=bias, aux=pre, epilogue=G.EPI_GELU)
    elif mode == "attention":
        from databricks_distributed_deep_learning_b200.ops import attention as A

        for S, B, h in ((128, 128, 12), (512, 16, 16)):
            qkv = torch.randn(B, S, 3, h, 64, device="cuda").to(bf).requires_grad_()
            for _ in range(2):
                out = A.attention_qkvpacked(qkv, None, 0.1)
                out.backward(torch.randn_like(out))
    elif mode == "bn":
        from databricks_distributed_deep_learning_b200.ops import functional as OF

        bn = torch.nn.BatchNorm2d(256).cuda()
        bn.weight.data = bn.weight.data.to(bf)
        bn.bias.data = bn.bias.data.to(bf)
        x = torch.randn(256, 256, 56, 56, device="cuda").to(bf).contiguous(memory_format=torch.channels_last).requires_grad_()
        res = torch.randn_like(x).contiguous(memory_format=torch.channels_last).requires_grad_()
        for _ in range(2):
            y = OF.bn_act(x, bn, relu=True, residual=res)
            y.backward(torch.randn_like(y))
    elif mode == "ln":
        from databricks_distributed_deep_learning_b200.ops import functional as OF

        H = 768
        x = torch.randn(16384, H, device="cuda").to(bf).requires_grad_()
        r = torch.randn(16384, H, device="cuda").to(bf).requires_grad_()
        g = torch.ones(H, device="cuda", dtype=bf).requires_grad_()
        b = torch.zeros(H, device="cuda", dtype=bf).requires_grad_()
        for _ in range(2):
            y = OF.dropout_add_layernorm(x, r, g, b, 1e-12, 0.1)
            y.backward(torch.randn_like(y))
    elif mode == "adamw":
        from databricks_distributed_deep_learning_b200.ops import _lib as L

        n = 8 * 4_000_000
        gr = torch.randn(n, device="cuda").to(bf)
        w, m, v = torch.randn(n, device="cuda"), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
        p = torch.empty(n, device="cuda", dtype=bf)
        hyper = torch.tensor([1e-3, 0.9, 0.999, 1e-8, 0.01, 1.0, 0, 0], device="cuda")
        for _ in range(3):
            L.call("ddl_adamw", gr.data_ptr(), 1, w.data_ptr(), m.data_ptr(), v.data_ptr(), p.data_ptr(), 1, n, hyper.data_ptr(), 1.0, L.stream_ptr())
    elif mode == "stem":
        from databricks_distributed_deep_learning_b200.ops import functional as OF

        w = torch.randn(64, 3, 7, 7, device="cuda").to(bf).contiguous(memory_format=torch.channels_last).requires_grad_()
        x = torch.randn(256, 3, 224, 224, device="cuda")
        for _ in range(2):
            y, _ = OF.stem_conv(x, w)
            y.backward(torch.randn_like(y))
    elif mode == "stempool":
        # the whole ImageNet-style stem: space-to-depth + 7x7/2 convolution (+ statistics), BatchNorm + ReLU + max pool in one pass, backward
        from databricks_distributed_deep_learning_b200.ops import functional as OF

        w = torch.randn(64, 3, 7, 7, device="cuda").to(bf).contiguous(memory_format=torch.channels_last).requires_grad_()
        bn = torch.nn.BatchNorm2d(64).cuda()
        bn.weight.data = bn.weight.data.to(bf)
        bn.bias.data = bn.bias.data.to(bf)
        x = torch.randn(256, 3, 224, 224, device="cuda")
        for _ in range(2):
            y, part = OF.stem_conv(x, w)
            p = OF.bn_relu_maxpool(y, bn, part)
            p.backward(torch.randn_like(p))
    elif mode == "conv1x1":
        from databricks_distributed_deep_learning_b200.ops import functional as OF

        for cin, cout, hw in ((64, 256, 56), (256, 64, 56), (128, 512, 28)):
            conv = torch.nn.Conv2d(cin, cout, 1, bias=False).cuda().to(bf).to(memory_format=torch.channels_last)
            x = torch.randn(256, cin, hw, hw, device="cuda").to(bf).contiguous(memory_format=torch.channels_last)
            for _ in range(2):
                OF._Conv1x1WS.apply(x, conv.weight)
    elif mode == "conv":
        # implicit-GEMM convolutions: forward (+ statistics), data gradient, weight gradient on three layer shapes
        from databricks_distributed_deep_learning_b200.ops import _lib as L

        lib = L.lib()
        if os.environ.get("DDL_CONV_PAIR"):
            L.call("ddl_conv_set_tuning", 0, int(os.environ["DDL_CONV_PAIR"]))
        shapes = [(256, 14, 1024, 1, 1), (256, 14, 256, 3, 1), (64, 56, 64, 3, 1)]
        if len(sys.argv) > 2:
            shapes = [tuple(int(v) for v in sh.split(",")) for sh in sys.argv[2].split(";")]
        for cin, hw, cout, k, s in shapes:
            N, pad = 256, k // 2
            oh = (hw + 2 * pad - k) // s + 1
            cl = torch.channels_last
            x = torch.randn(N, cin, hw, hw, device="cuda").to(bf).contiguous(memory_format=cl)
            w = torch.randn(cout, cin, k, k, device="cuda").to(bf).contiguous(memory_format=cl)
            dy = torch.randn(N, cout, oh, oh, device="cuda").to(bf).contiguous(memory_format=cl)
            y, dx, dw = torch.empty_like(dy), torch.empty_like(x), torch.empty_like(w)
            geom = (N, hw, hw, cin, cout, k, k, s, pad)
            part = torch.empty(lib.ddl_conv_fwd_groups(*geom) * 2 * cout, dtype=torch.float32, device="cuda")
            scr = torch.empty(max(lib.ddl_conv_wgrad_scratch_bytes(*geom), 16), dtype=torch.uint8, device="cuda")
            st = L.stream_ptr()
            for _ in range(1 if len(sys.argv) > 3 else 2):
                L.call("ddl_conv_fwd", x.data_ptr(), w.data_ptr(), y.data_ptr(), part.data_ptr(), *geom, st)
                L.call("ddl_conv_dgrad", dy.data_ptr(), w.data_ptr(), dx.data_ptr(), *geom, st)
                L.call("ddl_conv_wgrad", x.data_ptr(), dy.data_ptr(), dw.data_ptr(), *dw.stride(), scr.data_ptr(), scr.numel(), *geom, st)
    torch.cuda.synchronize()
    print("ok", mode)


if __name__ == "__main__":
    main()
