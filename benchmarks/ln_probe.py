"""Per-shape device time of the LayerNorm forward / backward / finalize kernels (torch.profiler; shares only)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from databricks_distributed_deep_learning_b200.ops import functional as OF
from torch.profiler import ProfilerActivity, profile
bf = torch.bfloat16
for rows, H in ((16384, 768), (8192, 1024), (16384, 1024), (8192, 768)):
    x = torch.randn(rows, H, device="cuda").to(bf).requires_grad_()
    r = torch.randn(rows, H, device="cuda").to(bf).requires_grad_()
    g = torch.ones(H, device="cuda", dtype=bf).requires_grad_()
    b = torch.zeros(H, device="cuda", dtype=bf).requires_grad_()
    for _ in range(3):
        y = OF.dropout_add_layernorm(x, r, g, b, 1e-12, 0.1); y.backward(torch.randn_like(y))
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(10):
            y = OF.dropout_add_layernorm(x, r, g, b, 1e-12, 0.1); y.backward(torch.randn_like(y))
        torch.cuda.synchronize()
    for e in prof.key_averages():
        if "ln_" in e.key:
            print(rows, H, e.key[:60], round(e.device_time_total / e.count, 2), "us")
