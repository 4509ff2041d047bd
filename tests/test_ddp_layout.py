"""Layout and gradient-collection details of the DDP wrapper that do not need a process group:
channels-last parameters keep their layout inside the flat buffers, and the
``zero_grad(set_to_none=True)`` path (one multi-tensor copy per bucket) gives the same gradients
as the memset + accumulate path."""
import copy

import torch
import torch.nn as nn

from databricks_distributed_deep_learning_b200.optim import FusedAdamW
from databricks_distributed_deep_learning_b200.parallel.ddp import DistributedDataParallel, flat_view


class _Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = nn.Conv2d(3, 8, 3, padding=1, bias=False)
        self.bn = nn.BatchNorm2d(8)
        self.fc = nn.Linear(8, 5)
        self.unused = nn.Linear(4, 4)

    def forward(self, x):
        y = torch.relu(self.bn(self.conv(x)))
        return self.fc(y.mean((2, 3)))


def _model(channels_last: bool):
    torch.manual_seed(0)
    m = _Net()
    if channels_last:
        m = m.to(memory_format=torch.channels_last)
    return m


def test_flat_view_keeps_channels_last():
    w = torch.randn(8, 3, 3, 3).contiguous(memory_format=torch.channels_last)
    flat = torch.zeros(1024)
    v = flat_view(flat, 64, w)
    assert v.shape == w.shape and v.stride() == w.stride()
    v.copy_(w)
    assert torch.equal(v, w)
    # the flat storage holds the channels-last element order
    assert torch.equal(flat[64 : 64 + w.numel()], w.permute(0, 2, 3, 1).reshape(-1))
    # plain parameters: ordinary view
    b = torch.randn(7)
    assert flat_view(flat, 512, b).stride() == (1,)


def test_flatten_params_preserves_layout_and_values():
    m = _model(True)
    ref = copy.deepcopy(m)
    ddp = DistributedDataParallel(m, flatten_params=True)
    assert m.conv.weight.is_contiguous(memory_format=torch.channels_last)
    assert m.conv.weight.grad is None or m.conv.weight.grad.stride() == m.conv.weight.stride()
    for (k, a), (_, b) in zip(m.state_dict().items(), ref.state_dict().items()):
        assert torch.equal(a, b), k
    assert all(k.startswith("module.") for k in ddp.state_dict())


def _grads(set_to_none: bool, channels_last: bool, accumulate: bool):
    m = _model(channels_last)
    ddp = DistributedDataParallel(m, flatten_params=True, bucket_cap_mb=0.0005)
    torch.manual_seed(1)
    xs = [torch.randn(4, 3, 6, 6) for _ in range(3)]
    out = []
    for step in range(2):
        ddp.zero_grad(set_to_none=set_to_none)
        if accumulate:
            with ddp.no_sync():
                ddp(xs[0]).square().mean().backward()
        ddp(xs[1 + step]).square().mean().backward()
        out.append({n: None if p.grad is None else p.grad.detach().clone() for n, p in m.named_parameters()})
        # gradients live in the flat buckets after a synchronised backward
        for b in ddp.buckets:
            for p, v in zip(b.params, b.views):
                if p.grad is not None:
                    assert p.grad.data_ptr() == v.data_ptr()
    return out


def test_set_to_none_matches_memset_path():
    for cl in (False, True):
        for acc in (False, True):
            a = _grads(False, cl, acc)
            b = _grads(True, cl, acc)
            for ga, gb in zip(a, b):
                for n in ga:
                    if n.startswith("unused"):
                        # never receives a gradient: zero slot in both modes
                        assert gb[n] is not None and float(gb[n].abs().sum()) == 0.0
                        assert float(ga[n].abs().sum()) == 0.0
                    else:
                        torch.testing.assert_close(ga[n], gb[n], rtol=1e-6, atol=1e-7, msg=n)


def test_fused_adamw_state_roundtrip_channels_last():
    m = _model(True)
    ddp = DistributedDataParallel(m, flatten_params=True)
    opt = FusedAdamW(ddp, lr=1e-2, overlap=False)
    x = torch.randn(4, 3, 6, 6)
    for _ in range(2):
        opt.zero_grad(set_to_none=True)
        ddp(x).square().mean().backward()
        opt.step()
    sd = opt.state_dict()
    idx = [i for i, (n, _) in enumerate(m.named_parameters()) if n == "conv.weight"][0]
    # the per-parameter entry is in the parameter's logical (NCHW-indexed) order
    torch.testing.assert_close(sd["state"][idx]["master"], m.conv.weight.detach().float().contiguous(), rtol=0, atol=1e-6)
    m2 = _model(True)
    ddp2 = DistributedDataParallel(m2, flatten_params=True)
    opt2 = FusedAdamW(ddp2, lr=1e-2, overlap=False)
    opt2.load_state_dict(sd)
    torch.testing.assert_close(m2.conv.weight, m.conv.weight)
    torch.testing.assert_close(m2.fc.weight, m.fc.weight)


def test_fsdp_set_to_none_matches_memset_path():
    from databricks_distributed_deep_learning_b200.parallel.fsdp import FullyShardedDataParallel

    def run(set_to_none: bool, accumulate: bool):
        m = _model(False)
        fsdp = FullyShardedDataParallel(m, lr=1e-2)
        torch.manual_seed(1)
        xs = [torch.randn(4, 3, 6, 6) for _ in range(3)]
        for step in range(2):
            fsdp.zero_grad(set_to_none=set_to_none)
            if accumulate:
                with fsdp.no_sync():
                    fsdp(xs[0]).square().mean().backward()
            fsdp(xs[1 + step]).square().mean().backward()
            fsdp.step()
        return {k: v.clone() for k, v in fsdp.full_state_dict().items()}

    for acc in (False, True):
        a, b = run(False, acc), run(True, acc)
        for k in a:
            torch.testing.assert_close(a[k], b[k], rtol=1e-5, atol=1e-6, msg=k)
