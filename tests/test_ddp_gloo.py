"""BASELINE.json:7 - ResNet-18 on synthetic 32x32 batches, DDP on CPU/gloo, world_size=2."""
import torch

import _workers as W
from databricks_distributed_deep_learning_b200 import DDP, Distributor
from databricks_distributed_deep_learning_b200.models import resnet18
from databricks_distributed_deep_learning_b200.parallel.ddp import assign_buckets


def test_ddp_matches_single_process_on_summed_batch():
    steps, b = 3, 4
    out = Distributor(num_processes=2, use_gpu=False, timeout_s=600).run(W.w_ddp_resnet18, steps, b, 1.0)
    p0, p1 = out["params"]
    assert torch.equal(p0, p1), "ranks diverged"
    ref = W.single_process_resnet18(steps, b, 2)
    assert torch.allclose(p0, ref, rtol=1e-3, atol=3e-4), float((p0 - ref).abs().max())
    assert out["nbuckets"] > 1
    assert all(k.startswith("module.") for k in out["keys"])


def test_no_sync_accumulates_then_reduces():
    steps, b = 2, 4
    out = Distributor(num_processes=2, use_gpu=False, timeout_s=600).run(W.w_ddp_resnet18, steps, b, 25.0, True)
    p0, p1 = out["params"]
    assert torch.equal(p0, p1)
    ref = W.single_process_resnet18(steps, b, 2)
    assert torch.allclose(p0, ref, rtol=1e-3, atol=3e-4), float((p0 - ref).abs().max())


def test_bucket_assignment_reverse_order_and_caps():
    m = resnet18(num_classes=10, small_input=True)
    params = list(m.parameters())
    buckets = assign_buckets(params, bucket_cap_bytes=4 << 20, first_bucket_cap_bytes=1 << 20)
    seen = [p for b in buckets for p in b.params]
    assert [id(p) for p in seen] == [id(p) for p in reversed(params)]
    assert buckets[0].params[0] is params[-1]
    for b in buckets:
        assert b.numel * 4 <= (4 << 20) or len(b.params) == 1
        assert all(o % 64 == 0 for o in b.offsets)
        assert b.padded_numel >= b.numel and b.padded_numel % 1024 == 0
    assert sum(p.numel() for p in seen) == sum(p.numel() for p in params)


def test_single_process_wrapper_and_grad_views():
    m = resnet18(num_classes=10, small_input=True)
    ddp = DDP(m, bucket_cap_mb=1.0)
    x = torch.randn(2, 3, 32, 32)
    ddp(x).sum().backward()
    for b in ddp.buckets:
        for p, v in zip(b.params, b.views):
            assert p.grad.data_ptr() == v.data_ptr()
    g = ddp.buckets[0].buffer.clone()
    assert g.abs().sum() > 0
    # zero_grad(set_to_none=True) by a user optimizer must be repaired by the hook
    for p in ddp.parameters():
        p.grad = None
    ddp(x).sum().backward()
    assert torch.allclose(ddp.buckets[0].buffer, g)
    ddp.zero_grad()
    assert all(float(b.buffer.abs().sum()) == 0 for b in ddp.buckets)
    assert set(k[len("module."):] for k in ddp.state_dict()) == set(m.state_dict())


def test_profile_comm_reports_every_bucket_and_rolls_back():
    """SURVEY.md section 5 "metrics": per-bucket communication time / bytes / bus bandwidth from a few eager steps."""
    out = Distributor(num_processes=2, use_gpu=False, timeout_s=600).run(W.w_profile_comm, 2)
    st = out["stats"]
    assert out["rolled_back"] and out["nbuckets"] > 1
    assert [b["bucket"] for b in st["buckets"]] == list(range(out["nbuckets"]))
    assert all(b["bytes"] > 0 and b["us"] > 0 and b["busbw_gbs"] > 0 for b in st["buckets"])
    assert st["bytes_per_step"] == sum(b["bytes"] for b in st["buckets"]) and st["world"] == 2
    assert abs(st["comm_us_per_step"] - sum(b["us"] for b in st["buckets"])) < 1e-6


def test_fit_logs_comm_and_throughput(tmp_path):
    import json

    path = str(tmp_path / "run.jsonl")
    out = Distributor(num_processes=2, use_gpu=False, timeout_s=600).run(W.w_fit_with_logging, path)
    assert out["steps"] == 4 and out["nlosses"] == 4
    recs = [json.loads(l) for l in open(path)]
    comm = [r for r in recs if r.get("kind") == "comm"]
    steps = [r for r in recs if r.get("kind") == "step"]
    assert len(comm) == 1 and comm[0]["world"] == 2 and len(comm[0]["buckets"]) > 1
    assert [r["step"] for r in steps] == [2, 4] and all(r["samples_per_s"] > 0 and r["loss"] > 0 for r in steps)
