"""FSDP / ZeRO-3 wrapper on the CPU/gloo plumbing path (world_size = 2)."""
import pytest
import torch

import _workers as W
from databricks_distributed_deep_learning_b200 import Distributor
from databricks_distributed_deep_learning_b200.models import bert_tiny
from databricks_distributed_deep_learning_b200.parallel.fsdp import FullyShardedDataParallel, default_units


@pytest.mark.parametrize("reshard", [False, True])
def test_fsdp_matches_single_process_adamw(reshard):
    steps = 3
    out = Distributor(num_processes=2, use_gpu=False, timeout_s=600).run(W.w_fsdp_bert_tiny, steps, reshard)
    f0, f1 = out["flat"]
    assert torch.equal(f0, f1), "ranks disagree on the gathered parameters"
    ref = W.single_process_bert_tiny(steps, 2)
    assert torch.allclose(f0, ref, rtol=1e-3, atol=2e-4), float((f0 - ref).abs().max())
    assert out["units"] == 3  # root + 2 transformer layers
    assert "bert.encoder.layer.0.attention.self.query.weight" in out["keys"]  # HF layout survives the gather


def test_units_and_single_process_wrapper():
    m = bert_tiny(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    groups, names = default_units(m)
    assert names[0] == "<root>" and names[1:] == ["bert.encoder.layer.0", "bert.encoder.layer.1"]
    assert sum(len(g) for g in groups) == len(list(m.parameters()))
    ref = bert_tiny(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    ref.load_state_dict(m.state_dict())
    fsdp = FullyShardedDataParallel(m, lr=1e-3, weight_decay=0.0)
    opt = torch.optim.AdamW(ref.parameters(), lr=1e-3, weight_decay=0.0)
    ids = torch.randint(0, 512, (2, 16))
    batch = dict(input_ids=ids, attention_mask=torch.ones_like(ids), labels=torch.tensor([0, 1]))
    for _ in range(2):
        fsdp.zero_grad()
        fsdp(**batch)[0].backward()
        fsdp.step()
        opt.zero_grad()
        ref(**batch)[0].backward()
        opt.step()
    sd = fsdp.full_state_dict()
    for k, v in ref.state_dict().items():
        assert torch.allclose(sd[k], v, rtol=1e-4, atol=1e-5), k
    # checkpoint round trip through the full state dict
    m2 = FullyShardedDataParallel(bert_tiny(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0))
    m2.load_full_state_dict(sd)
    for k, v in m2.full_state_dict().items():
        assert torch.equal(v, sd[k]), k
