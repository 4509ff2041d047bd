"""CPU unit tests: all-reduce variant selector (K4) table logic and the Trainer on the plumbing path."""
import json

import torch

from databricks_distributed_deep_learning_b200.engine import Trainer
from databricks_distributed_deep_learning_b200.models import bert_tiny, resnet18
from databricks_distributed_deep_learning_b200.ops import _lib as L
from databricks_distributed_deep_learning_b200.parallel.collectives import VariantSelector
from databricks_distributed_deep_learning_b200.utils.data import SyntheticTokenLoader, cifar_shape_loader


def test_selector_follows_the_measured_table(tmp_path, monkeypatch):
    table = [{"max_bytes": 8192, "algo": "oneshot", "blocks": 4}, {"max_bytes": 1 << 20, "algo": "nvls", "blocks": 16},
             {"max_bytes": 1 << 62, "algo": "nvls", "blocks": 8}]
    p = tmp_path / "t.json"
    p.write_text(json.dumps({"world": 8, "table": table}))
    monkeypatch.setenv("DDL_ALLREDUCE_TABLE", str(p))
    sel = VariantSelector(8, nvls=True)
    assert sel.pick(1024, True) == (L.ALGO_ONESHOT, 4)
    assert sel.pick(8192, True) == (L.ALGO_ONESHOT, 4)
    assert sel.pick(8193, True) == (L.ALGO_NVLS, 16)
    assert sel.pick(64 << 20, True) == (L.ALGO_NVLS, 8)
    # the switch path needs multicast and keeps the dtype: otherwise two-shot with enough CTAs to move the bytes
    assert sel.pick(64 << 20, False) == (L.ALGO_TWOSHOT, 64)
    no_mc = VariantSelector(8, nvls=False)
    assert no_mc.pick(1 << 20, True)[0] == L.ALGO_TWOSHOT and no_mc.pick(1 << 20, True)[1] >= 32


def test_selector_default_and_committed_tables():
    monkey_free = VariantSelector(3, nvls=False)  # no table for 3 ranks -> built-in default
    a, nb = monkey_free.pick(1024, True)
    assert a == L.ALGO_ONESHOT and nb >= 1
    for w in (2, 4, 8):  # the tables measured on B200 and committed under profiles/
        t = VariantSelector._load(w)
        assert t is not None and t[-1]["max_bytes"] >= (1 << 40)
        assert all(r["algo"] in ("oneshot", "twoshot", "nvls") for r in t)
        assert [r["max_bytes"] for r in t] == sorted(r["max_bytes"] for r in t)


def test_trainer_cpu_ddp_and_fsdp_learn():
    torch.manual_seed(0)
    for par in ("ddp", "fsdp"):
        tr = Trainer(bert_tiny(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0), lr=2e-3, device=torch.device("cpu"),
                     parallelism=par)
        out = tr.fit(SyntheticTokenLoader(4, seq_len=16, vocab_size=512, steps=8, pin_memory=False, pool=1))
        assert out["steps"] == 8 and out["losses"][-1] < out["losses"][0], (par, out["losses"])
    tr = Trainer(resnet18(num_classes=10, small_input=True), lr=1e-3, device=torch.device("cpu"))
    out = tr.fit(cifar_shape_loader(8, steps=4, pin_memory=False, pool=1))
    assert out["losses"][-1] < out["losses"][0]
    sd = tr.opt.state_dict()
    assert sd["step"] == 5.0 and len(sd["state"]) == len(list(tr.model.parameters()))
