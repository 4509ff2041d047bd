"""Checkpoint correctness into a LIVE trainer (SURVEY.md section 5 "Checkpoint / resume"; the reference only saves,
notebooks/cv/onnx_experiments.py:198): weights-only loads must re-seed the optimizer's fp32 masters, resume must
continue exactly, FSDP must save/restore its sharded optimizer state and load correctly with shared parameter buffers."""
import pytest
import torch

import _workers as W
from databricks_distributed_deep_learning_b200 import Distributor
from databricks_distributed_deep_learning_b200.engine import Trainer
from databricks_distributed_deep_learning_b200.models import bert_tiny, resnet18
from databricks_distributed_deep_learning_b200.utils import checkpoint as ck


def _batches(n, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(4, 3, 32, 32, generator=g), torch.randint(0, 10, (4,), generator=g)) for _ in range(n)]


def _flat(tr):
    return torch.cat([p.detach().reshape(-1).float() for p in tr.model.parameters()])


def _trainer(seed):
    torch.manual_seed(seed)
    return Trainer(resnet18(num_classes=10, small_input=True), lr=1e-2, device=torch.device("cpu"), use_graph=False)


def test_weights_only_load_into_live_trainer_moves_from_loaded_values(tmp_path):
    src = _trainer(1)
    for b in _batches(2):
        src.step(b)
    path = str(tmp_path / "w.pt")
    ck.save_checkpoint(path, src.ddp)  # no optimizer state in the file
    loaded = _flat(src).clone()
    dst = _trainer(2)  # different weights, optimizer (and its fp32 masters) already constructed
    assert (_flat(dst) - loaded).abs().max() > 0.1
    ck.load_checkpoint(path, dst.model)  # bare module: the wrapper is found through the module
    assert torch.equal(_flat(dst), loaded)
    dst.step(_batches(1, seed=5)[0])
    # one AdamW step of lr 1e-2 moves every weight by at most ~lr (+ weight decay): still next to the LOADED values
    assert (_flat(dst) - loaded).abs().max() < 0.05
    # safetensors entry point, wrapper passed
    st = str(tmp_path / "w.safetensors")
    ck.save_safetensors(st, src.ddp)
    dst2 = _trainer(3)
    ck.load_safetensors(st, dst2.ddp)
    dst2.step(_batches(1, seed=5)[0])
    assert (_flat(dst2) - loaded).abs().max() < 0.05


def test_ddp_resume_equals_uninterrupted_run(tmp_path):
    data = _batches(6)
    ref = _trainer(1)
    for b in data:
        ref.step(b)
    a = _trainer(1)
    for b in data[:3]:
        a.step(b)
    path = str(tmp_path / "resume.pt")
    a.save_checkpoint(path)
    b2 = _trainer(9)
    meta = b2.load_checkpoint(path)
    assert meta["step"] == 3 and b2.steps_done == 3
    for b in data[3:]:
        b2.step(b)
    assert torch.allclose(_flat(b2), _flat(ref), rtol=0, atol=1e-6), float((_flat(b2) - _flat(ref)).abs().max())


def test_fsdp_single_process_optimizer_state_roundtrip(tmp_path):
    def make(seed):
        torch.manual_seed(seed)
        return Trainer(bert_tiny(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0), lr=1e-3, device=torch.device("cpu"),
                       use_graph=False, parallelism="fsdp")

    ids = torch.randint(0, 512, (2, 16))
    batch = dict(input_ids=ids, attention_mask=torch.ones_like(ids), labels=torch.tensor([0, 1]))
    ref = make(1)
    for _ in range(4):
        ref.step(batch)
    a = make(1)
    for _ in range(2):
        a.step(batch)
    path = str(tmp_path / "fsdp.pt")
    a.save_checkpoint(path)
    blob = torch.load(path, weights_only=False)
    assert blob["optimizer"]["format"] == "fsdp/by-name/v1" and blob["optimizer"]["step"] == 3.0
    assert set(blob["optimizer"]["state"]) == set(dict(a.model.named_parameters()))
    b = make(5)
    b.load_checkpoint(path)
    for _ in range(2):
        b.step(batch)
    sa, sb = ref.ddp.full_state_dict(), b.ddp.full_state_dict()
    for k in sa:
        assert torch.allclose(sa[k].float(), sb[k].float(), rtol=0, atol=1e-6), k


@pytest.mark.parametrize("reshard", [False, True])
def test_fsdp_gloo_resume_and_load(reshard):
    out = Distributor(num_processes=2, use_gpu=False, timeout_s=600).run(W.w_fsdp_resume, reshard)
    assert out["resume_err"] < 1e-6, out
    assert out["load_err"] == 0.0, out
